# per-launch durations of one bench step (ncu, serialised + cold cache: shares only, not absolutes)
set -x
CELLS=${1:-128}
python bench.py --cells $CELLS --steps 1 --warmup 1 --no-cpu --no-e2e > gpurun_out/ll_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --cells $CELLS --steps 1 --warmup 1 --no-cpu --no-e2e > gpurun_out/ll_ncu.log 2>&1
tail -2 gpurun_out/ll_plain.log
