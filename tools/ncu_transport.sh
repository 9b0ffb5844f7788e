# ncu capture of the transport kernel on the 64^3 twin (one launch, full set, source import)
set -x
K=${1:-k_transport_imc}
ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -s 1 -o gpurun_out/ncu_transport -f \
  python bench.py --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_run.log 2>&1
tail -3 gpurun_out/ncu_run.log
ls -la gpurun_out/
