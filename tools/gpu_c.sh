#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -8 $O/pytest_gpu.log
# launch list of one default step (cold-cache, serialised: shares only)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches256.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/launches.log 2>&1
# full captures of the two streaming kernels at 128^3
N="ncu --set full --clock-control none --import-source on -c 1"
timeout 900 $N -k regex:k_source_emit -s 1 -o $O/ncu_source python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/ncu_source.log 2>&1
timeout 900 $N -k regex:k_tally -s 2 -o $O/ncu_tally python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/ncu_tally.log 2>&1
for r in source tally; do ncu -i $O/ncu_$r.ncu-rep --page raw --csv > $O/ncu_${r}_raw.csv 2>/dev/null; done
ls -la $O
