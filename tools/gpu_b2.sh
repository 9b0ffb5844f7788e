#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/b2; mkdir -p $O
timeout 900 python bench.py --steps 3 --warmup 3 > $O/bench_default.json 2> $O/bench_default.err
echo "bench rc=$?"; tail -c 600 $O/bench_default.err
N="ncu --set full --clock-control none --import-source on -k regex:k_transport_pool -c 1 -s 1"
timeout 900 $N -o $O/ncu_pool64 python bench.py --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/ncu64.log 2>&1
timeout 3000 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -5 $O/pytest_gpu.log
ls -la $O
