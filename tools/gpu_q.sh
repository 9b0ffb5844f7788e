#!/bin/bash
# 1 GPU: parity + headline at HEAD
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/q3; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -5 $O/pytest_gpu.log
timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants > $O/imc.json 2> $O/bench.err
python tools/bench_brief.py $O/imc.json; tail -3 $O/bench.err
