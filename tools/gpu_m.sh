#!/bin/bash
# 1 GPU: parity + DDMC variant after moving the carry block of the refill into the deferred phase
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/m2; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
D="timeout 600 python bench.py --ddmc --cells 128 --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
$D > $O/ddmc_a.json 2>> $O/bench.err
$D > $O/ddmc_b.json 2>> $O/bench.err
python tools/bench_brief.py $O/ddmc_*.json
tail -3 $O/bench.err
