#!/bin/bash
# 1 GPU: parity + default bench line at HEAD (5 CTAs x 64 slots, 96 registers)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/q; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -5 $O/pytest_gpu.log
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench.err
python tools/bench_brief.py $O/bench_default.json; tail -3 $O/bench.err
