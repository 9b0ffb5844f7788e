#!/bin/bash
# 1 GPU: parity with the ordered pool + order-preserving squeeze, default bench with variants
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/h; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -15 $O/pytest_gpu.log
timeout 900 python bench.py --no-cpu > $O/bench_variants.json 2> $O/bench.err
python tools/bench_brief.py $O/bench_variants.json; tail -3 $O/bench.err
