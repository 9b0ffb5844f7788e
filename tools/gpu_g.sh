#!/bin/bash
# 1 GPU: parity at HEAD, bench (A/B of the round-key change vs call f), launch list of a sorted step
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/g; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
$B > $O/bench_base.json 2> $O/bench.err
python tools/bench_brief.py $O/bench_base.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_sorted256.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants --sort-interval 1 > $O/launches_sorted.log 2>&1
python tools/launch_summary.py $O/launches_sorted256.csv 2>/dev/null | head -30
