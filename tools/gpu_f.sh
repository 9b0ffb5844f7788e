#!/bin/bash
# 1 GPU: sanitizer logs, VALIDATE at 128^3, launch list of the default step, default bench line
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/f; mkdir -p $O
timeout 1500 python tools/sanitize.py $O inf stepdiff_smr_hybrid smr3d_hybrid > $O/sanitize_summary.txt 2>&1
cat $O/sanitize_summary.txt
JB_IMC_VALIDATE=1 timeout 600 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/validate128.json 2> $O/validate128.err
echo "validate rc=$?"; tail -2 $O/validate128.err
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches256.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/launches.log 2>&1
