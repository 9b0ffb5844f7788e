#!/bin/bash
# 1 GPU: registers-vs-warps A/B for the hybrid pool kernel (4/5 CTAs) and the sourcing kernel (3/4 CTAs of 256)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/r; mkdir -p $O
D="timeout 600 python bench.py --ddmc --cells 128 --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
JB_HPOOL_CTAS=5 $D > $O/ddmc_hpool5.json 2>> $O/bench.err
JB_HPOOL_CTAS=4 $D > $O/ddmc_hpool4.json 2>> $O/bench.err
python tools/bench_brief.py $O/ddmc_*.json
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
JB_SOURCE_CTAS=4 $B > $O/imc_src4.json 2>> $O/bench.err
JB_SOURCE_CTAS=3 $B > $O/imc_src3.json 2>> $O/bench.err
python tools/bench_brief.py $O/imc_*.json
tail -3 $O/bench.err
