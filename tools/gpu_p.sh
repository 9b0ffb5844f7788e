#!/bin/bash
# 1 GPU: pool size / L1 trade A/B of k_transport_pool at 256^3
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/p; mkdir -p $O
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
$B > $O/imc_s64.json 2>> $O/bench.err
JB_POOL_SLOTS=48 $B > $O/imc_s48.json 2>> $O/bench.err
JB_POOL_SLOTS=48 JB_POOL_REFILL_MIN=16 JB_POOL_READY_LOW=24 $B > $O/imc_s48_16_24.json 2>> $O/bench.err
JB_POOL_SLOTS=664 $B > $O/imc_s64_6ctas.json 2>> $O/bench.err
JB_POOL_SLOTS=56 $B > $O/imc_s56_8ctas.json 2>> $O/bench.err
python tools/bench_brief.py $O/imc_*.json
tail -3 $O/bench.err
