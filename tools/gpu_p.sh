#!/bin/bash
# 1 GPU: 96-register configurations of k_transport_pool, 256^3 — third sweep
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/p4; mkdir -p $O
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
run() { n=$1; shift; env "$@" $B > $O/imc_$n.json 2>> $O/bench.err; }
run s64_7x4 JB_X=0
run s64_5x4 JB_POOL_SLOTS=564
run s64_7x3 JB_POOL_SLOTS=764
run s64_5x4_rf16 JB_POOL_SLOTS=564 JB_POOL_REFILL_MIN=16
run s64_5x4_rf32 JB_POOL_SLOTS=564 JB_POOL_REFILL_MIN=32
run s64_5x4_rl24 JB_POOL_SLOTS=564 JB_POOL_READY_LOW=24
run s64_7x3_rl24 JB_POOL_SLOTS=764 JB_POOL_READY_LOW=24
python tools/bench_brief.py $O/imc_*.json
tail -3 $O/bench.err
