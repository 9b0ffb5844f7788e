#!/bin/bash
# 1 GPU: parity with the hybrid pool kernel; DDMC variant at 5/6/7 CTAs per SM and old kernel; A/B guided vs fixed chunk
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/k; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -25 $O/pytest_gpu.log
D="timeout 600 python bench.py --ddmc --cells 128 --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
for c in 5 6 7; do JB_HPOOL_CTAS=$c $D > $O/ddmc_hpool$c.json 2>> $O/bench.err; done
JB_TRANSPORT_VARIANT=1 $D > $O/ddmc_old.json 2>> $O/bench.err
JB_HPOOL_CTAS=6 JB_HPOOL_REFILL_MIN=16 $D > $O/ddmc_hpool6_rf16.json 2>> $O/bench.err
JB_HPOOL_CTAS=6 JB_HPOOL_READY_LOW=16 $D > $O/ddmc_hpool6_rl16.json 2>> $O/bench.err
python tools/bench_brief.py $O/ddmc_*.json
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
$B > $O/imc_guided.json 2>> $O/bench.err
JB_POOL_FIXED_CHUNK=1 $B > $O/imc_fixed.json 2>> $O/bench.err
$B > $O/imc_guided2.json 2>> $O/bench.err
python tools/bench_brief.py $O/imc_*.json
tail -3 $O/bench.err
