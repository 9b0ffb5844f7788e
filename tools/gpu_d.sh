#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/d; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
$B > $O/bench_base.json 2> $O/bench.err
JB_POOL_PREFETCH=1 $B > $O/bench_prefetch.json 2>> $O/bench.err
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], "value %.4g"%d["value"], "ms/step %.1f"%d["ms_per_step"], "transport ms %.1f"%d["transport_ms_per_step"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches256.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/launches.log 2>&1
JB_POOL_PREFETCH=1 timeout 900 ncu --set full --clock-control none -k regex:k_transport_pool -c 1 -s 1 -o $O/ncu_pool_prefetch python bench.py --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/ncu_pf.log 2>&1
ncu -i $O/ncu_pool_prefetch.ncu-rep --page raw --csv > $O/ncu_pool_prefetch_raw.csv 2>/dev/null; rm -f $O/ncu_pool_prefetch.ncu-rep
