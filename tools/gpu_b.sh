#!/bin/bash
# round-2 GPU call B: ncu of the pool kernel, more tuning points, 3D SMR parity tests
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/b; mkdir -p $O
timeout 1200 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "smr3d or golden_runs or ddmc or hybrid" > $O/pytest_smr3d.log 2>&1
echo "pytest rc=$?" >> $O/pytest_smr3d.log
tail -5 $O/pytest_smr3d.log
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e"
for cfg in "64 24 32" "64 32 32" "64 16 32" "56 24 32" "56 32 32"; do
  set -- $cfg
  JB_TRANSPORT_VARIANT=2 JB_POOL_SLOTS=$1 JB_POOL_REFILL_MIN=$2 JB_POOL_READY_LOW=$3 $B > $O/bench_v10_$1_$2_$3.json 2>> $O/bench.err
done
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], "value %.4g"%d["value"], "ms/step %.1f"%d["ms_per_step"], "transport ms %.1f"%d["transport_ms_per_step"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
N="ncu --set full --clock-control none --import-source on -k regex:k_transport_pool -c 1 -s 1"
JB_TRANSPORT_VARIANT=2 JB_POOL_SLOTS=56 JB_POOL_REFILL_MIN=24 JB_POOL_READY_LOW=32 timeout 900 $N -o $O/ncu_pool56 python bench.py --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/ncu56.log 2>&1
JB_TRANSPORT_VARIANT=2 JB_POOL_SLOTS=64 JB_POOL_REFILL_MIN=24 JB_POOL_READY_LOW=32 timeout 900 $N -o $O/ncu_pool64 python bench.py --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e > $O/ncu64.log 2>&1
ls -la $O
