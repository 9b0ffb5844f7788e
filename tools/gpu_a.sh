#!/bin/bash
# round-2 GPU call A: pool kernel (v10) correctness + first A/B against v9
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > $O/gpu.txt 2>&1
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "variants or pool_kernel or validate_build_finds" > $O/pytest_variants.log 2>&1
echo "pytest rc=$?" >> $O/pytest_variants.log
tail -5 $O/pytest_variants.log
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e"
JB_TRANSPORT_VARIANT=1 $B > $O/bench_v9.json 2> $O/bench_v9.err
JB_TRANSPORT_VARIANT=2 $B > $O/bench_v10_56_16_24.json 2> $O/bench_v10.err
JB_TRANSPORT_VARIANT=2 JB_POOL_SLOTS=64 $B > $O/bench_v10_64_16_24.json 2>> $O/bench_v10.err
JB_TRANSPORT_VARIANT=2 JB_POOL_REFILL_MIN=8 $B > $O/bench_v10_56_8_24.json 2>> $O/bench_v10.err
JB_TRANSPORT_VARIANT=2 JB_POOL_REFILL_MIN=24 $B > $O/bench_v10_56_24_24.json 2>> $O/bench_v10.err
JB_TRANSPORT_VARIANT=2 JB_POOL_READY_LOW=16 $B > $O/bench_v10_56_16_16.json 2>> $O/bench_v10.err
JB_TRANSPORT_VARIANT=2 JB_POOL_READY_LOW=32 $B > $O/bench_v10_56_16_32.json 2>> $O/bench_v10.err
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], "value %.4g"%d["value"], "ms/step %.1f"%d["ms_per_step"], "transport ms %.1f"%d["transport_ms_per_step"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
