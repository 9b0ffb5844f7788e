#!/bin/bash
# 1 GPU: parity + DDMC variant + headline after the census round / source kernel trims
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/o; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -5 $O/pytest_gpu.log
D="timeout 600 python bench.py --ddmc --cells 128 --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
JB_HPOOL_CTAS=5 $D > $O/ddmc_hpool5.json 2>> $O/bench.err
JB_HPOOL_CTAS=6 $D > $O/ddmc_hpool6.json 2>> $O/bench.err
JB_HPOOL_CTAS=5 JB_HPOOL_READY_LOW=16 $D > $O/ddmc_hpool5_rl16.json 2>> $O/bench.err
JB_HPOOL_CTAS=5 JB_HPOOL_REFILL_MIN=32 $D > $O/ddmc_hpool5_rf32.json 2>> $O/bench.err
python tools/bench_brief.py $O/ddmc_*.json
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants > $O/imc.json 2>> $O/bench.err
python tools/bench_brief.py $O/imc.json
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/o/imc.json') if l.startswith('{')][-1]); print(d["timeline"]["rank0_ms"])
PY
tail -3 $O/bench.err
