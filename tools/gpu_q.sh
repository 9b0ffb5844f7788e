#!/bin/bash
# 1 GPU: parity (-x, as the round's end runs it) + default bench line at HEAD
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/q4; mkdir -p $O
timeout 3000 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench.err
python tools/bench_brief.py $O/bench_default.json; tail -3 $O/bench.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/q4/bench_default.json') if l.startswith('{')][-1])
print("e2e %.4g"%d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], "launches", d["gpu_launches"])
print([ (round(c["ms_total"],1), round(c["ms_transport"],1)) for c in d["variants"]["continuing"]["per_cycle_rank0"]])
PY
