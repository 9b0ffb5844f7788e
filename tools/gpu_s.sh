#!/bin/bash
# 1 GPU: refill loads with L1::no_allocate — headline + ddmc + continuing quick check
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/s4; mkdir -p $O
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
$B > $O/imc_na.json 2>> $O/bench.err
$B --continuing 4 > $O/imc_na_cont.json 2>> $O/bench.err
python tools/bench_brief.py $O/imc_*.json
python -c "
import json; d=json.loads([l for l in open('$O/imc_na_cont.json') if l.startswith('{')][-1]); print([(round(c['ms_total'],1), round(c['ms_transport'],1)) for c in d['continuing']['per_cycle_rank0']])"
timeout 600 python bench.py --ddmc --cells 128 --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants > $O/ddmc_na.json 2>> $O/bench.err
python tools/bench_brief.py $O/ddmc_na.json
