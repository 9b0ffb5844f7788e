#!/bin/bash
# 1 GPU: DDMC variant with the shared-memory carve-out set by need (more L1), then the DDMC / hybrid parity tests
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/m3; mkdir -p $O
D="timeout 600 python bench.py --ddmc --cells 128 --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
$D > $O/ddmc_a.json 2>> $O/bench.err
JB_HPOOL_CTAS=6 $D > $O/ddmc_6.json 2>> $O/bench.err
python tools/bench_brief.py $O/ddmc_*.json
timeout 1500 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "ddmc or hybrid or smr or stiff" > $O/pytest_ddmc.log 2>&1
echo "pytest rc=$?" >> $O/pytest_ddmc.log; tail -3 $O/pytest_ddmc.log
