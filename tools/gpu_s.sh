#!/bin/bash
# 1 GPU: parity + headline with the out-of-line general gray path
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/s3; mkdir -p $O
B="timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
$B > $O/imc_fast.json 2>> $O/bench.err
JB_GRAY_FAST=0 $B > $O/imc_general.json 2>> $O/bench.err
python tools/bench_brief.py $O/imc_*.json
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
JB_IMC_VALIDATE=1 timeout 600 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/validate128.json 2> $O/validate128.err
echo "validate rc=$?"; tail -2 $O/validate128.err
