#!/bin/bash
# 1 GPU: the no-allocation entry point's test alone (fresh process), then the whole GPU suite (-x, as the round's end runs it)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/t; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone.log 2>&1
echo "alone rc=$?" >> $O/pytest_noalloc_alone.log; tail -15 $O/pytest_noalloc_alone.log
timeout 3000 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -8 $O/pytest_gpu.log
