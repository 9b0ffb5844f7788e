#!/bin/bash
# 1 GPU: the whole GPU suite at HEAD (-x, as the round's end runs it) + smoke
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/q5; mkdir -p $O
timeout 3000 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
python -c "
import sys; sys.path.insert(0,'.')
import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
