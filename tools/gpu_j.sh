#!/bin/bash
# 2 GPUs at HEAD: multi-rank parity tests, VALIDATE at 128^3, default bench (N=1, CPU baseline + variants), reference arm, N=2, smoke
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/j3; mkdir -p $O
timeout 1500 python -m pytest tests/test_multirank_gpu.py tests/test_cpp_host.py -m gpu -q > $O/pytest_2gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_2gpu.log; tail -4 $O/pytest_2gpu.log
JB_IMC_VALIDATE=1 timeout 600 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/validate128.json 2> $O/validate128.err
echo "validate rc=$?"; tail -2 $O/validate128.err
timeout 900 python bench.py > $O/bench_n1.json 2> $O/bench_n1.err
python tools/bench_brief.py $O/bench_n1.json
timeout 900 python bench.py --impl reference --steps 1 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
python -c "
import json; d=json.loads(open('$O/bench_ref.json').read().strip().splitlines()[-1]); print('reference arm', d['value'], d['cpu_baseline']['cores'], d['config']['cells'], d['config']['cuda_library_mapped'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu --no-variants > $O/bench_n2.json 2> $O/bench_n2.err
python tools/bench_brief.py $O/bench_n2.json
python -c "
import sys; sys.path.insert(0,'.')
import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
