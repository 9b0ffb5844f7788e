#!/bin/bash
# 2 GPUs: full GPU parity suite at HEAD, bench N=1 (default, variants) and N=2 (direct peer)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/j; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu_2gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu_2gpu.log; tail -5 $O/pytest_gpu_2gpu.log
timeout 900 python bench.py --no-cpu > $O/bench_n1.json 2> $O/bench_n1.err
python tools/bench_brief.py $O/bench_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu --no-e2e --no-variants > $O/bench_n2.json 2> $O/bench_n2.err
python tools/bench_brief.py $O/bench_n2.json
