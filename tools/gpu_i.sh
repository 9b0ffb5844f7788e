#!/bin/bash
# 8 GPUs: the driver's scaling command at N = 8 at HEAD (bench only)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/i3; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu --no-variants > $O/bench_n8.json 2> $O/bench_n8.err
python tools/bench_brief.py $O/bench_n8.json; tail -2 $O/bench_n8.err
