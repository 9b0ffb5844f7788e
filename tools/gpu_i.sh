#!/bin/bash
# N GPUs (first argument): the driver's scaling command at HEAD (bench only)
cd "$GRAFT_REPO_ROOT" || exit 1
N=${1:-4}
O=gpurun_out/i4; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu --no-variants > $O/bench_n$N.json 2> $O/bench_n$N.err
python tools/bench_brief.py $O/bench_n$N.json; tail -1 $O/bench_n$N.err
