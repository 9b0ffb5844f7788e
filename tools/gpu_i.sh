#!/bin/bash
# 8 GPUs: multi-rank parity tests at HEAD (NCCL + direct peer), bench at N=8 in both exchange modes and N=4
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/i; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 1500 python -m pytest tests/test_multirank_gpu.py tests/test_cpp_host.py -m gpu -q > $O/pytest_8gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_8gpu.log; tail -6 $O/pytest_8gpu.log
run() { # name nproc env
  $3 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $2 --steps 5 --warmup 3 --no-cpu --no-e2e --no-variants > $O/bench_$1.json 2> $O/bench_$1.err
  python tools/bench_brief.py $O/bench_$1.json
}
run n8_p2p 8 "env JB_P2P=1"
run n8_nccl 8 "env JB_P2P=0"
run n4_p2p 4 "env JB_P2P=1"
