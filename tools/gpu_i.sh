#!/bin/bash
# 8 GPUs: multi-rank parity tests at HEAD, then the driver's scaling commands (default arguments) at N = 8, 4, 2, 1
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/i2; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 1500 python -m pytest tests/test_multirank_gpu.py tests/test_cpp_host.py -m gpu -q > $O/pytest_8gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_8gpu.log; tail -4 $O/pytest_8gpu.log
run() { # name nproc
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $2 --steps 5 --warmup 3 --no-cpu --no-variants > $O/bench_$1.json 2> $O/bench_$1.err
  python tools/bench_brief.py $O/bench_$1.json
}
run n8 8
JB_P2P=0 run n8_nccl 8
run n4 4
run n2 2
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-variants > $O/bench_n1.json 2> $O/bench_n1.err
python tools/bench_brief.py $O/bench_n1.json
