#!/bin/bash
# 2 GPUs: NCCL and direct-peer exchange parity, then bench at N=2 in both modes
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/e; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 1500 python -m pytest tests/test_multirank_gpu.py tests/test_cpp_host.py -m gpu -q -x > $O/pytest_2gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_2gpu.log; tail -6 $O/pytest_2gpu.log
T="timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
JB_P2P=0 $T > $O/bench_n2_nccl.json 2> $O/bench_n2_nccl.err
JB_P2P=1 $T > $O/bench_n2_p2p.json 2> $O/bench_n2_p2p.err
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith("{")][-1])
    print(sys.argv[1].split('/')[-1], "value %.4g"%d["value"], "ms/step %.1f"%d["ms_per_step"], "transport ms %.1f"%d["transport_ms_per_step"], "iters", d["last_step_rank0"]["n_iterations"], "n_seg", d["invariants"]["n_segments"], d["invariants"]["n_end"], d["invariants"]["e_end_hex"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
tail -3 $O/bench_n2_p2p.err
