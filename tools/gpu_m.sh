#!/bin/bash
# 1 GPU: parity + DDMC variant A/B after shrinking the hybrid pool kernel's hot code
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/m; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -5 $O/pytest_gpu.log
D="timeout 600 python bench.py --ddmc --cells 128 --steps 3 --warmup 2 --no-cpu --no-e2e --no-variants"
for c in 5 6; do JB_HPOOL_CTAS=$c $D > $O/ddmc_hpool$c.json 2>> $O/bench.err; done
JB_TRANSPORT_VARIANT=1 $D > $O/ddmc_old.json 2>> $O/bench.err
python tools/bench_brief.py $O/ddmc_*.json
N="ncu --set full --clock-control none --import-source on -c 1 -s 1"
JB_HPOOL_CTAS=5 timeout 900 $N -k regex:k_transport_hpool -o $O/ncu_hpool -f python bench.py --ddmc --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/ncu_hpool.log 2>&1
ncu -i $O/ncu_hpool.ncu-rep --page raw --csv > $O/ncu_hpool_raw.csv 2>/dev/null
ncu -i $O/ncu_hpool.ncu-rep --page source --csv > $O/ncu_hpool_source.csv 2>/dev/null
rm -f $O/ncu_hpool.ncu-rep
tail -3 $O/bench.err
