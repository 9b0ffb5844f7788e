#!/bin/bash
# 1 GPU: ncu --set full of k_transport_pool at HEAD (64^3 twin), exported to CSV on the box
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/l2; mkdir -p $O
N="ncu --set full --clock-control none --import-source on -c 1 -s 1"
timeout 900 $N -k regex:k_transport_pool -o $O/ncu_pool -f python bench.py --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/ncu_pool.log 2>&1
ncu -i $O/ncu_pool.ncu-rep --page raw --csv > $O/ncu_pool_raw.csv 2>/dev/null
ncu -i $O/ncu_pool.ncu-rep --page source --csv > $O/ncu_pool_source.csv 2>/dev/null
rm -f $O/ncu_pool.ncu-rep
ls -la $O
