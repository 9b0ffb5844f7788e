#!/bin/bash
# 1 GPU: ncu --set full of k_transport_hpool at HEAD (C5-ddmc 64^3 twin), exported to CSV on the box
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/l3; mkdir -p $O
N="ncu --set full --clock-control none --import-source on -c 1 -s 1"
timeout 900 $N -k regex:k_transport_hpool -o $O/ncu_hpool -f python bench.py --ddmc --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/ncu_hpool.log 2>&1
ncu -i $O/ncu_hpool.ncu-rep --page raw --csv > $O/ncu_hpool_raw.csv 2>/dev/null
ncu -i $O/ncu_hpool.ncu-rep --page source --csv > $O/ncu_hpool_source.csv 2>/dev/null
rm -f $O/ncu_hpool.ncu-rep
ls -la $O
