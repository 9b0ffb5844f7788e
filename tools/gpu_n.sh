#!/bin/bash
# 1 GPU: ncu --set full of the thermal k_source_emit launch (128^3: 1.25e8 photons)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/n; mkdir -p $O
N="ncu --set full --clock-control none --import-source on -c 1 -s 0"
timeout 900 $N -k regex:k_source_emit -o $O/ncu_source -f python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/ncu_source.log 2>&1
ncu -i $O/ncu_source.ncu-rep --page raw --csv > $O/ncu_source_raw.csv 2>/dev/null
ncu -i $O/ncu_source.ncu-rep --page source --csv > $O/ncu_source_source.csv 2>/dev/null
rm -f $O/ncu_source.ncu-rep
ls -la $O
