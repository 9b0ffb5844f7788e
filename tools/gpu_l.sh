#!/bin/bash
# 1 GPU: ncu --set full of the hybrid pool kernel, the old hybrid kernel (64^3 DDMC twin) and the IMC pool kernel (64^3)
# (reports are exported to CSV on the box: three .ncu-rep files exceed what gpurun copies back)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/l; mkdir -p $O
N="ncu --set full --clock-control none --import-source on -c 1 -s 1"
D="python bench.py --ddmc --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants"
cap() { # name regex env... -- command
  local name=$1 rx=$2; shift 2
  timeout 900 env "$@" > $O/ncu_$name.log 2>&1
  ncu -i $O/ncu_$name.ncu-rep --page raw --csv > $O/ncu_${name}_raw.csv 2>/dev/null
  ncu -i $O/ncu_$name.ncu-rep --page source --csv > $O/ncu_${name}_source.csv 2>/dev/null
  rm -f $O/ncu_$name.ncu-rep
}
cap hpool x JB_HPOOL_CTAS=5 $N -k regex:k_transport_hpool -o $O/ncu_hpool -f $D
cap hybrid x JB_TRANSPORT_VARIANT=1 $N -k regex:k_transport_hybrid -o $O/ncu_hybrid -f $D
cap pool x JB_X=0 $N -k regex:k_transport_pool -o $O/ncu_pool -f python bench.py --cells 64 --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants
ls -la $O
