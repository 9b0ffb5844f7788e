#!/bin/bash
# launch list of the default bench command (cold-cache, serialised: shares only) -> gpurun_out/ll/
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/ll; mkdir -p $O
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches256.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-variants > $O/launches.log 2>&1
python tools/launch_summary.py $O/launches256.csv | head -30
