#!/bin/bash
set -u
mkdir -p gpurun_out
for s in "veach_mi 16" "cbox 4"; do
  set -- $s
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_small.csv python tools/sweep_pool.py $1 $2 0 > gpurun_out/ncu_small.log 2>&1; echo "ncu $s rc=$?"
  python tools/launch_shares.py gpurun_out/launches_small.csv
  rm -f gpurun_out/launches_small.csv
done
