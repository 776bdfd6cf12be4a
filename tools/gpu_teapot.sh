#!/bin/bash
set -u
mkdir -p gpurun_out
for s in vol_cbox_teapot vol_cbox; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/launches_$s.csv python tools/sweep_pool.py $s 16 1048576 > gpurun_out/ncu_$s.log 2>&1; echo "ncu $s rc=$?"
  python tools/launch_shares.py gpurun_out/launches_$s.csv | tee gpurun_out/launch_shares_$s.txt
  rm -f gpurun_out/launches_$s.csv
done
for p in 524288 1048576 2097152 4194304 8388608; do echo "$(timeout 300 python tools/sweep_pool.py vol_cbox_teapot 64 $p | cut -c1-220)"; done
