#!/bin/bash
set -u
mkdir -p gpurun_out
for k in k_vol_main k_vol_nee k_vol_finish_medium; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -o gpurun_out/cap_$k -f python tools/sweep_pool.py vol_cbox 16 1048576 > gpurun_out/ncu_$k.log 2>&1; echo "ncu $k rc=$?"
  ncu -i gpurun_out/cap_$k.ncu-rep --page source --csv > gpurun_out/${k}_source.csv 2>/dev/null; gzip -f gpurun_out/${k}_source.csv
  ncu -i gpurun_out/cap_$k.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; v=rows[2]
for w in ['gpu__time_duration.sum','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__inst_executed.sum','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_sector_hit_rate.pct']:
    print('  ',w, v[h.index(w)])
"
  rm -f gpurun_out/cap_$k.ncu-rep
done
