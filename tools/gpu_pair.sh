#!/bin/bash
# paired density/albedo grid records: parity, then hetvol_colored with the pairing on / off; source-level capture of a full tracking launch
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_volpath.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/pair_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pair_pytest.log
for rep in 1 2; do
  echo "paired  : $(timeout 300 python tools/sweep_pool.py hetvol_colored 32 8388608)"
  echo "separate: $(LJ_NO_GRID_PAIR=1 timeout 300 python tools/sweep_pool.py hetvol_colored 32 8388608)"
done 2>&1 | tee gpurun_out/pair_ab.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_vol_track -s 2 -c 1 -o gpurun_out/track_col_full -f python tools/sweep_pool.py hetvol_colored 8 8388608 > gpurun_out/ncu_track_col.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/track_col_full.ncu-rep --page source --csv > gpurun_out/track_col_source.csv 2>/dev/null; gzip -f gpurun_out/track_col_source.csv
ncu -i gpurun_out/track_col_full.ncu-rep --page raw --csv > gpurun_out/track_col_raw.csv 2>/dev/null
rm -f gpurun_out/track_col_full.ncu-rep
