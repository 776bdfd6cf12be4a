#!/bin/bash
set -u
mkdir -p gpurun_out
for e in "LJ_PLOC_COOP=0 LJ_PLOC_NO_TAIL=1" "LJ_PLOC_COOP=0" "LJ_PLOC_STOP=1024" "LJ_X=1" "LJ_PLOC_STOP=64" "LJ_PLOC_GRID=148" "LJ_PLOC_GRID=1184"; do
  echo "== $e"; env $e timeout 300 python tools/bvh_trace.py sponza matpreview 2>&1 | grep "fingerprint\|steady\|rounds"
done 2>&1 | tee gpurun_out/bvh_fingerprint.log
timeout 600 python -m pytest tests/test_gpu_raycast.py tests/test_gpu_seam.py tests/test_gpu_robustness.py tests/test_gpu_parity.py tests/test_gpu_volpath.py -m gpu -q -x > gpurun_out/bvh_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/bvh_pytest.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_vol_track -s 60 -c 1 -o gpurun_out/track_col_full -f python tools/sweep_pool.py hetvol_colored 8 8388608 > gpurun_out/ncu_track_col.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/track_col_full.ncu-rep --page source --csv > gpurun_out/track_col_source.csv 2>/dev/null; gzip -f gpurun_out/track_col_source.csv
ncu -i gpurun_out/track_col_full.ncu-rep --page raw --csv > gpurun_out/track_col_raw.csv 2>/dev/null
rm -f gpurun_out/track_col_full.ncu-rep
