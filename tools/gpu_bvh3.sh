#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_raycast.py tests/test_gpu_seam.py tests/test_gpu_robustness.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/bvh_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/bvh_pytest.log
for e in "LJ_X=0" "LJ_SORT_BITS=63" "LJ_PLOC_STOP=256"; do
  echo "== $e"; env $e timeout 300 python tools/bvh_trace.py sponza matpreview 2>&1 | grep -v "PLOC after"
done 2>&1 | tee gpurun_out/bvh_trace_sort.log
for s in sponza matpreview cbox; do for e in "LJ_X=0" "LJ_SORT_BITS=63"; do
  env $e timeout 300 python tools/profile_stages.py $s 64 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin); s=d['stage_ms']
print('%-10s %-18s render %7.2f ms | extend %6.2f shade %6.2f shadow %6.2f' % (sys.argv[2], sys.argv[1], d['render_ms_plain'], s['extend'], s['shade'], s['shadow+finish']))" "$e" $s
done; done 2>&1 | tee gpurun_out/bvh_sort_quality.log
for s in hetvol hetvol_colored; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_$s.csv python tools/sweep_pool.py $s 8 8388608 > gpurun_out/ncu_$s.log 2>&1; echo "ncu $s rc=$?"
  python tools/launch_shares.py gpurun_out/launches_$s.csv | tee gpurun_out/launch_shares_$s.txt
  rm -f gpurun_out/launches_$s.csv
done
