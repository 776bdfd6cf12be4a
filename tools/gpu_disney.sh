#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_volpath.py -m gpu -x -q > gpurun_out/vol5_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/vol5_pytest.log
timeout 300 python tools/bench_scenes.py hetvol hetvol_colored 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin)
for k,v in d.items(): print(k, 'default pool: %.1f Msamples/s, %.1f ms' % (v['wavefront']['msamples_per_s'], v['wavefront']['render_ms']))"
for s in disney_bsdf matpreview; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 3 -c 1 -o gpurun_out/shade_$s -f python tools/profile_stages.py $s 16 > gpurun_out/ncu_shade_$s.log 2>&1; echo "ncu $s rc=$?"
  ncu -i gpurun_out/shade_$s.ncu-rep --page source --csv > gpurun_out/shade_${s}_source.csv 2>/dev/null; gzip -f gpurun_out/shade_${s}_source.csv
  rm -f gpurun_out/shade_$s.ncu-rep
done
