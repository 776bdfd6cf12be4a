#!/bin/bash
# zero-numerator division shortcut (div_z): the whole parity suite, then the stage profile and the scene rates
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/divz_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/divz_pytest.log
for s in "sponza 64" "matpreview 64" "disney_bsdf 64" "cbox 64" "veach_mi 64"; do
  timeout 300 python tools/profile_stages.py $s 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin); s=d['stage_ms']
print('%-12s %7.1f Mrays/s %6.1f Msamples/s render %7.2f ms | extend %6.2f shade %6.2f shadow %6.2f gen %5.2f' % (d['scene'], d['mrays_per_s'], d['msamples_per_s'], d['render_ms_plain'], s['extend'], s['shade'], s['shadow+finish'], s['generate']))"
done 2>&1 | tee gpurun_out/divz_stages.log
for s in hetvol hetvol_colored vol_cbox_teapot vol_cbox volpath_test3; do echo "$(timeout 300 python tools/sweep_pool.py $s 32 8388608 | cut -c1-200)"; done 2>&1 | tee gpurun_out/divz_vol.log
