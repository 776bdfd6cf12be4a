#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_volpath.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/ab2_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/ab2_pytest.log
for rep in 1 2; do
timeout 300 python tools/profile_stages.py sponza 64 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin); s=d['stage_ms']
print('%-12s %7.1f Mrays/s render %7.2f ms | extend %6.2f shade %6.2f shadow %6.2f gen %5.2f' % (d['scene'], d['mrays_per_s'], d['render_ms_plain'], s['extend'], s['shade'], s['shadow+finish'], s['generate']))"
done
timeout 300 python tools/bench_scenes.py hetvol hetvol_colored 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin)
for k,v in d.items(): print(k, 'default pool: %.1f Msamples/s, %.1f ms' % (v['wavefront']['msamples_per_s'], v['wavefront']['render_ms']))"
