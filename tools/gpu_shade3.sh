#!/bin/bash
set -u
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_raycast.py tests/test_gpu_seam.py -m gpu -x -q 2>&1 | tail -1
for s in "sponza 64" "cbox 64"; do for rep in 1 2; do
  timeout 300 python tools/profile_stages.py $s 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin); s=d['stage_ms']
print('%-12s %3d spp %7.1f Msamples/s render %7.2f ms | extend %6.2f shade %6.2f shadow %6.2f gen %5.2f' % (d['scene'], d['spp'], d['msamples_per_s'], d['render_ms_plain'], s['extend'], s['shade'], s['shadow+finish'], s['generate']))"
done; done
