#!/bin/bash
# A/B of an environment switch on the stage profile: tools/gpu_ab.sh "<env A>" "<env B>" [scene spp]...
# (quick parity first: the path + seam tests)
set -u
A="$1"; B="$2"; shift 2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_volpath.py -m gpu -q -x > gpurun_out/ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/ab_pytest.log
show() { python -c "
import json,sys
d=json.load(sys.stdin); s=d['stage_ms']
print('%-38s %s %3d spp: %7.1f Mrays/s %6.1f Msamples/s render %7.2f ms | extend %6.2f shade %6.2f shadow %6.2f gen %5.2f | iters %d' % (sys.argv[1], d['scene'], d['spp'], d['mrays_per_s'], d['msamples_per_s'], d['render_ms_plain'], s['extend'], s['shade'], s['shadow+finish'], s['generate'], d['iterations']))" "$1"; }
while [ $# -ge 2 ]; do
  scene=$1; spp=$2; shift 2
  for rep in 1 2; do
    env $A timeout 300 python tools/profile_stages.py $scene $spp 2>/dev/null | show "A: $A"
    env $B timeout 300 python tools/profile_stages.py $scene $spp 2>/dev/null | show "B: $B"
  done
done
