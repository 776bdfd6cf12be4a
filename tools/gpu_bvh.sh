#!/bin/bash
# BVH build visit: parity of the rebuilt trees, then per-phase build times with the cooperative kernels on and off.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_raycast.py tests/test_gpu_seam.py tests/test_gpu_robustness.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/bvh_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/bvh_pytest.log
for rep in 1 2; do
  echo "== cooperative (default)"; timeout 300 python tools/bvh_trace.py sponza matpreview cbox 2>&1
  echo "== launch sequences (LJ_PLOC_COOP=0 LJ_BVH8_COOP=0)"; LJ_PLOC_COOP=0 LJ_BVH8_COOP=0 timeout 300 python tools/bvh_trace.py sponza matpreview cbox 2>&1
done > gpurun_out/bvh_trace_ab.log
cat gpurun_out/bvh_trace_ab.log | grep -v "PLOC after"
for e in "LJ_X=0" "LJ_PLOC_COOP=0 LJ_BVH8_COOP=0"; do
  env $e timeout 300 python tools/profile_stages.py sponza 64 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin); s=d['stage_ms']
print('%-34s render %7.2f ms | extend %6.2f shade %6.2f shadow %6.2f' % (sys.argv[1], d['render_ms_plain'], s['extend'], s['shade'], s['shadow+finish']))" "$e"
done
