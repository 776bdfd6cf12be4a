#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_raycast.py tests/test_gpu_seam.py tests/test_gpu_robustness.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/bvh_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/bvh_pytest.log
for g in 148 296 592 1184; do
  echo "== LJ_PLOC_GRID=$g"; LJ_PLOC_GRID=$g timeout 300 python tools/bvh_trace.py sponza 2>&1 | grep -v "PLOC after"
done 2>&1 | tee gpurun_out/bvh_trace_grid.log
echo "== matpreview"; timeout 300 python tools/bvh_trace.py matpreview 2>&1
