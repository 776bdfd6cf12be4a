#!/bin/bash
# Quick GPU visit for a kernel experiment: ray-cast + path parity tests, then the sponza stage profile (twice).
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_raycast.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/quick_pytest.log
for i in 1 2; do timeout 300 env "$@" python tools/profile_stages.py sponza 64 > gpurun_out/quick_stages_$i.json 2>&1; python - <<PY
import json
d=json.load(open("gpurun_out/quick_stages_$i.json")); s=d["stage_ms"]
print("Mrays/s %.1f render %.1f ms extend %.1f shade %.1f shadow %.1f gen %.1f nodes/ray %.2f prims/ray %.2f iters %d" % (d["mrays_per_s"], d["render_ms_plain"], s["extend"], s["shade"], s["shadow+finish"], s["generate"], d["closest"]["nodes_per_ray"], d["closest"]["prims_per_ray"], d["iterations"]))
PY
done
