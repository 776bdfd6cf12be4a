#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_volpath.py tests/test_gpu_fullsize.py tests/test_gpu_partition.py -m gpu -x -q > gpurun_out/vol6_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/vol6_pytest.log
for rep in 1 2 3; do timeout 300 python tools/bench_scenes.py hetvol hetvol_colored 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin)
for k,v in d.items(): print(k, 'default pool: %.1f Msamples/s, %.1f ms' % (v['wavefront']['msamples_per_s'], v['wavefront']['render_ms']))"; done
