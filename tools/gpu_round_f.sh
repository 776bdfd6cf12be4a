#!/bin/bash
# GPU visit r01f: full GPU test suite (with the asset decoders and the per-config 1024-spp RMSE tests), smoke,
# per-scene throughput + counting table.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python tools/bench_scenes.py --cpu > gpurun_out/scenes.json 2> gpurun_out/scenes.err; echo "scenes rc=$?"
grep -E "passed|failed|error" gpurun_out/pytest_gpu.log | tail -5
