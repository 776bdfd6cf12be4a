#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_volpath.py -m gpu -x -q > gpurun_out/vol5_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/vol5_pytest.log
for rep in 1 2; do for s in hetvol hetvol_colored; do echo "$(timeout 300 python tools/sweep_pool.py $s 32 8388608 16777216 | cut -c1-330)"; done; done 2>&1 | tee gpurun_out/vol5.log
for b in 8 32; do echo "budget $b: $(LJ_TRACK_BUDGET=$b timeout 300 python tools/sweep_pool.py hetvol 32 8388608 | cut -c1-200)"; done
for r in 16 28; do echo "refill $r: $(LJ_TRACK_REFILL_BELOW=$r timeout 300 python tools/sweep_pool.py hetvol_colored 32 8388608 | cut -c1-200)"; done
