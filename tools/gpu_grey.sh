#!/bin/bash
# grey-medium tracking: parity of the volumetric scenes, then hetvol with the one-channel kernels on / off
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_volpath.py tests/test_gpu_fullsize.py tests/test_gpu_partition.py -m gpu -q -x > gpurun_out/grey_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/grey_pytest.log
for rep in 1 2; do
  for s in hetvol vol_cbox_teapot; do
    echo "grey  : $(timeout 300 python tools/sweep_pool.py $s 32 8388608 16777216)"
    echo "colour: $(LJ_NO_GREY_TRACK=1 timeout 300 python tools/sweep_pool.py $s 32 8388608 16777216)"
  done
done 2>&1 | tee gpurun_out/grey_ab.log
