#!/bin/bash
# N-GPU visit: multi-device tests through lj_init_devices, then the bench under torchrun (weak headline + strong block).
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus_$N.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi_device.py tests/test_gpu_partition.py -m gpu -q -s > gpurun_out/pytest_multi_$N.log 2>&1; echo "pytest multi rc=$?"
grep -E "passed|failed|FAILED|ERROR|devices" gpurun_out/pytest_multi_$N.log | tail -30
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 \
    > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"; cat gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
