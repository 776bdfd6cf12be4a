#!/bin/bash
# One GPU visit: parity tests, smoke, bench, then (each only after its command exited 0 without ncu)
# the ncu launch list of the bench command and one --set full capture of the dominant kernel.
# Everything lands in gpurun_out/ (merged back by gpurun).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; rc=$?; echo "bench rc=$rc"; cat gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 300 python tools/profile_stages.py sponza 64 > gpurun_out/stages_sponza.json 2>&1
timeout 900 python tools/bench_scenes.py --cpu > gpurun_out/scenes.json 2> gpurun_out/scenes.err; echo "scenes rc=$?"
if [ $rc -eq 0 ]; then
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_wf_extend -s 6 -c 1 -o gpurun_out/extend_full -f \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 6 -c 1 -o gpurun_out/shade_full -f \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_shade.log 2>&1; echo "ncu shade rc=$?"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_wf_shadow -s 6 -c 1 -o gpurun_out/shadow_full -f \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_shadow.log 2>&1; echo "ncu shadow rc=$?"
fi
tail -5 gpurun_out/pytest_gpu.log
