#!/bin/bash
# Round-2 GPU visit: whole -m gpu suite, smoke, bench (+ reference arm), scenes table, then (each only after its own
# command exited 0 without ncu) the ncu launch list of the bench command and --set full captures of the stage kernels.
set -u
TAG=${1:-r02}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; rc=$?; echo "bench rc=$rc"; cut -c1-900 gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"; cut -c1-400 gpurun_out/bench_ref.json
timeout 300 python tools/profile_stages.py sponza 64 > gpurun_out/stages_sponza.json 2>&1
timeout 900 python tools/bench_scenes.py --cpu > gpurun_out/scenes.json 2> gpurun_out/scenes.err; echo "scenes rc=$?"
timeout 300 python tools/step_overhead.py > gpurun_out/step_overhead.log 2>&1; echo "overhead rc=$?"
# the same probe on the round-1 library (commit 70eb13a, built by hand into tools/variants/): before / after of the host loop
if [ -f tools/variants/lib_r01.so ]; then LJ_LIB=$PWD/tools/variants/lib_r01.so LJ_OVERHEAD_TAG=_r01 timeout 300 python tools/step_overhead.py > gpurun_out/step_overhead_r01.log 2>&1; echo "overhead r01 rc=$?"; fi
timeout 300 python tools/bvh_trace.py sponza matpreview cbox > gpurun_out/bvh_trace.log 2>&1
if [ $rc -eq 0 ]; then
  NB="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-strong"
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $NB > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
  for k in extend shade shadow; do
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_$k -s 6 -c 1 -o gpurun_out/${k}_full -f $NB > gpurun_out/ncu_full_$k.log 2>&1; echo "ncu $k rc=$?"
  done
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_vol_track -s 2 -c 1 -o gpurun_out/vol_track_full -f \
      python tools/sweep_pool.py hetvol 8 16777216 > gpurun_out/ncu_full_track.log 2>&1; echo "ncu track rc=$?"
fi
# the reports are too big to travel (gpurun_out is capped at 64 MiB): summarise them here, keep only the summaries
mkdir -p gpurun_out/profiles_$TAG
python tools/summarize_profiles.py $TAG gpurun_out gpurun_out/profiles_$TAG > gpurun_out/summarize.log 2>&1; echo "summarize rc=$?"
for k in extend shade; do ncu -i gpurun_out/${k}_full.ncu-rep --page source --csv > gpurun_out/${k}_source.csv 2>/dev/null; gzip -f gpurun_out/${k}_source.csv; done
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out
