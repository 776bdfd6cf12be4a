#!/bin/bash
# volumetric + BVH-build visit: parity of everything the two changes touch, then throughput
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_volpath.py tests/test_gpu_raycast.py tests/test_gpu_seam.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_partition.py -m gpu -q -x > gpurun_out/vol_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/vol_pytest.log
timeout 900 python tools/bench_scenes.py hetvol hetvol_colored vol_cbox vol_cbox_teapot volpath_test3 > gpurun_out/vol_scenes.json 2> gpurun_out/vol_scenes.err; echo "scenes rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/vol_scenes.json"))
for k,v in d.items():
    print(k, "wavefront %.1f Msamples/s" % v["wavefront"]["msamples_per_s"], "bvh build %.2f ms" % v["bvh_build_ms"])
PY
for s in hetvol hetvol_colored; do timeout 300 python tools/sweep_pool.py $s 32 1048576 2097152 4194304 8388608 16777216; done
for s in sponza cbox matpreview; do LJ_PLOC_NO_TAIL=1 timeout 200 python tools/profile_stages.py $s 8 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('$s no-tail bvh', d['bvh'])"; timeout 200 python tools/profile_stages.py $s 8 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('$s tail    bvh', d['bvh'], 'nodes/ray', d['closest'])"; done
