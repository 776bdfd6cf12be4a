#!/usr/bin/env python
"""Steady-state scene-create cost (upload + BVH build) and the per-phase breakdown of the build (LJ_BVH_TRACE).
usage: tools/bvh_trace.py [scene ...]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lajolla_public_b200 import LajollaGPU  # noqa: E402
from lajolla_public_b200.api import HostSceneDesc  # noqa: E402

gpu = LajollaGPU(0)
for name in (sys.argv[1:] or ["sponza", "cbox", "matpreview"]):
    blob = next(p for p in (os.path.join(ROOT, d, name + ".ljs") for d in ("scenes_flat", "tests/golden")) if os.path.exists(p))
    desc = HostSceneDesc(blob, pin=True)
    for i in range(4):
        t0 = time.perf_counter()
        sc = gpu.scene_from_desc(desc)
        dt = (time.perf_counter() - t0) * 1e3
        st = sc.build_stats
        sc.close()
    print(f"{name}: steady-state create {dt:.3f} ms wall; upload {st.upload_ms:.3f} ms, bvh build {st.bvh_build_ms:.3f} ms ({st.bvh_nodes} nodes, {st.bvh_prims} prims)", flush=True)
    os.environ["LJ_BVH_TRACE"] = "1"
    sys.stderr.flush()
    gpu.scene_from_desc(desc).close()
    os.environ.pop("LJ_BVH_TRACE")
    desc.close()
gpu.close()
