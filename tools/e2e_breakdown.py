#!/usr/bin/env python
"""Where the end-to-end step goes: scene creation (upload + BVH build), render, read-back, destroy.
usage: tools/e2e_breakdown.py [scene] [spp] [paths_in_flight]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lajolla_public_b200 import LajollaGPU, RenderParams, RNG_SAMPLE_STREAMS  # noqa: E402
from lajolla_public_b200.api import HostSceneDesc  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "sponza"
    spp = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    pif = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    blob = next(p for p in (os.path.join(ROOT, d, name + ".ljs") for d in ("scenes_flat", "tests/golden")) if os.path.exists(p))
    import numpy as np
    import torch
    gpu = LajollaGPU(0)
    desc = HostSceneDesc(blob, pin=True)
    rows = []
    img = None
    for i in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sc = gpu.scene_from_desc(desc)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        if img is None:
            img = np.empty((sc.height, sc.width, 3), dtype=np.float32)
        _, st = sc.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, paths_in_flight=pif), out=img)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        bs = sc.build_stats
        sc.close()
        torch.cuda.synchronize(); t3 = time.perf_counter()
        rows.append({"create_ms": (t1 - t0) * 1e3, "upload_ms": bs.upload_ms, "bvh_build_ms": bs.bvh_build_ms,
                     "render_call_ms": (t2 - t1) * 1e3, "render_ms": st.render_ms, "destroy_ms": (t3 - t2) * 1e3})
    print(json.dumps({"scene": name, "spp": spp, "paths_in_flight": pif, "passes": rows}, indent=1))


if __name__ == "__main__":
    main()
