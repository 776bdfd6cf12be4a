#!/usr/bin/env python
"""Per-stage breakdown of one wavefront render (CUDA events around each stage) and the
traversal counting pass (nodes / primitives visited per ray): the inputs of bench.py's roofline."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lajolla_public_b200 import LajollaGPU, RenderParams, RNG_SAMPLE_STREAMS  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "sponza"
    spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    blob = next(p for p in (os.path.join(ROOT, d, name + ".ljs") for d in ("scenes_flat", "tests/golden")) if os.path.exists(p))
    g = LajollaGPU(0)
    sc = g.load_scene(blob)
    import numpy as np
    out = np.empty((sc.height, sc.width, 3), dtype=np.float32)
    sc.render(RenderParams(spp=spp), out=out)   # warm-up
    _, st = sc.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, flags=2), out=out)
    _, ct = sc.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, flags=1), out=out)
    _, pl = sc.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS), out=out)
    ext = ct.rays_primary + ct.rays_extension
    res = {"scene": name, "spp": spp, "render_ms_plain": pl.render_ms, "render_ms_timed": st.render_ms,
           "mrays_per_s": pl.rays / pl.render_ms * 1e-3, "msamples_per_s": pl.samples / pl.render_ms * 1e-3,
           "stage_ms": {"generate": st.generate_ms, "extend": st.extend_ms, "shade": st.shade_ms, "shadow+finish": st.shadow_ms},
           "iterations": st.iterations, "launches": pl.kernel_launches,
           "rays": {"primary": pl.rays_primary, "extension": pl.rays_extension, "shadow": pl.rays_shadow},
           "closest": {"nodes_per_ray": ct.nodes_visited / max(ext, 1), "prims_per_ray": ct.prims_tested / max(ext, 1)},
           "shadow": {"nodes_per_ray": ct.shadow_nodes_visited / max(ct.rays_shadow, 1), "prims_per_ray": ct.shadow_prims_tested / max(ct.rays_shadow, 1)},
           "bvh": {"nodes": sc.build_stats.bvh_nodes, "prims": sc.build_stats.bvh_prims, "build_ms": sc.build_stats.bvh_build_ms, "upload_ms": sc.build_stats.upload_ms}}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
