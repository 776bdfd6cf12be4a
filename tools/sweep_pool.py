#!/usr/bin/env python
"""Throughput of one scene against the number of path slots kept in flight by the wavefront engines.
usage: tools/sweep_pool.py scene spp [slots ...]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lajolla_public_b200 import LajollaGPU, RenderParams, RNG_SAMPLE_STREAMS  # noqa: E402


def main():
    name, spp = sys.argv[1], int(sys.argv[2])
    slots = [int(a) for a in sys.argv[3:]] or [1 << 20, 1 << 21, 1 << 22, 1 << 23]
    blob = next(p for p in (os.path.join(ROOT, d, name + ".ljs") for d in ("scenes_flat", "tests/golden")) if os.path.exists(p))
    sc = LajollaGPU(0).load_scene(blob)
    import numpy as np
    out = np.empty((sc.height, sc.width, 3), dtype=np.float32)
    res = {}
    for n in slots:
        rp = RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, paths_in_flight=n)
        sc.render(rp, out=out)
        best = None
        for _ in range(2):
            _, st = sc.render(rp, out=out)
            best = st if best is None or st.render_ms < best.render_ms else best
        res[n] = {"render_ms": best.render_ms, "msamples_per_s": best.samples / best.render_ms * 1e-3,
                  "mrays_per_s": best.rays / best.render_ms * 1e-3, "iterations": best.iterations}
    print(json.dumps({"scene": name, "spp": spp, "slots": res}))


if __name__ == "__main__":
    main()
