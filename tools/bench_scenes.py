#!/usr/bin/env python
"""Throughput of every config scene (BASELINE.json configs 1-5) at its own film size through the C ABI:
Msamples/s and Mrays/s per engine, plus (optionally) the oracle on the host cores for a bounded spp.
usage: tools/bench_scenes.py [--cpu] [scene ...]   -> JSON on stdout"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lajolla_public_b200 import LajollaGPU, RenderParams, RNG_SAMPLE_STREAMS, ENGINE_WAVEFRONT, ENGINE_MEGAKERNEL  # noqa: E402

# scene -> spp used for the GPU measurement (the XML's own sampleCount, capped so a run stays in seconds)
SCENES = {"cbox": 64, "veach_mi": 64, "matpreview": 64, "disney_bsdf": 64, "sponza": 64,
          "volpath_test1": 64, "volpath_test2": 256, "volpath_test3": 256, "volpath_test4": 256, "volpath_test5": 256,
          "volpath_test6": 256, "vol_cbox": 64, "vol_cbox_teapot": 64, "hetvol": 32, "hetvol_colored": 32}


def blob(name):
    for d in ("scenes_flat", os.path.join("tests", "golden")):
        p = os.path.join(ROOT, d, name + ".ljs")
        if os.path.exists(p):
            return p
    return None


def cpu(name, manifest, spp):
    drv = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    xml = os.path.join(ROOT, "oracle", "_ref", "scenes", manifest[name]["xml"])
    if not (os.path.exists(drv) and os.path.exists(xml)):
        return None
    out = subprocess.run([drv, "render", xml, "-o", "/tmp/lj_bs.pfm", "--mode", "B", "--spp", str(spp)],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    lines = [l for l in out.splitlines() if l.startswith("{")]
    return json.loads(lines[-1]) if lines else None


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    want_cpu = "--cpu" in sys.argv
    manifest = json.load(open(os.path.join(ROOT, "tests", "golden", "manifest.json")))
    gpu = LajollaGPU(0)
    res = {}
    for name, spp in SCENES.items():
        if args and name not in args:
            continue
        b = blob(name)
        if b is None:
            continue
        sc = gpu.load_scene(b)
        entry = {"width": sc.width, "height": sc.height, "spp": spp, "bvh_build_ms": sc.build_stats.bvh_build_ms}
        for ename, eng in (("wavefront", ENGINE_WAVEFRONT), ("megakernel", ENGINE_MEGAKERNEL)):
            rp = RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, engine=eng)
            sc.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, engine=eng))   # warm-up at the same spp (the slot arena is sized from it)
            _, st = sc.render(rp)
            entry[ename] = {"render_ms": st.render_ms, "msamples_per_s": st.samples / st.render_ms * 1e-3,
                            "mrays_per_s": st.rays / st.render_ms * 1e-3, "rays_per_sample": st.rays / max(st.samples, 1)}
        # counting pass (flags bit0: nodes / primitives per ray; bit1: per-stage CUDA-event sums) of the wavefront engine:
        # the per-scene N-bar figures and the algorithmic bytes per ray of SURVEY 8(d) / BASELINE.md section 4
        _, st = sc.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, engine=ENGINE_WAVEFRONT, flags=3))
        closest = st.rays_primary + st.rays_extension
        if closest:
            nn, npr = st.nodes_visited / closest, st.prims_tested / closest
            sn, sp = (st.shadow_nodes_visited / st.rays_shadow, st.shadow_prims_tested / st.rays_shadow) if st.rays_shadow else (0.0, 0.0)
            bpr, sbpr = 32 + 20 + 80 * nn + 48 * npr, 32 + 4 + 80 * sn + 48 * sp
            wf = entry["wavefront"]
            entry["counting"] = {"rays_primary": st.rays_primary, "rays_extension": st.rays_extension, "rays_shadow": st.rays_shadow,
                                 "nodes_per_ray": nn, "prims_per_ray": npr, "bytes_per_ray": bpr,
                                 "shadow_nodes_per_ray": sn, "shadow_prims_per_ray": sp, "shadow_bytes_per_ray": sbpr,
                                 "extend_ms": st.extend_ms, "shadow_ms": st.shadow_ms, "shade_ms": st.shade_ms, "generate_ms": st.generate_ms,
                                 "iterations": st.iterations, "bvh_nodes": st.bvh_nodes, "bvh_prims": st.bvh_prims,
                                 # whole-render algorithmic GB/s: (closest-hit rays x B/ray + shadow rays x B/ray) / timed render
                                 "algorithmic_gbps_whole_render": (closest * bpr + st.rays_shadow * sbpr) / (wf["render_ms"] * 1e-3) * 1e-9}
        sc.close()
        if want_cpu:
            r = cpu(name, manifest, max(1, spp // 16))
            if r:
                entry["cpu_oracle"] = {"threads": r["threads"], "spp": r["spp"], "render_s": r["render_s"],
                                       "msamples_per_s": r["msamples_per_s"], "mrays_per_s": r["mrays_per_s"]}
        res[name] = entry
        print(name, json.dumps(entry), file=sys.stderr)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
