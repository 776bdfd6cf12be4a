#!/usr/bin/env python
"""Mint test fixtures from the oracle (reference sources + BVH shim), in the authoring
container (needs /root/reference, oracle/_ref/ref_driver and lajolla_public_b200/bin/lj_flatten).

For every case:  <name>.ljs   flattened scene blob (reference parser -> lj_scene_desc)
                 <name>_<mode><spp>.npy   oracle image, float32 (h, w, 3)
Small fixtures go to tests/golden/ (committed); big ones (textures, grids) to scenes_flat/
(git-ignored, but it travels to the GPU box with gpurun).
"""
import json
import os
import subprocess
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("LJ_REFERENCE", "/root/reference")
DRIVER = os.path.join(ROOT, "oracle/_ref/ref_driver")
FLATTEN = os.path.join(ROOT, "lajolla_public_b200/bin/lj_flatten")
sys.path.insert(0, ROOT)
from lajolla_public_b200.api import read_pfm  # noqa: E402

# name, xml (relative to REF/scenes), test film size (w, h), [(mode, spp)], committed?
# One blob per scene (at the XML's own film size); tests override the film size when loading it.
CASES = [
    ("cbox", "cbox/cbox.xml", (128, 128), [("A", 1), ("B", 1), ("B", 64), ("A", 1024), ("B", 1024)], True),
    ("veach_mi", "veach_mi/mi.xml", (192, 128), [("A", 1), ("B", 1), ("B", 64)], True),
    ("sponza", "sponza/sponza.xml", (192, 144), [("A", 1), ("B", 1), ("B", 64)], False),
    ("matpreview", "matpreview/matpreview.xml", (171, 128), [("A", 1), ("B", 1), ("B", 64)], False),
    # scenes/disney_bsdf_test: the matpreview geometry with each Disney lobe / the full Disney BSDF
    ("disney_diffuse", "disney_bsdf_test/disney_diffuse.xml", (171, 128), [("A", 1), ("B", 1), ("B", 64)], False),
    ("disney_metal", "disney_bsdf_test/disney_metal.xml", (171, 128), [("A", 1), ("B", 1), ("B", 64)], False),
    ("disney_glass", "disney_bsdf_test/disney_glass.xml", (171, 128), [("A", 1), ("B", 1), ("B", 64)], False),
    ("disney_clearcoat", "disney_bsdf_test/disney_clearcoat.xml", (171, 128), [("A", 1), ("B", 1), ("B", 64)], False),
    ("disney_sheen", "disney_bsdf_test/disney_sheen.xml", (171, 128), [("A", 1), ("B", 1), ("B", 64)], False),
    ("disney_bsdf", "disney_bsdf_test/disney_bsdf.xml", (171, 128), [("A", 1), ("B", 1), ("B", 64)], False),
    ("simple_sphere", "disney_bsdf_test/simple_sphere.xml", (128, 128), [("A", 1), ("B", 1), ("B", 64)], True),
    ("pixel_filter_test", "pixel_filter_test/pixel_filter_test.xml", (128, 128), [("A", 1), ("B", 1), ("B", 64)], True),
    # vol_path_tracing.h: versions 1, 2 and the final integrator (3..6), homogeneous + heterogeneous media
    ("volpath_test1", "volpath_test/volpath_test1.xml", (64, 64), [("A", 1), ("B", 1), ("B", 16)], True),
    ("volpath_test2", "volpath_test/volpath_test2.xml", (64, 64), [("A", 1), ("B", 1), ("B", 16)], True),
    ("volpath_test3", "volpath_test/volpath_test3.xml", (64, 64), [("A", 1), ("B", 1), ("B", 16)], True),
    ("volpath_test4", "volpath_test/volpath_test4.xml", (64, 64), [("A", 1), ("B", 1), ("B", 16)], True),
    ("volpath_test4_2", "volpath_test/volpath_test4_2.xml", (64, 64), [("A", 1), ("B", 1), ("B", 16)], True),
    ("volpath_test5", "volpath_test/volpath_test5.xml", (64, 64), [("A", 1), ("B", 1), ("B", 16)], True),
    ("volpath_test5_2", "volpath_test/volpath_test5_2.xml", (64, 64), [("A", 1), ("B", 1), ("B", 16)], True),
    ("volpath_test6", "volpath_test/volpath_test6.xml", (64, 64), [("A", 1), ("B", 1), ("B", 16)], True),
    ("vol_cbox", "volpath_test/vol_cbox.xml", (96, 96), [("A", 1), ("B", 1), ("B", 16)], True),
    ("vol_cbox_teapot", "volpath_test/vol_cbox_teapot.xml", (96, 96), [("A", 1), ("B", 1), ("B", 16)], False),
    ("hetvol", "volpath_test/hetvol.xml", (96, 72), [("A", 1), ("B", 1), ("B", 16)], False),
    ("hetvol_colored", "volpath_test/hetvol_colored.xml", (96, 72), [("A", 1), ("B", 1), ("B", 16)], False),
]


# north-star threshold (ii): 1024-spp oracle renders (mode A = the reference's render(), mode B = per-sample
# streams) at a reduced film size; images only (a few hundred KB each), always committed to tests/golden/.
RMSE_CASES = [
    ("veach_mi", "veach_mi/mi.xml", (96, 64)),
    ("sponza", "sponza/sponza.xml", (128, 96)),
    ("matpreview", "matpreview/matpreview.xml", (114, 85)),
    ("disney_bsdf", "disney_bsdf_test/disney_bsdf.xml", (114, 85)),
    ("volpath_test3", "volpath_test/volpath_test3.xml", (64, 64)),
    ("vol_cbox", "volpath_test/vol_cbox.xml", (64, 64)),
    ("hetvol", "volpath_test/hetvol.xml", (64, 48)),
]


def run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        print(r.stdout[-2000:])
        raise SystemExit(f"command failed: {' '.join(cmd)}")
    return r.stdout


def imwrite_kat():
    """Known-answer vector for the output step: a small f32 image full of rounding / overflow / denormal /
    non-finite cases, written by the reference's own imwrite() as .exr (tinyexr, fp16, ZIP) and .pfm."""
    import struct
    rng = np.random.default_rng(7)
    w, h = 37, 21
    img = rng.standard_normal((h, w, 3)).astype(np.float32) * np.exp(rng.uniform(-20, 12, (h, w, 3))).astype(np.float32)
    special = np.array([0.0, -0.0, 1.0, -1.0, 65504.0, 65520.0, 65519.99, 1e-8, 5.96e-8, 2.98e-8, 2.9802322e-08, 6.1e-5,
                        6.0975552e-05, np.inf, -np.inf, np.nan, 1e-40, -1e-40, 1.0009766, 1.00048828125, 1.000488,
                        0.33333334, 1e5, -1e5, 3.0e-8], dtype=np.float32)
    img.reshape(-1)[:len(special)] = special
    gold = os.path.join(ROOT, "tests/golden")
    src = os.path.join(gold, "imwrite_kat.f32img")
    with open(src, "wb") as f:
        f.write(struct.pack("<ii", w, h) + img.tobytes())
    for ext in ("exr", "pfm"):
        run([DRIVER, "imwrite", src, os.path.join(gold, "imwrite_kat." + ext)])
    print(f"  imwrite_kat: {w}x{h} -> .exr {os.path.getsize(os.path.join(gold, 'imwrite_kat.exr'))} B, .pfm {os.path.getsize(os.path.join(gold, 'imwrite_kat.pfm'))} B")


def rmse_goldens(only):
    """python tools/make_goldens.py --rmse [names]: mint <name>_{A,B}1024.npy + manifest[name]["rmse1024"]."""
    manifest_path = os.path.join(ROOT, "tests/golden/manifest.json")
    manifest = json.load(open(manifest_path))
    for name, xml, res in RMSE_CASES:
        if only and name not in only:
            continue
        xml_path = os.path.join(REF, "scenes", xml)
        entry = {"res": list(res), "images": {}}
        for mode in ("A", "B"):
            pfm = f"/tmp/lj_golden_{name}_{mode}1024.pfm"
            out = run([DRIVER, "render", xml_path, "-o", pfm, "--mode", mode, "--spp", "1024", "--res", str(res[0]), str(res[1])])
            stats = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
            img = read_pfm(pfm)
            np.save(os.path.join(ROOT, "tests/golden", f"{name}_{mode}1024.npy"), img)
            entry["images"][f"{mode}1024"] = {"intersect_rays": stats["intersect_rays"], "occluded_rays": stats["occluded_rays"],
                                              "mean": float(img.mean()), "render_s": stats["render_s"], "threads": stats["threads"]}
            os.remove(pfm)
            print(f"  {name} {mode}1024 {res}: mean {img.mean():.6f}, {stats['render_s']:.1f} s on {stats['threads']} threads", flush=True)
        manifest[name]["rmse1024"] = entry
        json.dump(manifest, open(manifest_path, "w"), indent=1, sort_keys=True)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--rmse":
        return rmse_goldens(set(sys.argv[2:]))
    only = set(sys.argv[1:])
    if not only or "imwrite_kat" in only:
        imwrite_kat()
    os.makedirs(os.path.join(ROOT, "scenes_flat"), exist_ok=True)
    manifest_path = os.path.join(ROOT, "tests/golden/manifest.json")
    manifest = json.load(open(manifest_path)) if os.path.exists(manifest_path) else {}
    for name, xml, res, renders, committed in CASES:
        if only and name not in only:
            continue
        outdir = os.path.join(ROOT, "tests/golden" if committed else "scenes_flat")
        xml_path = os.path.join(REF, "scenes", xml)
        blob = os.path.join(outdir, name + ".ljs")
        res_args = ["--res", str(res[0]), str(res[1])]
        print(run([FLATTEN, xml_path, blob]).strip().splitlines()[-1])
        entry = {"xml": xml, "res": res, "committed": committed, "images": {}}
        for mode, spp in renders:
            pfm = f"/tmp/lj_golden_{name}_{mode}{spp}.pfm"
            out = run([DRIVER, "render", xml_path, "-o", pfm, "--mode", mode, "--spp", str(spp)] + res_args)
            stats = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
            img = read_pfm(pfm)
            key = f"{mode}{spp}"
            np.save(os.path.join(outdir, f"{name}_{key}.npy"), img)
            entry["images"][key] = {"intersect_rays": stats["intersect_rays"], "occluded_rays": stats["occluded_rays"],
                                    "mean": float(img.mean())}
            os.remove(pfm)
            print(f"  {name} {key}: mean {img.mean():.6f}, rays {stats['intersect_rays']}+{stats['occluded_rays']}")
        if name in manifest and "rmse1024" in manifest[name]:
            entry["rmse1024"] = manifest[name]["rmse1024"]
        manifest[name] = entry
    json.dump(manifest, open(manifest_path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
