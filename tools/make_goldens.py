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


# ---- round 2: parity at the configs' OWN film size (VERDICT r01 "parity is proven on thumbnails") ----
# name, xml, [(mode, spp)]; images go to scenes_flat/<name>_full_<mode><spp>.npy (git-ignored, re-minted by build()).
FULL_CASES = [
    ("cbox", "cbox/cbox.xml", [("A", 1), ("B", 1), ("B", 4)]),                      # 512x512, shipped spp 4
    ("veach_mi", "veach_mi/mi.xml", [("A", 1), ("B", 1), ("B", 16)]),               # 768x512, shipped spp 16
    ("sponza", "sponza/sponza.xml", [("A", 1), ("B", 1), ("B", 8)]),                # 768x575
    ("matpreview", "matpreview/matpreview.xml", [("A", 1), ("B", 1), ("B", 8)]),    # 683x512
    ("disney_bsdf", "disney_bsdf_test/disney_bsdf.xml", [("B", 1)]),                # 683x512
    ("disney_bsdf_array", "disney_bsdf_test/disney_bsdf_array.xml", [("B", 1)]),    # 1366x1024
    ("hetvol", "volpath_test/hetvol.xml", [("A", 1), ("B", 1), ("B", 8)]),          # 768x576
    ("vol_cbox", "volpath_test/vol_cbox.xml", [("B", 1)]),                          # 512x512
]
# blobs that exist only for the full-size cases
FULL_ONLY_BLOBS = {"disney_bsdf_array"}

# ---- round 2: AOV integrators (render.cpp:12-69, selected at parse_scene.cpp:295-304); RNG-free ----
AOV_TYPES = ["depth", "shadingNormal", "meanCurvature", "rayDifferential", "mipmapLevel"]
# name, xml, film size or None (= the XML's own), committed?
AOV_CASES = [
    ("cbox", "cbox/cbox.xml", (128, 128), True),
    ("veach_mi", "veach_mi/mi.xml", (192, 128), True),       # sphere lights: curvature / normals of sphere.inl
    ("sponza", "sponza/sponza.xml", None, False),            # textured: mipmapLevel, uv_screen_size
    ("matpreview", "matpreview/matpreview.xml", None, False),
]

# ---- round 2: Box / Tent pixel filters (filters/box.inl, tent.inl; pixel_filter_test.xml:23-24) ----
# variant name, the <rfilter> element that replaces the gaussian one
FILTER_CASES = [
    ("pixel_filter_box", '<rfilter type="box"/>'),
    ("pixel_filter_tent", '<rfilter type="tent"/>'),
    ("pixel_filter_box3", '<rfilter type="box"><float name="width" value="3"/></rfilter>'),
    ("pixel_filter_tent3", '<rfilter type="tent"><float name="width" value="3.5"/></rfilter>'),
]

# ---- round 2: ray-cast seam against the reference's intersect() / occluded() (ref_driver hits) ----
SEAM_CASES = [
    ("cbox_small", "cbox/cbox.xml", (128, 128), True),
    ("veach_mi", "veach_mi/mi.xml", None, False),
    ("sponza", "sponza/sponza.xml", None, False),
    ("matpreview", "matpreview/matpreview.xml", None, False),
]
SEAM_REC = np.dtype([("org", "<f4", 3), ("tnear", "<f4"), ("dir", "<f4", 3), ("tfar", "<f4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4"),
                     ("geom", "<i4"), ("prim", "<i4"), ("ng", "<f4", 3), ("valid", "<i4"), ("occl", "<i4")])


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



def variant_xml(xml_rel, edit):
    """A temp copy of a scene folder (symlinks to every asset) whose XML went through `edit(text) -> text`:
    the reference's parser then produces the variant exactly as it would from a hand-edited file."""
    import tempfile
    src_dir = os.path.dirname(os.path.join(REF, "scenes", xml_rel))
    tmp = tempfile.mkdtemp(prefix="lj_variant_")
    for e in os.listdir(src_dir):
        os.symlink(os.path.join(src_dir, e), os.path.join(tmp, e))
    xml = os.path.join(tmp, "variant_" + os.path.basename(xml_rel))
    text = open(os.path.join(REF, "scenes", xml_rel)).read()
    new = edit(text)
    assert new != text, "variant edit did not change the XML"
    open(xml, "w").write(new)
    return xml


def render_oracle(xml_path, mode, spp, res=None):
    pfm = f"/tmp/lj_golden_{os.getpid()}.pfm"
    cmd = [DRIVER, "render", xml_path, "-o", pfm, "--mode", mode, "--spp", str(spp)]
    if res:
        cmd += ["--res", str(res[0]), str(res[1])]
    out = run(cmd)
    stats = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    img = read_pfm(pfm)
    os.remove(pfm)
    return img, stats


def load_manifest():
    p = os.path.join(ROOT, "tests/golden/manifest.json")
    return json.load(open(p)) if os.path.exists(p) else {}


def save_manifest(m):
    json.dump(m, open(os.path.join(ROOT, "tests/golden/manifest.json"), "w"), indent=1, sort_keys=True)


# second manifest: the round-2 fixture kinds (full-size images, AOVs, seam dumps), keyed by kind then scene
def load_manifest2():
    p = os.path.join(ROOT, "tests/golden/manifest_r02.json")
    return json.load(open(p)) if os.path.exists(p) else {}


def save_manifest2(m):
    json.dump(m, open(os.path.join(ROOT, "tests/golden/manifest_r02.json"), "w"), indent=1, sort_keys=True)


def full_goldens(only, missing_only=False):
    """python tools/make_goldens.py --full [names]: oracle images at the XML's own film size."""
    manifest = load_manifest2()
    flat = os.path.join(ROOT, "scenes_flat")
    os.makedirs(flat, exist_ok=True)
    sec = manifest.setdefault("full", {})
    for name, xml, renders in FULL_CASES:
        if only and name not in only:
            continue
        xml_path = os.path.join(REF, "scenes", xml)
        if name in FULL_ONLY_BLOBS and not (missing_only and os.path.exists(os.path.join(flat, name + ".ljs"))):
            print(run([FLATTEN, xml_path, os.path.join(flat, name + ".ljs")]).strip().splitlines()[-1])
        entry = sec.setdefault(name, {"xml": xml, "images": {}})
        for mode, spp in renders:
            key = f"{mode}{spp}"
            out = os.path.join(flat, f"{name}_full_{key}.npy")
            if missing_only and os.path.exists(out) and key in entry["images"]:
                continue
            img, stats = render_oracle(xml_path, mode, spp)
            np.save(out, img)
            entry["res"] = [stats["width"], stats["height"]]
            entry["images"][key] = {"intersect_rays": stats["intersect_rays"], "occluded_rays": stats["occluded_rays"], "mean": float(img.mean())}
            print(f"  full {name} {key} {stats['width']}x{stats['height']}: mean {img.mean():.6f}, {stats['render_s']:.1f} s", flush=True)
            save_manifest2(manifest)


def aov_goldens(only, missing_only=False):
    """--aov: the five debug integrators through the reference's render() (aux_render), via a temp copy of the XML
    whose <integrator> element names the AOV."""
    import re
    manifest = load_manifest2()
    sec = manifest.setdefault("aov", {})
    for name, xml, res, committed in AOV_CASES:
        if only and name not in only:
            continue
        outdir = os.path.join(ROOT, "tests/golden" if committed else "scenes_flat")
        os.makedirs(outdir, exist_ok=True)
        out = os.path.join(outdir, f"{name}_aov.npz")
        if missing_only and os.path.exists(out):
            continue
        imgs = {}
        for t in AOV_TYPES:
            def edit(text, t=t):
                # replace the (first, active) <integrator ...> ... </integrator> element or self-closing tag
                text2, n = re.subn(r'<integrator type="[A-Za-z]+"\s*>.*?</integrator>', f'<integrator type="{t}"/>', text, count=1, flags=re.S)
                if n == 0:
                    text2, n = re.subn(r'<integrator type="[A-Za-z]+"\s*/>', f'<integrator type="{t}"/>', text, count=1)
                return text2
            vx = variant_xml(xml, edit)
            img, stats = render_oracle(vx, "A", 1, res)
            imgs[t] = img
            print(f"  aov {name} {t} {stats['width']}x{stats['height']}: mean {np.nanmean(img):.6f}", flush=True)
        np.savez_compressed(out, **imgs)
        sec[name] = {"xml": xml, "res": [int(imgs["depth"].shape[1]), int(imgs["depth"].shape[0])], "committed": committed}
        save_manifest2(manifest)


def filter_goldens(only, missing_only=False):
    """--filters: pixel_filter_test.xml with its <rfilter> swapped for box / tent (the variants the XML keeps commented out)."""
    import re
    manifest = load_manifest()
    gold = os.path.join(ROOT, "tests/golden")
    res = (128, 128)
    for name, element in FILTER_CASES:
        if only and name not in only:
            continue
        if missing_only and os.path.exists(os.path.join(gold, name + ".ljs")) and name in manifest:
            continue
        def edit(text):
            return re.sub(r'<rfilter type="gaussian">.*?</rfilter>', element, text, count=1, flags=re.S)
        vx = variant_xml("pixel_filter_test/pixel_filter_test.xml", edit)
        print(run([FLATTEN, vx, os.path.join(gold, name + ".ljs")]).strip().splitlines()[-1])
        entry = {"xml": "pixel_filter_test/pixel_filter_test.xml", "variant": element, "res": list(res), "committed": True, "images": {}}
        for mode, spp in [("A", 1), ("B", 1), ("B", 64)]:
            img, stats = render_oracle(vx, mode, spp, res)
            np.save(os.path.join(gold, f"{name}_{mode}{spp}.npy"), img)
            entry["images"][f"{mode}{spp}"] = {"intersect_rays": stats["intersect_rays"], "occluded_rays": stats["occluded_rays"], "mean": float(img.mean())}
            print(f"  {name} {mode}{spp}: mean {img.mean():.6f}")
        manifest[name] = entry
        save_manifest(manifest)


def seam_goldens(only, missing_only=False):
    """--seam: ref_driver hits -> compact .npz per scene (valid records only, one array per field and ray set)."""
    manifest = load_manifest2()
    sec = manifest.setdefault("seam", {})
    for name, xml, res, committed in SEAM_CASES:
        if only and name not in only:
            continue
        outdir = os.path.join(ROOT, "tests/golden" if committed else "scenes_flat")
        os.makedirs(outdir, exist_ok=True)
        out = os.path.join(outdir, f"{name}_seam.npz")
        if missing_only and os.path.exists(out):
            continue
        raw_path = f"/tmp/lj_seam_{os.getpid()}.bin"
        cmd = [DRIVER, "hits", os.path.join(REF, "scenes", xml), "-o", raw_path]
        if res:
            cmd += ["--res", str(res[0]), str(res[1])]
        print("  " + run(cmd).strip().splitlines()[-1], flush=True)
        raw = open(raw_path, "rb").read()
        os.remove(raw_path)
        w, h, version, nsets = np.frombuffer(raw, dtype="<i4", count=4)
        assert version == 3 and nsets == 3
        sets = np.frombuffer(raw, dtype=SEAM_REC, offset=16).reshape(3, h * w)
        arrays = {}
        for si, sn in enumerate(("primary", "bounce", "shadow")):
            recs = sets[si][sets[si]["valid"] != 0]
            for f in ("org", "tnear", "dir", "tfar"):
                arrays[f"{sn}_{f}"] = recs[f]
            if sn == "shadow":
                arrays["shadow_occl"] = recs["occl"].astype(np.uint8)
            else:
                for f in ("t", "u", "v", "geom", "prim", "ng"):
                    arrays[f"{sn}_{f}"] = recs[f]
        np.savez_compressed(out, **arrays)
        blob = name[:-6] if name.endswith("_small") else name
        sec[name] = {"xml": xml, "blob": blob, "res": [int(w), int(h)], "committed": committed,
                     "counts": {sn: int((sets[si]["valid"] != 0).sum()) for si, sn in enumerate(("primary", "bounce", "shadow"))}}
        print(f"  seam {name}: {os.path.getsize(out) / 1e6:.1f} MB")
        save_manifest2(manifest)


def mint_missing():
    """--missing: everything that is git-ignored (or not minted yet) and absent -- called by __graft_entry__.build()."""
    manifest = load_manifest()
    miss = [n for n, e in manifest.items() if not e.get("committed", True)
            and not os.path.exists(os.path.join(ROOT, "scenes_flat", n + ".ljs"))]
    if miss:
        base_goldens(set(miss))
    full_goldens(set(), missing_only=True)
    aov_goldens(set(), missing_only=True)
    filter_goldens(set(), missing_only=True)
    seam_goldens(set(), missing_only=True)



def main():
    if len(sys.argv) > 1 and sys.argv[1].startswith("--"):
        fn = {"--rmse": rmse_goldens, "--full": full_goldens, "--aov": aov_goldens, "--filters": filter_goldens, "--seam": seam_goldens}.get(sys.argv[1])
        if sys.argv[1] == "--missing":
            return mint_missing()
        if fn is None:
            raise SystemExit(f"unknown option {sys.argv[1]}")
        return fn(set(sys.argv[2:]))
    return base_goldens(set(sys.argv[1:]))


def base_goldens(only):
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
