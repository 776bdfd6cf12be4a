"""GPU parity: CUDA path (through the C ABI) vs the oracle's images of the same scene,
same film size, same pcg32 streams.  Oracle = unmodified reference sources + BVH shim
(oracle/), images minted by tools/make_goldens.py.

Thresholds (BASELINE.json north_star):
  * 1 spp: fraction of pixels with |gpu - ref| <= 1e-4 * max(1, |ref|) on every channel.
  * many spp: RMSE(gpu, ref) on identical sample streams (mode B), expected ~0, and
    RMSE between independent streams bounded by the Monte Carlo noise floor.
"""
import os

import numpy as np
import pytest

from conftest import fixture_dir

pytestmark = pytest.mark.gpu

TOL = 1e-4


def match_fraction(img, ref):
    ok = np.abs(img - ref) <= TOL * np.maximum(1.0, np.abs(ref))
    return float(np.all(ok, axis=2).mean())


def tile_fraction(img, ref):
    ok = np.all(np.abs(img - ref) <= TOL * np.maximum(1.0, np.abs(ref)), axis=2)
    h, w = ok.shape
    tiles = [ok[y:y + 16, x:x + 16].all() for y in range(0, h, 16) for x in range(0, w, 16)]
    return float(np.mean(tiles))


def load_case(gpu, manifest, name):
    d = fixture_dir(name)
    if d is None:
        pytest.skip(f"fixture {name}.ljs not present (big fixtures live in git-ignored scenes_flat/)")
    entry = manifest[name]
    scene = gpu.load_scene(os.path.join(d, name + ".ljs"), res=entry["res"])
    return d, entry, scene


# stated 1-spp pixel-match bars per scene (mode A = reference tile streams, mode B = per-sample streams)
BARS = {"cbox": 0.99, "veach_mi": 0.99, "sponza": 0.99, "matpreview": 0.95, "simple_sphere": 0.99, "pixel_filter_test": 0.99,
        "disney_diffuse": 0.95, "disney_metal": 0.95, "disney_glass": 0.95, "disney_clearcoat": 0.95, "disney_sheen": 0.95,
        "disney_bsdf": 0.95,
        # Box / Tent pixel filters (filters/box.inl, tent.inl): pixel_filter_test.xml with the <rfilter> variants it keeps
        # commented out (:23-24), default widths and explicit ones
        "pixel_filter_box": 0.99, "pixel_filter_tent": 0.99, "pixel_filter_box3": 0.99, "pixel_filter_tent3": 0.99}
PATH_SCENES = list(BARS)


@pytest.mark.parametrize("name", PATH_SCENES)
@pytest.mark.parametrize("mode", ["A", "B"])
def test_one_spp_path_match(gpu, manifest, name, mode):
    from lajolla_public_b200 import RenderParams, RNG_TILE_STREAMS, RNG_SAMPLE_STREAMS, ENGINE_MEGAKERNEL
    d, entry, scene = load_case(gpu, manifest, name)
    if f"{mode}1" not in entry["images"]:
        pytest.skip("no golden")
    ref = np.load(os.path.join(d, f"{name}_{mode}1.npy"))
    rp = RenderParams(spp=1, rng_mode=RNG_TILE_STREAMS if mode == "A" else RNG_SAMPLE_STREAMS, engine=ENGINE_MEGAKERNEL)
    img, st = scene.render(rp)
    scene.close()
    frac = match_fraction(img, ref)
    print(f"{name} mode {mode}: pixel match {frac:.5f}, tile match {tile_fraction(img, ref):.4f}, "
          f"rays gpu {st.rays_primary + st.rays_extension}+{st.rays_shadow} ref "
          f"{entry['images'][mode + '1']['intersect_rays']}+{entry['images'][mode + '1']['occluded_rays']}")
    assert frac >= BARS[name]


@pytest.mark.parametrize("name", PATH_SCENES)
@pytest.mark.parametrize("engine", ["megakernel", "wavefront"])
def test_many_spp_same_streams(gpu, manifest, name, engine):
    """64 spp on identical per-sample streams: images must agree far below the noise level."""
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS, ENGINE_MEGAKERNEL, ENGINE_WAVEFRONT
    d, entry, scene = load_case(gpu, manifest, name)
    if "B64" not in entry["images"]:
        pytest.skip("no golden")
    ref = np.load(os.path.join(d, f"{name}_B64.npy"))
    rp = RenderParams(spp=64, rng_mode=RNG_SAMPLE_STREAMS, engine=ENGINE_MEGAKERNEL if engine == "megakernel" else ENGINE_WAVEFRONT)
    img, st = scene.render(rp)
    scene.close()
    rmse = float(np.sqrt(np.mean((img - ref) ** 2)))
    frac = match_fraction(img, ref)
    print(f"{name} {engine}: RMSE {rmse:.3e}, pixel match {frac:.5f}, mean gpu {img.mean():.6f} ref {ref.mean():.6f}")
    # a pixel differs only if one of its 64 paths took a different discrete decision
    assert frac >= 0.97
    assert abs(float(img.mean()) - float(ref.mean())) <= 0.01 * float(ref.mean())


RMSE_SCENES = ["cbox", "veach_mi", "sponza", "matpreview", "disney_bsdf", "volpath_test3", "vol_cbox", "hetvol"]


@pytest.mark.parametrize("name", RMSE_SCENES)
def test_1024spp_rmse_vs_reference_render(gpu, manifest, name):
    """BASELINE.json north star, threshold (ii), one case per config: at 1024 spp the image RMSE against the
    reference's own CPU render (mode A = the reference's render(), per-tile streams) is below a stated bound.
    Bound: 2 x the noise floor RMSE(ref_A1024, ref_B1024) of two independent oracle renders of that scene
    (cbox at 128x128: floor 0.0178 => bound 0.0356; the others are printed and recorded in BASELINE.md);
    secondary: relMSE with 0.1 % trimming <= 2 x its floor.  The GPU renders on per-sample streams with the
    wavefront engine, i.e. independent of render A; against render B (identical streams) it must be ~0."""
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS, ENGINE_WAVEFRONT
    d = fixture_dir(name)
    if d is None:
        pytest.skip(f"fixture {name}.ljs not present (big fixtures live in git-ignored scenes_flat/)")
    entry = manifest[name]
    res = entry["res"] if "A1024" in entry["images"] else entry["rmse1024"]["res"]
    scene = gpu.load_scene(os.path.join(d, name + ".ljs"), res=res)
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    ref_a = np.load(os.path.join(gold, f"{name}_A1024.npy"))
    ref_b = np.load(os.path.join(gold, f"{name}_B1024.npy"))
    img, st = scene.render(RenderParams(spp=1024, rng_mode=RNG_SAMPLE_STREAMS, engine=ENGINE_WAVEFRONT))
    scene.close()

    def rmse(x, r):
        return float(np.sqrt(np.mean((x.astype(np.float64) - r) ** 2)))

    def relmse(x, r):
        e = np.sort(((x.astype(np.float64) - r) ** 2 / (r.astype(np.float64) ** 2 + 1e-2)).ravel())
        return float(e[: int(len(e) * 0.999)].mean())

    floor, floor_rel = rmse(ref_a, ref_b), relmse(ref_a, ref_b)
    got, got_rel = rmse(img, ref_a), relmse(img, ref_a)
    same = rmse(img, ref_b)
    print(f"{name} {res[0]}x{res[1]} 1024 spp: RMSE vs reference render {got:.5f} (floor {floor:.5f}, bound {2 * floor:.5f}); "
          f"relMSE {got_rel:.3e} (floor {floor_rel:.3e}); RMSE vs oracle on identical streams {same:.3e}; "
          f"{st.rays / st.render_ms * 1e-3:.0f} Mrays/s, {res[0] * res[1] * 1024 / st.render_ms * 1e-3:.1f} Msamples/s")
    assert got <= 2 * floor
    assert got_rel <= 2 * floor_rel
    # identical streams: only rare flipped decisions (last-ulp differences of libm functions) may differ
    assert same <= max(1e-3 * float(ref_b.mean()), 0.05 * floor)
