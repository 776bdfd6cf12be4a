"""The five debug AOV integrators (aux_render, /root/reference/src/render.cpp:12-69; selected at
parse_scene.cpp:295-304) on the GPU (k_aux) against the reference's own render() of the same scene with the
<integrator> element swapped (tools/make_goldens.py --aov).  They are RNG-free, so the bar is bit equality
of the f32 images; they expose PathVertex fields directly (intersection.cpp:38-62): distance, shading normal,
mean curvature, the ray differential and uv_screen_size through the mip level."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, FLAT

pytestmark = pytest.mark.gpu

AOV_IDS = {"depth": 0, "shadingNormal": 1, "meanCurvature": 2, "rayDifferential": 3, "mipmapLevel": 4}   # scene.h:14-22
M2 = json.load(open(os.path.join(GOLDEN, "manifest_r02.json")))["aov"]


def ulp_distance(a, b):
    """distance in units of the last place between two f32 arrays (NaN == NaN, +0 == -0)"""
    ia = a.view(np.int32).astype(np.int64); ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia); ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    d = np.abs(ia - ib)
    d[np.isnan(a) & np.isnan(b)] = 0
    return d


@pytest.mark.parametrize("aov", list(AOV_IDS))
@pytest.mark.parametrize("name", list(M2))
def test_aov_bit_exact(gpu, name, aov):
    from lajolla_public_b200 import RenderParams, RNG_TILE_STREAMS
    e = M2[name]
    d = GOLDEN if e["committed"] else FLAT
    npz = os.path.join(d, name + "_aov.npz")
    blob = None
    for dd in (GOLDEN, FLAT):
        if os.path.exists(os.path.join(dd, name + ".ljs")):
            blob = os.path.join(dd, name + ".ljs")
    if blob is None or not os.path.exists(npz):
        pytest.skip("fixture not present (big fixtures live in git-ignored scenes_flat/)")
    ref = np.load(npz)[aov]
    scene = gpu.load_scene(blob, res=e["res"], integrator=AOV_IDS[aov])
    img, st = scene.render(RenderParams(spp=1, rng_mode=RNG_TILE_STREAMS))
    scene.close()
    assert img.shape == ref.shape
    d_ulp = ulp_distance(img, ref)
    bad = int((d_ulp > 0).sum())
    print(f"{name} {aov} {e['res'][0]}x{e['res'][1]}: {bad} of {img.size} values differ, max {int(d_ulp.max())} ulp; "
          f"mean gpu {np.nanmean(img):.6f} ref {np.nanmean(ref):.6f}")
    assert bad == 0
