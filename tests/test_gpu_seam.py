"""Ray-cast seam against the ORACLE (not against the repo's own binary tree): the f32 rays the reference handed to
rtcIntersect1 / rtcOccluded1 through its own intersect() / occluded() (intersection.cpp:7-85) -- pixel-centre
primary rays, one incoherent bounce ray and one NEE shadow ray per pixel, dumped by `ref_driver hits`
(oracle/ref_driver.cpp) at the configs' own film size -- go through lj_intersect_batch and must return the same
geomID, primID and bit-identical t (tfar), u, v, Ng; occlusion must agree ray by ray."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, FLAT

pytestmark = pytest.mark.gpu
M2 = json.load(open(os.path.join(GOLDEN, "manifest_r02.json")))["seam"]


def rays_of(z, which):
    from lajolla_public_b200.api import RAY_DTYPE
    n = z[which + "_tnear"].shape[0]
    rays = np.zeros(n, dtype=RAY_DTYPE)
    for f in ("org", "tnear", "dir", "tfar"):
        rays[f] = z[f"{which}_{f}"]
    return rays


@pytest.mark.parametrize("name", list(M2))
def test_seam_matches_reference(gpu, name):
    e = M2[name]
    d = GOLDEN if e["committed"] else FLAT
    npz = os.path.join(d, name + "_seam.npz")
    blob = None
    for dd in (GOLDEN, FLAT):
        if os.path.exists(os.path.join(dd, e["blob"] + ".ljs")):
            blob = os.path.join(dd, e["blob"] + ".ljs")
    if blob is None or not os.path.exists(npz):
        pytest.skip("fixture not present (big fixtures live in git-ignored scenes_flat/)")
    z = np.load(npz)
    scene = gpu.load_scene(blob)
    try:
        for which in ("primary", "bounce"):
            rays = rays_of(z, which)
            hits = scene.intersect(rays)
            ref_geom = z[which + "_geom"]
            nhit = int((ref_geom >= 0).sum())
            assert np.array_equal(hits["geom_id"], ref_geom), f"{which}: geomID differs on {(hits['geom_id'] != ref_geom).sum()} rays"
            assert np.array_equal(hits["prim_id"], z[which + "_prim"]), f"{which}: primID differs"
            hit = ref_geom >= 0
            for f in ("t", "u", "v", "ng"):
                a, b = np.ascontiguousarray(hits[f][hit]), np.ascontiguousarray(z[f"{which}_{f}"][hit])
                same = a.view(np.uint32) == b.view(np.uint32)
                assert same.all(), f"{which}: {f} differs on {(~same).sum()} of {nhit} hits (max abs {np.abs(a - b).max():.3e})"
            print(f"{name} {which}: {len(rays)} rays, {nhit} hits: geomID, primID, t, u, v, Ng identical")
        rays = rays_of(z, "shadow")
        occ = scene.intersect(rays, any_hit=True)
        got = np.isneginf(occ["t"])
        ref = z["shadow_occl"] != 0
        assert np.array_equal(got, ref), f"shadow: occlusion differs on {(got != ref).sum()} of {len(rays)} rays"
        print(f"{name} shadow: {len(rays)} rays, {int(ref.sum())} occluded: identical")
    finally:
        scene.close()
