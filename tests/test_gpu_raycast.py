"""Ray-cast seam on the GPU (lj_intersect_batch = rtcIntersect1 / rtcOccluded1 for a batch):
the compressed 8-wide BVH must return exactly what the binary tree returns, for closest hits
(t, u, v, ids bit-exact: same primitive arithmetic, order-independent tie rule) and occlusion."""
import os

import numpy as np
import pytest

from conftest import fixture_dir

pytestmark = pytest.mark.gpu


def random_rays(n, lo, hi, seed):
    from lajolla_public_b200.api import RAY_DTYPE
    rng = np.random.default_rng(seed)
    rays = np.zeros(n, dtype=RAY_DTYPE)
    rays["org"] = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    # a share of axis-aligned / zero-component directions (slab tests with infinite inverse)
    d[: n // 16, 0] = 0.0
    d[n // 16: n // 8, 1] = 0.0
    rays["dir"] = d.astype(np.float32)
    rays["tnear"] = 1e-4
    rays["tfar"] = np.where(rng.uniform(size=n) < 0.5, np.inf, rng.uniform(0.1, 2.0, size=n) * np.linalg.norm(hi - lo)).astype(np.float32)
    return rays


def scene_box(blob):
    """bounding sphere stored in the blob header: lj_scene_desc.bounds_radius / bounds_center"""
    with open(blob, "rb") as f:
        hdr = f.read(4 + 400)
    radius, cx, cy, cz = np.frombuffer(hdr, dtype="<f8", count=4, offset=4 + 304)
    c = np.array([cx, cy, cz])
    return c - radius, c + radius


@pytest.mark.parametrize("name", ["cbox", "veach_mi", "sponza"])
def test_bvh8_matches_bvh2(gpu, name):
    d = fixture_dir(name)
    if d is None:
        pytest.skip("fixture missing")
    blob = os.path.join(d, name + ".ljs")
    lo, hi = scene_box(blob)
    rays = random_rays(200_000, lo, hi, seed=7)
    os.environ["LJ_BVH"] = "2"
    s2 = gpu.load_scene(blob)
    os.environ.pop("LJ_BVH")
    s8 = gpu.load_scene(blob)
    try:
        h2, h8 = s2.intersect(rays), s8.intersect(rays)
        o2, o8 = s2.intersect(rays, any_hit=True), s8.intersect(rays, any_hit=True)
    finally:
        s2.close(); s8.close()
    nhit = int((h2["geom_id"] >= 0).sum())
    print(f"{name}: {nhit} of {len(rays)} rays hit; bvh8 nodes {s8.build_stats.bvh_nodes}, build {s8.build_stats.bvh_build_ms:.2f} ms")
    assert nhit > len(rays) // 20
    for f in ("geom_id", "prim_id", "t", "u", "v", "ng"):
        assert np.array_equal(h2[f], h8[f]), f"closest-hit field {f} differs between BVH2 and BVH8"
    assert np.array_equal(o2["t"], o8["t"])
    # occlusion agrees with closest hit: occluded <=> a hit exists within (tnear, tfar]
    assert np.array_equal(np.isneginf(o8["t"]), h8["geom_id"] >= 0)


@pytest.mark.parametrize("name", ["cbox", "sponza", "matpreview"])
def test_cooperative_build_equals_launch_sequence(gpu, name, monkeypatch):
    """The cooperative PLOC / collapse kernels (lj_bvh_build.cu, lj_bvh8_build.cu: the default) and the launch-sequence
    builders behind LJ_PLOC_COOP=0 / LJ_BVH8_COOP=0 implement the same algorithm: same node count, same hits bit for bit
    (closest hit and occlusion), also with the hand-over to the one-block tail moved."""
    d = fixture_dir(name)
    if d is None:
        pytest.skip("fixture missing")
    blob = os.path.join(d, name + ".ljs")
    lo, hi = scene_box(blob)
    rays = random_rays(100_000, lo, hi, seed=11)
    results = {}
    for tag, env in (("cooperative", {}), ("launches", {"LJ_PLOC_COOP": "0", "LJ_BVH8_COOP": "0"}), ("handover64", {"LJ_PLOC_STOP": "64"})):
        for k in ("LJ_PLOC_COOP", "LJ_BVH8_COOP", "LJ_PLOC_STOP"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        sc = gpu.load_scene(blob)
        try:
            results[tag] = (sc.build_stats.bvh_nodes, sc.intersect(rays), sc.intersect(rays, any_hit=True), sc.build_stats.bvh_build_ms)
        finally:
            sc.close()
    n0, h0, o0, ms0 = results["cooperative"]
    print(f"{name}: {n0} nodes; build ms " + ", ".join(f"{k} {v[3]:.2f}" for k, v in results.items()))
    for tag in ("launches", "handover64"):
        n, h, o, _ = results[tag]
        assert n == n0, f"{tag}: {n} nodes against {n0}"
        for f in ("geom_id", "prim_id", "t", "u", "v", "ng"):
            assert np.array_equal(h0[f], h[f]), f"{tag}: closest-hit field {f} differs"
        assert np.array_equal(o0["t"], o["t"])
