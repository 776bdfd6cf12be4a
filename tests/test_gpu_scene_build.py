"""Scene build step on the device (lj_scene_create_ex + LJ_SCENE_BUILD_TABLES) against the tables the
reference's own Scene::Scene built on the host (they are what the flattened blobs carry): mip levels >= 1, per-mesh
area pmf/cdf + total_area, the envmap TableDist2D + total_values, the light distribution.  The derived ranges are
overwritten with NaN on the device before the kernels run (LJ_SCENE_POISON_DERIVED), so equality proves the kernels
wrote every value.  Bar: bit-exact, except values downstream of the envmap's sin(pi v) (device sin vs glibc sin):
relative 1e-13."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN, FLAT

pytestmark = pytest.mark.gpu

BLOBS = sorted(glob.glob(os.path.join(GOLDEN, "*.ljs")) + glob.glob(os.path.join(FLAT, "*.ljs")))
NAMES = [os.path.splitext(os.path.basename(b))[0] for b in BLOBS]


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float64).view(np.uint64)


@pytest.mark.parametrize("blob", BLOBS, ids=NAMES)
def test_device_built_tables_match_the_reference(gpu, blob):
    from lajolla_public_b200.api import HostSceneDesc, SCENE_BUILD_TABLES, SCENE_POISON_DERIVED
    desc = HostSceneDesc(blob)
    ref = desc.arrays()
    sc = gpu.scene_from_desc(desc, flags=SCENE_BUILD_TABLES | SCENE_POISON_DERIVED)
    tables, texels, shapes, lights = sc.read_derived(len(ref["tables"]), len(ref["texels"]), len(ref["shapes"]), len(ref["lights"]))
    sc.close()
    # mip chains.  Levels below an image dimension of 1 are excluded: there the reference's box filter reads past
    # the end of the image (mipmap.h:38-41 with 2y+1 >= height), so the blob holds whatever was on the host heap.
    undefined = np.zeros(len(texels), dtype=bool)
    for cls in ("images1", "images3"):
        for im in ref[cls]:
            bad = False
            for l in range(1, int(im["num_levels"])):
                bad = bad or 2 * int(im["width"][l]) > int(im["width"][l - 1]) or 2 * int(im["height"][l]) > int(im["height"][l - 1])
                if bad:
                    o = int(im["offset"][l])
                    undefined[o:o + int(im["width"][l]) * int(im["height"][l]) * int(im["channels"])] = True
    assert np.array_equal(bits(texels)[~undefined], bits(ref["texels"])[~undefined]), "mip levels differ"
    # mesh tables
    n_mesh = 0
    for i, sh in enumerate(ref["shapes"]):
        if sh["kind"] != 1:
            continue
        n_mesh += 1
        n = int(sh["num_triangles"])
        for off, cnt, what in ((int(sh["tri_pmf_offset"]), n, "pmf"), (int(sh["tri_cdf_offset"]), n + 1, "cdf")):
            assert np.array_equal(bits(tables[off:off + cnt]), bits(ref["tables"][off:off + cnt])), f"shape {i} {what}"
        assert bits(shapes[i]["total_area"]) == bits(sh["total_area"]), f"shape {i} total_area"
    # envmap distribution
    has_env = False
    for i, l in enumerate(ref["lights"]):
        if l["kind"] != 1 or l["dist_width"] <= 0:
            continue
        has_env = True
        w, h = int(l["dist_width"]), int(l["dist_height"])
        for key, cnt in (("cdf_rows_offset", h * (w + 1)), ("pdf_rows_offset", h * w), ("cdf_marginals_offset", h + 1), ("pdf_marginals_offset", h)):
            off = int(l[key])
            np.testing.assert_allclose(tables[off:off + cnt], ref["tables"][off:off + cnt], rtol=1e-13, atol=0, err_msg=f"light {i} {key}")
        np.testing.assert_allclose(lights[i]["total_values"], l["total_values"], rtol=1e-13)
    # light distribution
    nl = len(ref["lights"])
    for off, cnt in ((ref["light_pmf_offset"], nl), (ref["light_cdf_offset"], nl + 1)):
        if has_env:
            np.testing.assert_allclose(tables[off:off + cnt], ref["tables"][off:off + cnt], rtol=1e-13)
        else:
            assert np.array_equal(bits(tables[off:off + cnt]), bits(ref["tables"][off:off + cnt])), "light distribution"
    assert not np.isnan(tables).any() and not np.isnan(texels).any()
    desc.close()
    print(f"{os.path.basename(blob)}: {n_mesh} meshes, {len(ref['texels'])} texels, {len(ref['tables'])} table entries, envmap {has_env}")


@pytest.mark.parametrize("name", ["cbox", "volpath_test5"])
def test_render_of_a_device_built_scene_is_unchanged(gpu, manifest, name):
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS
    from lajolla_public_b200.api import HostSceneDesc, SCENE_BUILD_TABLES, SCENE_POISON_DERIVED
    from conftest import fixture_dir
    d = fixture_dir(name)
    if d is None or "B1" not in manifest.get(name, {}).get("images", {}):
        pytest.skip("fixture missing")
    desc = HostSceneDesc(os.path.join(d, name + ".ljs"), res=manifest[name]["res"])
    sc = gpu.scene_from_desc(desc, flags=SCENE_BUILD_TABLES | SCENE_POISON_DERIVED)
    img, _ = sc.render(RenderParams(spp=1, rng_mode=RNG_SAMPLE_STREAMS))
    sc.close(); desc.close()
    ref = np.load(os.path.join(d, f"{name}_B1.npy"))
    assert np.array_equal(img, ref)
