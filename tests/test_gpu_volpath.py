"""GPU parity of the volumetric integrators (vol_path_tracing.h: versions 1, 2 and the final
integrator that versions 3-6 forward to) against the oracle's images on identical pcg32 streams.

Bars: 1 spp, fraction of pixels with |gpu - ref| <= 1e-4 * max(1, |ref|) on every
channel >= 0.90 (one flipped free-flight decision -- exp/log/pow differ in the last bit between
CUDA's libm and glibc -- changes that pixel and, in mode A, de-synchronises the rest of its tile);
16 spp on per-sample streams: image mean within 2 %, pixel match >= 0.85; ray counts reported."""
import os

import numpy as np
import pytest

from conftest import fixture_dir
from test_gpu_parity import match_fraction, tile_fraction

pytestmark = pytest.mark.gpu

SCENES = ["volpath_test1", "volpath_test2", "volpath_test3", "volpath_test4", "volpath_test4_2", "volpath_test5",
          "volpath_test5_2", "volpath_test6", "vol_cbox", "vol_cbox_teapot", "hetvol", "hetvol_colored"]


def load(gpu, manifest, name):
    d = fixture_dir(name)
    if d is None:
        pytest.skip(f"fixture {name}.ljs not present (big fixtures live in git-ignored scenes_flat/)")
    e = manifest[name]
    return d, e, gpu.load_scene(os.path.join(d, name + ".ljs"), res=e["res"])


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("mode", ["A", "B"])
def test_volpath_one_spp(gpu, manifest, name, mode):
    from lajolla_public_b200 import RenderParams, RNG_TILE_STREAMS, RNG_SAMPLE_STREAMS
    d, e, scene = load(gpu, manifest, name)
    ref = np.load(os.path.join(d, f"{name}_{mode}1.npy"))
    img, st = scene.render(RenderParams(spp=1, rng_mode=RNG_TILE_STREAMS if mode == "A" else RNG_SAMPLE_STREAMS))
    scene.close()
    frac = match_fraction(img, ref)
    meta = e["images"][mode + "1"]   # mode B goes through the default engine (wavefront for versions >= 3)
    print(f"{name} mode {mode}: pixel match {frac:.5f}, tile match {tile_fraction(img, ref):.4f}, rays gpu "
          f"{st.rays_primary + st.rays_extension}+{st.rays_shadow} ref {meta['intersect_rays']}+{meta['occluded_rays']}, "
          f"{st.render_ms:.2f} ms")
    assert np.isfinite(img).all()
    assert frac >= 0.90


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("engine", ["wavefront", "megakernel"])
def test_volpath_many_spp_same_streams(gpu, manifest, name, engine):
    """16 spp on per-sample streams through both engines: the staged wavefront engine (lj_volwave.cu;
    versions 1 and 2 stay on the per-lane kernel) and the one-lane-per-path megakernel"""
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS, ENGINE_WAVEFRONT, ENGINE_MEGAKERNEL
    d, e, scene = load(gpu, manifest, name)
    ref = np.load(os.path.join(d, f"{name}_B16.npy"))
    img, st = scene.render(RenderParams(spp=16, rng_mode=RNG_SAMPLE_STREAMS,
                                        engine=ENGINE_WAVEFRONT if engine == "wavefront" else ENGINE_MEGAKERNEL))
    scene.close()
    frac = match_fraction(img, ref)
    rmse = float(np.sqrt(np.mean((img - ref) ** 2)))
    print(f"{name} {engine}: RMSE {rmse:.3e}, pixel match {frac:.5f}, mean gpu {img.mean():.6f} ref {ref.mean():.6f}, "
          f"{st.samples / st.render_ms * 1e-3:.2f} Msamples/s, {st.rays / st.render_ms * 1e-3:.1f} Mrays/s")
    assert frac >= 0.85
    assert abs(float(img.mean()) - float(ref.mean())) <= 0.02 * float(ref.mean())


@pytest.mark.parametrize("name", ["hetvol", "hetvol_colored"])
def test_grey_media_tracking_is_bit_identical(gpu, manifest, name, monkeypatch):
    """Scenes whose media all have equal channels track ONE channel (lj_volpath.cuh, `sp_*` over double): the image must
    be the three-channel kernels' image bit for bit (hetvol is grey; hetvol_colored is not and runs the same kernels twice)."""
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS, ENGINE_WAVEFRONT
    d, e, scene = load(gpu, manifest, name)
    rp = RenderParams(spp=4, rng_mode=RNG_SAMPLE_STREAMS, engine=ENGINE_WAVEFRONT)
    monkeypatch.delenv("LJ_NO_GREY_TRACK", raising=False)
    grey, st_g = scene.render(rp)
    monkeypatch.setenv("LJ_NO_GREY_TRACK", "1")
    full, st_f = scene.render(rp)
    scene.close()
    print(f"{name}: grey {st_g.render_ms:.2f} ms, three channels {st_f.render_ms:.2f} ms, rays {st_g.rays} vs {st_f.rays}")
    assert st_g.rays == st_f.rays
    assert np.array_equal(grey.view(np.uint32), full.view(np.uint32))
