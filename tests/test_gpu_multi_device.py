"""Multi-GPU inside the product: lj_init_devices + lj_render on a context that spans several
devices of the box -- scene replicated, one host thread per device, tile or sample-slice partition, ONE ncclReduce of
the f64 sums (lajolla_public_b200/csrc/lj_multi.cu).  Needs >= 2 GPUs (gpurun --gpus 2); skipped on a one-GPU box,
where tests/test_gpu_partition.py checks the same partitions rank by rank."""
import os
import subprocess

import numpy as np
import pytest

from conftest import fixture_dir, ROOT

pytestmark = pytest.mark.gpu


def gpu_count():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(gpu_count() < 2, reason="needs at least 2 GPUs")


@pytest.fixture(scope="module")
def multi():
    from lajolla_public_b200 import LajollaGPU
    g = LajollaGPU("all")
    yield g
    g.close()


@needs2
@pytest.mark.parametrize("name,res,spp", [("cbox", (128, 128), 8), ("sponza", None, 4), ("vol_cbox", (96, 96), 6), ("hetvol", (96, 72), 4)])
def test_tiles_over_devices_bit_identical(gpu, multi, name, res, spp):
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS, PARTITION_TILES
    d = fixture_dir(name)
    if d is None:
        pytest.skip("fixture missing")
    blob = os.path.join(d, name + ".ljs")
    one = gpu.load_scene(blob, res=res)
    many = multi.load_scene(blob, res=res)
    try:
        ref, st1 = one.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS))
        img, stn = many.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, partition=PARTITION_TILES))
        print(f"{name}: {stn.n_devices} devices, tiles: max |N-way - 1-way| = {np.abs(img - ref).max():.3e}; rays {stn.rays} vs {st1.rays}; "
              f"device ms {stn.device_ms_min:.2f}..{stn.device_ms_max:.2f}, reduce {stn.reduce_ms:.3f} ms, one device {st1.render_ms:.2f} ms")
        assert stn.n_devices == multi.n_devices >= 2
        assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))
        assert stn.rays == st1.rays
    finally:
        one.close(); many.close()


@needs2
@pytest.mark.parametrize("name,res,spp", [("cbox", (128, 128), 16), ("sponza", (192, 144), 16)])
def test_spp_slices_over_devices(gpu, multi, name, res, spp):
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS, PARTITION_SPP
    d = fixture_dir(name)
    if d is None:
        pytest.skip("fixture missing")
    blob = os.path.join(d, name + ".ljs")
    one = gpu.load_scene(blob, res=res)
    many = multi.load_scene(blob, res=res)
    try:
        ref, st1 = one.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS))
        img, stn = many.render(RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, partition=PARTITION_SPP))
        ulp = np.abs(img.view(np.int32).astype(np.int64) - ref.view(np.int32).astype(np.int64))
        print(f"{name}: {stn.n_devices} devices, spp slices: max {int(ulp.max())} ulp, rays {stn.rays} vs {st1.rays}, samples {stn.samples}")
        assert stn.partition == PARTITION_SPP
        assert stn.rays == st1.rays and stn.samples == st1.samples
        assert int(ulp.max()) <= 1
    finally:
        one.close(); many.close()


@needs2
def test_mode_a_and_aov_over_devices(gpu, multi):
    """per-tile streams and an RNG-free AOV under the tile partition across devices"""
    from lajolla_public_b200 import RenderParams, RNG_TILE_STREAMS
    blob = os.path.join(fixture_dir("cbox"), "cbox.ljs")
    for integ in (None, 1):
        one = gpu.load_scene(blob, res=(128, 128), integrator=integ)
        many = multi.load_scene(blob, res=(128, 128), integrator=integ)
        try:
            ref, _ = one.render(RenderParams(spp=2, rng_mode=RNG_TILE_STREAMS))
            img, stn = many.render(RenderParams(spp=2, rng_mode=RNG_TILE_STREAMS))
            assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))
        finally:
            one.close(); many.close()


@needs2
def test_cli_uses_all_devices_byte_identical(tmp_path):
    """`./lajolla scene.xml` on the multi-GPU box: all visible devices by default, LJ_DEVICE pins one; the PFM files
    are byte-identical (tile partition)."""
    exe = os.path.join(ROOT, "lajolla_public_b200", "bin", "lajolla")
    xml = os.path.join(ROOT, "oracle", "_ref", "scenes", "cbox", "cbox.xml")
    if not (os.path.exists(exe) and os.path.exists(xml)):
        pytest.skip("drop-in binary / scene assets not present")
    outs = []
    for env_extra in ({"LJ_DEVICE": "0"}, {}):
        out = tmp_path / f"cbox_{len(outs)}.pfm"
        env = dict(os.environ, LJ_STATS="1", **env_extra)
        env.pop("LJ_DEVICE", None) if not env_extra else None
        r = subprocess.run([exe, "-o", str(out), xml], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, cwd=str(tmp_path))
        assert r.returncode == 0, r.stdout[-2000:]
        print([l for l in r.stdout.splitlines() if "lajolla_b200" in l][-1])
        outs.append(open(out, "rb").read())
    assert outs[0] == outs[1]


def test_init_devices_argument_errors():
    """lj_init_devices refuses ordinals out of range and duplicates; one listed device gives a plain one-device context"""
    import ctypes as C
    from lajolla_public_b200 import LajollaGPU, LjError
    with pytest.raises(LjError, match="out of range"):
        LajollaGPU([0, 999])
    with pytest.raises(LjError, match="listed twice"):
        LajollaGPU([0, 0])
    g = LajollaGPU([0])
    try:
        assert g.n_devices == 1
    finally:
        g.close()
