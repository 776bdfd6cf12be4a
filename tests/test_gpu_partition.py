"""The work partitions of SURVEY 8(e) on the GPU itself ("no N-way-equals-1-way check on a GPU").

One device: for world in {2, 3, 8} every rank's share (the row-rotated round-robin deal of the 16x16 tiles of render.cpp:75-82) is
rendered into its own f64 sums; their sum must equal the unpartitioned render BIT FOR BIT (the accumulators are laid
out so that the per-pixel summation order does not depend on the partition or the pool size, lj_wave_common.cuh
SlotMap).  Sample slices [0,a) [a,b) [b,spp) add up to the full render to within 1 ulp of the f32 image.
Several devices (needs >= 2 GPUs, otherwise skipped): the same through lj_init_devices -- the library's own per-device
threads and ncclReduce -- against the one-device render of the same streams."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import fixture_dir

pytestmark = pytest.mark.gpu

# scene, film override, spp: surface path tracer (sponza at its own film size, several material kinds on veach_mi),
# volumetric wavefront (vol_cbox, hetvol), and the whole-path engines
# (the 150-spp case has more sample lanes than 64 and a sample count that is not a power of two: K0 = 256, K < K0)
CASES = [("pixel_filter_test", (48, 48), 150), ("cbox", (128, 128), 8), ("veach_mi", (192, 128), 6), ("sponza", None, 4), ("vol_cbox", (96, 96), 6), ("hetvol", (96, 72), 5)]


def device_sums(scene, rp):
    import torch
    sums = torch.full((scene.height * scene.width * 3,), float("nan"), dtype=torch.float64, device="cuda:0")
    st = scene.render_into_device(sums.data_ptr(), rp)
    torch.cuda.synchronize()
    return sums, st


def load(gpu, name, res):
    d = fixture_dir(name)
    if d is None:
        pytest.skip(f"fixture {name}.ljs not present (big fixtures live in git-ignored scenes_flat/)")
    return gpu.load_scene(os.path.join(d, name + ".ljs"), res=res)


@pytest.mark.parametrize("engine", ["wavefront", "megakernel"])
@pytest.mark.parametrize("name,res,spp", CASES)
def test_tile_partition_is_bit_identical(gpu, name, res, spp, engine):
    import torch
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS, ENGINE_MEGAKERNEL, ENGINE_WAVEFRONT
    from lajolla_public_b200.distributed import tile_owner_mask
    if engine == "megakernel" and name in ("sponza", "hetvol"):
        pytest.skip("whole-path engine covered on the small scenes")
    eng = ENGINE_MEGAKERNEL if engine == "megakernel" else ENGINE_WAVEFRONT
    scene = load(gpu, name, res)
    try:
        full, st_full = device_sums(scene, RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, engine=eng))
        assert bool(torch.isfinite(full).all())
        for world in (2, 3, 8):
            total = torch.zeros_like(full)
            rays = 0
            for r in range(world):
                part, st = device_sums(scene, RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, engine=eng, tile_rank=r, tile_world=world))
                # a rank leaves the other ranks' pixels at exactly zero, so adding is exact; its own pixels are the ones the
                # documented deal gives it (distributed.tile_owner_mask restates the library's tile_dealt_index)
                touched = (part.reshape(scene.height, scene.width, 3) != 0).any(dim=2).cpu().numpy()
                assert not (touched & ~tile_owner_mask(scene.width, scene.height, r, world)).any()
                total += part
                rays += st.rays
            same = bool(torch.equal(total, full))
            diff = float((total - full).abs().max())
            print(f"{name} {engine} world {world}: bit-identical {same} (max |sum of ranks - full| = {diff:.3e}), rays {rays} vs {st_full.rays}")
            assert same
            assert rays == st_full.rays
    finally:
        scene.close()


@pytest.mark.parametrize("name,res,spp", [("cbox", (128, 128), 4), ("pixel_filter_test", (128, 128), 2)])
def test_tile_partition_mode_a(gpu, name, res, spp):
    """the reference's own per-tile streams (mode A, tile-serial engine) under the tile partition"""
    import torch
    from lajolla_public_b200 import RenderParams, RNG_TILE_STREAMS
    scene = load(gpu, name, res)
    try:
        full, _ = device_sums(scene, RenderParams(spp=spp, rng_mode=RNG_TILE_STREAMS))
        total = torch.zeros_like(full)
        for r in range(3):
            part, _ = device_sums(scene, RenderParams(spp=spp, rng_mode=RNG_TILE_STREAMS, tile_rank=r, tile_world=3))
            total += part
        assert bool(torch.equal(total, full))
    finally:
        scene.close()


@pytest.mark.parametrize("name,res,spp", [("cbox", (128, 128), 16), ("sponza", (192, 144), 12), ("vol_cbox", (96, 96), 10)])
def test_sample_slices_add_up(gpu, name, res, spp):
    import torch
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS
    scene = load(gpu, name, res)
    try:
        full, st_full = device_sums(scene, RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS))
        a, b = spp // 3, (2 * spp) // 3 + 1
        total = torch.zeros_like(full)
        rays = 0
        for lo, hi in ((0, a), (a, b), (b, spp)):
            part, st = device_sums(scene, RenderParams(spp=spp, rng_mode=RNG_SAMPLE_STREAMS, sample_begin=lo, sample_end=hi))
            total += part
            rays += st.rays
        img = (total * (1.0 / spp)).to(torch.float32).cpu().numpy()
        ref = (full * (1.0 / spp)).to(torch.float32).cpu().numpy()
        ulp = np.abs(img.view(np.int32).astype(np.int64) - ref.view(np.int32).astype(np.int64))
        print(f"{name}: slices [0,{a}) [{a},{b}) [{b},{spp}): max {int(ulp.max())} ulp in the f32 image, {int((ulp > 0).sum())} of {ulp.size} values differ, "
              f"rays {rays} vs {st_full.rays}")
        assert rays == st_full.rays                     # same paths, only the summation order differs
        assert int(ulp.max()) <= 1
    finally:
        scene.close()


def test_slice_needs_sample_streams(gpu):
    """a sample slice combined with per-tile streams (or an AOV integrator) is an error, not a brighter image"""
    from lajolla_public_b200 import RenderParams, RNG_TILE_STREAMS, LjError
    scene = load(gpu, "cbox", (64, 64))
    try:
        with pytest.raises(LjError, match="sample slice"):
            scene.render(RenderParams(spp=4, rng_mode=RNG_TILE_STREAMS, sample_begin=0, sample_end=2))
    finally:
        scene.close()
