"""Output step: lj_image_encode / lj_accum_* against the files the reference's own imwrite()
(image.cpp:135-173) wrote for the same pixels (tests/golden/imwrite_kat.*, minted by tools/make_goldens.py).
PFM must be byte-identical; EXR must decode to the same HALF bit patterns (the reference deflates, we store
the scanlines uncompressed) and carry the same header attributes."""
import os
import struct

import numpy as np
import pytest

from exr_util import half_bits_tinyexr, read_exr_half

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def kat_image():
    raw = open(os.path.join(GOLD, "imwrite_kat.f32img"), "rb").read()
    w, h = struct.unpack_from("<ii", raw)
    return np.frombuffer(raw, dtype="<f4", offset=8).reshape(h, w, 3).copy()


def test_pfm_is_byte_identical_to_the_reference(gpu):
    from lajolla_public_b200.api import IMAGE_PFM
    img = kat_image()
    mine = gpu.encode_image(img, IMAGE_PFM)
    ref = open(os.path.join(GOLD, "imwrite_kat.pfm"), "rb").read()
    assert mine == ref
    assert len(mine) == gpu.lib.lj_image_encoded_size(IMAGE_PFM, img.shape[1], img.shape[0])


def test_exr_decodes_to_the_reference_pixels(gpu):
    from lajolla_public_b200.api import IMAGE_EXR
    img = kat_image()
    mine = gpu.encode_image(img, IMAGE_EXR)
    assert len(mine) == gpu.lib.lj_image_encoded_size(IMAGE_EXR, img.shape[1], img.shape[0])
    ch_m, at_m = read_exr_half(mine)
    ch_r, at_r = read_exr_half(open(os.path.join(GOLD, "imwrite_kat.exr"), "rb").read())
    assert list(ch_m) == list(ch_r) == ["B", "G", "R"]
    for c in "BGR":
        assert np.array_equal(ch_m[c], ch_r[c]), f"channel {c}: {(ch_m[c] != ch_r[c]).sum()} half values differ"
    # same attributes in the same order; only the compression byte differs (0 = none, reference 3 = ZIP)
    assert list(at_m) == list(at_r)
    for k in at_r:
        if k != "compression":
            assert at_m[k] == at_r[k], k
    assert at_m["compression"][1] == b"\0"


def test_exr_half_conversion_large_random(gpu):
    """size-independent check at a film-sized image: every half equals the tinyexr restatement"""
    from lajolla_public_b200.api import IMAGE_EXR
    rng = np.random.default_rng(3)
    h, w = 575, 768
    bits = rng.integers(0, 2 ** 32, size=(h, w, 3), dtype=np.uint64).astype(np.uint32)
    img = bits.view(np.float32)                 # all exponents, denormals, infinities, NaNs
    ch, _ = read_exr_half(gpu.encode_image(img, IMAGE_EXR))
    want = half_bits_tinyexr(img)
    for ci, c in enumerate("RGB"):
        assert np.array_equal(ch[c], want[:, :, ci])


def test_write_by_extension_and_unsupported_format(gpu, tmp_path):
    from lajolla_public_b200 import LjError
    img = kat_image()
    p = tmp_path / "a.pfm"
    gpu.write_image(str(p), img)
    assert p.read_bytes() == open(os.path.join(GOLD, "imwrite_kat.pfm"), "rb").read()
    q = tmp_path / "a.exr"
    gpu.write_image(str(q), img)
    ch, _ = read_exr_half(q.read_bytes())
    assert np.array_equal(ch["R"], half_bits_tinyexr(img)[:, :, 0])
    with pytest.raises(LjError) as ei:
        gpu.write_image(str(tmp_path / "a.png"), img)
    assert "Unsupported image format" in str(ei.value)      # image.cpp:130 wording


def test_accumulation_finalize_encode_checkpoint(gpu, tmp_path):
    """device f64 sums -> f32 image / encoded file without a host round trip; checkpoint save / load; a render
    resumed from a checkpoint equals the one-pass render up to f64 summation order"""
    import torch
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS
    from lajolla_public_b200.api import IMAGE_PFM, IMAGE_EXR
    from conftest import fixture_dir
    sc = gpu.load_scene(os.path.join(fixture_dir("cbox"), "cbox.ljs"))
    w, h = sc.width, sc.height
    sums = torch.zeros((h, w, 3), dtype=torch.float64, device="cuda:0")
    rp = RenderParams(spp=4, rng_mode=RNG_SAMPLE_STREAMS)
    sc.render_into_device(sums.data_ptr(), rp)
    img, _ = sc.render(rp)
    fin = gpu.accum_finalize(sums.data_ptr(), w, h, 4)
    assert np.array_equal(fin, img)
    assert gpu.accum_encode(sums.data_ptr(), w, h, 4, IMAGE_PFM) == gpu.encode_image(img, IMAGE_PFM)
    assert gpu.accum_encode(sums.data_ptr(), w, h, 4, IMAGE_EXR) == gpu.encode_image(img, IMAGE_EXR)
    out = tmp_path / "cbox.exr"
    gpu.accum_write(str(out), sums.data_ptr(), w, h, 4)
    assert out.read_bytes() == gpu.encode_image(img, IMAGE_EXR)

    # two passes with a checkpoint in between
    first = torch.zeros_like(sums)
    sc.render_into_device(first.data_ptr(), RenderParams(spp=4, rng_mode=RNG_SAMPLE_STREAMS, sample_begin=0, sample_end=2))
    ck = tmp_path / "cbox.ljacc"
    gpu.accum_save(str(ck), first.data_ptr(), w, h, 0, 2)
    restored = torch.full_like(sums, -1.0)
    assert gpu.accum_load(str(ck), restored.data_ptr(), w, h) == (0, 2)
    assert torch.equal(restored, first)
    second = torch.zeros_like(sums)
    sc.render_into_device(second.data_ptr(), RenderParams(spp=4, rng_mode=RNG_SAMPLE_STREAMS, sample_begin=2, sample_end=4))
    total = restored + second
    assert torch.allclose(total, sums, rtol=1e-12, atol=1e-300)
    with pytest.raises(Exception):
        gpu.accum_load(str(ck), restored.data_ptr(), w + 1, h)
    sc.close()
