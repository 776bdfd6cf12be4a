"""lj_accum_finalize_static (allocation-free variant of lj_accum_finalize, include/lajolla_gpu.h): its f32 staging image
is a file-scope static __device__ array instead of an allocation.  Checked through the C ABI against the existing
entry point, bit for bit (the tolerance of the other output tests), on two calls in a row with DIFFERENT inputs (the
scratch keeps the previous call's contents) and a different image size, plus the refusal beyond the stated maximum.
Collected last on purpose."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def test_static_finalize_matches_allocating_finalize(gpu):
    import torch
    from lajolla_public_b200 import RenderParams, RNG_SAMPLE_STREAMS
    sc = gpu.load_scene(os.path.join(GOLDEN, "cbox.ljs"), res=(160, 96))
    try:
        w, h = sc.width, sc.height
        sums = torch.zeros((h, w, 3), dtype=torch.float64, device="cuda:0")
        sc.render_into_device(sums.data_ptr(), RenderParams(spp=3, rng_mode=RNG_SAMPLE_STREAMS))
        torch.cuda.synchronize()
        ref1 = gpu.accum_finalize(sums.data_ptr(), w, h, 3)
        got1 = gpu.accum_finalize_static(sums.data_ptr(), w, h, 3)
        assert np.array_equal(got1.view(np.uint32), ref1.view(np.uint32))
        assert float(np.abs(got1).max()) > 0
        # second call in a row: other values, other spp; nothing of the first call may leak through the shared scratch
        sums2 = sums * 0.37 + 1.25
        torch.cuda.synchronize()
        ref2 = gpu.accum_finalize(sums2.data_ptr(), w, h, 7)
        got2 = gpu.accum_finalize_static(sums2.data_ptr(), w, h, 7)
        assert np.array_equal(got2.view(np.uint32), ref2.view(np.uint32))
        assert not np.array_equal(got2, got1)
        # a smaller image after a larger one: only its own w*h*3 values are produced
        small = sums2[:40, :64, :].contiguous()
        torch.cuda.synchronize()
        ref3 = gpu.accum_finalize(small.data_ptr(), 64, 40, 2)
        got3 = gpu.accum_finalize_static(small.data_ptr(), 64, 40, 2)
        assert np.array_equal(got3.view(np.uint32), ref3.view(np.uint32))
    finally:
        sc.close()


def test_static_finalize_refuses_images_beyond_its_scratch(gpu):
    import torch
    from lajolla_public_b200 import LjError
    dummy = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    out = np.empty((1, 1, 3), dtype=np.float32)
    with pytest.raises(LjError, match="exceed the static scratch"):
        gpu.accum_finalize_static(dummy.data_ptr(), 2048, 1024, 1, out=out)      # 2^21 pixels > 2^20
