"""N>1 host logic on CPU: world_size-2 (and 3) process groups over gloo.  The partition planner and
the reduction of lajolla_public_b200/distributed.py are the code under test; the per-rank "render"
is the oracle (reference code, oracle/_ref/ref_driver) rendering that rank's share, so the reduced
image can be compared with the oracle's single-process image of the same streams."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT
from test_cpu_oracle import needs_oracle, oracle_render

RES = (48, 40)     # 3 x 3 tiles, ragged right/bottom edges
SPP = 5            # not divisible by 2: slices of 3 and 2 samples


def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_plan_partition_covers_everything_once():
    from lajolla_public_b200.distributed import plan_partition, tile_owner_mask
    for world in (1, 2, 3, 4, 8):
        for spp in (1, 5, 64):
            parts = [plan_partition("spp", r, world, spp) for r in range(world)]
            covered = sorted(s for p in parts for s in range(p.sample_begin, p.sample_end))
            assert covered == list(range(spp))
            assert max(p.local_spp for p in parts) - min(p.local_spp for p in parts) <= 1
        masks = [tile_owner_mask(RES[0], RES[1], r, world) for r in range(world)]
        assert np.array_equal(np.sum(masks, axis=0), np.ones((RES[1], RES[0]), dtype=int))
        t = plan_partition("tiles", world - 1, world, 4)
        assert (t.sample_begin, t.sample_end) == (0, 4) and t.tile_rank == world - 1
    with pytest.raises(ValueError):
        plan_partition("spp", 2, 2, 4)
    with pytest.raises(ValueError):
        plan_partition("rows", 0, 2, 4)


def _worker(rank, world, port, kind, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    from lajolla_public_b200.distributed import render_distributed, tile_owner_mask
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def render_fn(part, sums):
        # this rank's share, rendered by the oracle on per-(pixel,sample) streams
        img, st = oracle_render("cbox/cbox.xml", RES, "B", part.spp, threads=2,
                                extra=("--sample-range", str(part.sample_begin), str(part.sample_end)))
        s = img.astype(np.float64) * part.local_spp          # back to radiance sums
        if part.tile_world > 0:
            s *= tile_owner_mask(RES[0], RES[1], part.tile_rank, part.tile_world)[:, :, None]
        sums.copy_(torch.from_numpy(s.reshape(-1)))
        return st

    image, _ = render_distributed(render_fn, RES[0], RES[1], SPP, kind=kind, root=None)
    np.save(os.path.join(out_dir, f"{kind}_{rank}.npy"), image.numpy())
    dist.barrier()
    dist.destroy_process_group()


@needs_oracle
@pytest.mark.parametrize("kind,world", [("spp", 2), ("tiles", 2), ("spp", 3)])
def test_two_rank_render_matches_single_process(tmp_path, kind, world):
    import torch.multiprocessing as mp
    port = free_port()
    mp.spawn(_worker, args=(world, port, kind, str(tmp_path)), nprocs=world, join=True)
    full, _ = oracle_render("cbox/cbox.xml", RES, "B", SPP)
    imgs = [np.load(tmp_path / f"{kind}_{r}.npy") for r in range(world)]
    for im in imgs[1:]:
        assert np.array_equal(im, imgs[0])                   # all-reduce: every rank holds the image
    if kind == "tiles":
        # disjoint pixels: adding zeros is exact; only the PFM f32 rounding of the oracle remains
        np.testing.assert_allclose(imgs[0], full, rtol=3e-7, atol=0)
    else:
        np.testing.assert_allclose(imgs[0], full, rtol=3e-6, atol=1e-7)
