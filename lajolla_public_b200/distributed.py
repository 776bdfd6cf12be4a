"""Multi-GPU rendering: one process per GPU (torchrun), scene replicated, work partitioned, one
reduction of the f64 radiance sums over NCCL (NVLink / NVSwitch).  SURVEY.md 8(e).

The reference's unit of independent work is the 16x16 tile with its own pcg32 stream
(render.cpp:80-98); per-(pixel,sample) streams make every sample independent.  Two partitions:

  "spp"   rank r renders samples [begin_r, end_r) of EVERY pixel (balanced even when cost is
          spatially skewed); the per-pixel sums of all ranks are added.
  "tiles" tile (tx, ty) belongs to rank (ty*ntx + (tx + 3 ty) % ntx) % world (the library's deal: diagonals rather
          than vertical stripes); pixels are disjoint, so the reduced image is bit-identical to the 1-GPU image for any
          world size (adding zeros is exact).

There is no data-path exchange other than that single reduction.  The render itself is the
C-ABI call lj_render_device (include/lajolla_gpu.h); this module only plans the partition and
issues the collective.  `render_fn` is injectable so the planning + reduction logic is testable
on CPU with the gloo backend (tests/test_cpu_distributed.py).
"""
from dataclasses import dataclass

TILE = 16   # render.cpp:76


@dataclass
class Partition:
    kind: str            # "spp" | "tiles"
    rank: int
    world: int
    spp: int             # total samples per pixel of the final image
    sample_begin: int
    sample_end: int
    tile_rank: int
    tile_world: int      # 0 => this rank renders every tile

    @property
    def local_spp(self):
        return self.sample_end - self.sample_begin


def plan_partition(kind, rank, world, spp):
    """Split `spp` samples per pixel (kind "spp") or the tile grid (kind "tiles") over `world` ranks."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    if spp < 1:
        raise ValueError("spp must be positive")
    if kind == "spp":
        # contiguous slices whose sizes differ by at most one sample; ranks beyond spp get nothing
        base, extra = divmod(spp, world)
        begin = rank * base + min(rank, extra)
        end = begin + base + (1 if rank < extra else 0)
        return Partition(kind, rank, world, spp, begin, end, 0, 0)
    if kind == "tiles":
        return Partition(kind, rank, world, spp, 0, spp, rank, world if world > 1 else 0)
    raise ValueError(f"unknown partition kind {kind!r}")


def tile_owner_mask(width, height, rank, world):
    """boolean (h, w) mask of the pixels whose tile id (ty*ntx+tx, render.cpp:82) is dealt to `rank`"""
    import numpy as np
    ntx = (width + TILE - 1) // TILE
    ys, xs = np.mgrid[0:height, 0:width]
    # the library's deal (tile_dealt_index, csrc/lj_scene.cuh): each tile row rotated by three more columns than the
    # one above, so a rank's tiles form diagonals rather than vertical stripes
    ty, tx = ys // TILE, xs // TILE
    dealt = ty * ntx + (tx + 3 * ty) % ntx
    return (dealt % world) == rank if world > 1 else np.ones((height, width), dtype=bool)


def params_for(part, rng_mode, engine, flags=0):
    from .api import RenderParams
    return RenderParams(spp=part.spp, rng_mode=rng_mode, engine=engine, sample_begin=part.sample_begin,
                        sample_end=part.sample_end, tile_rank=part.tile_rank, tile_world=part.tile_world, flags=flags)


def reduce_sums(sums, world, root=None):
    """sum the per-rank radiance sums: all-reduce (root None) or reduce to `root`"""
    if world == 1:
        return sums
    import torch.distributed as dist
    if root is None:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    else:
        dist.reduce(sums, dst=root, op=dist.ReduceOp.SUM)
    return sums


def render_distributed(render_fn, width, height, spp, kind="spp", rank=None, world=None, root=None, device=None):
    """Render one image over all ranks of the default process group.

    render_fn(partition, sums) must ADD NOTHING ELSE than this rank's share: it fills `sums`
    (torch f64 tensor of h*w*3, on `device`) with per-pixel radiance sums over the partition's
    samples/tiles (zeros elsewhere) and returns its stats object (or None).
    Returns (image f32 tensor (h, w, 3) valid on `root` or on every rank, stats)."""
    import torch
    import torch.distributed as dist
    if world is None:
        world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if world > 1 else 0
    part = plan_partition(kind, rank, world, spp)
    sums = torch.zeros(height * width * 3, dtype=torch.float64, device=device)
    stats = None
    if part.local_spp > 0:
        stats = render_fn(part, sums)
    reduce_sums(sums, world, root)
    # render.cpp:94 divides by multiplying with the reciprocal (vector.h:193-197), and so does k_finalize
    image = (sums * (1.0 / float(spp))).to(torch.float32).reshape(height, width, 3)
    return image, stats


def gpu_render_fn(scene, rng_mode, engine, flags=0):
    """render_fn for a GPU-resident lajolla_public_b200.Scene (the product path)"""
    def fn(part, sums):
        return scene.render_into_device(sums.data_ptr(), params_for(part, rng_mode, engine, flags))
    return fn
