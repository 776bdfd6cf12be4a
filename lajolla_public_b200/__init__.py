"""lajolla_public_b200 -- B200-native render hot path for lajolla (UCSD CSE 272).

The product is `liblajolla_b200.so` (hand-written sm_100a CUDA kernels behind the
C ABI of include/lajolla_gpu.h).  This package is the thin Python binding used by
the tests, bench.py and __graft_entry__.py; it adds no compute of its own and has
NO CPU fallback: importing works anywhere, but every render call needs a CUDA
device and the built shared library.
"""
from .api import (LajollaGPU, Scene, RenderParams, Stats, LjError, lib_path, load_library,
                  read_pfm, write_pfm, RNG_TILE_STREAMS, RNG_SAMPLE_STREAMS,
                  ENGINE_WAVEFRONT, ENGINE_MEGAKERNEL, PARTITION_AUTO, PARTITION_TILES, PARTITION_SPP)

__all__ = ["LajollaGPU", "Scene", "RenderParams", "Stats", "LjError", "lib_path", "load_library",
           "read_pfm", "write_pfm", "RNG_TILE_STREAMS", "RNG_SAMPLE_STREAMS",
           "ENGINE_WAVEFRONT", "ENGINE_MEGAKERNEL", "PARTITION_AUTO", "PARTITION_TILES", "PARTITION_SPP"]
