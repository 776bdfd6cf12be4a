"""Host-side geometry of the work partition and the accumulators (lajolla_public_b200/csrc/lj_wave_common.cuh SlotMap,
lj_scene.cuh tile_dealt_index): compiled with nvcc as plain host code and run here, no GPU.  Checked:
  * the tile deal is a bijection, and for every film / world size each pixel belongs to exactly one rank's pool;
  * K0 (accumulators per pixel) depends on the film and the sample count only -- never on the partition or the pool
    size -- and K divides K0: the property that makes N-way renders bit-identical to the one-device render;
  * every (pixel, accumulator) pair of a rank maps to its own cell, and a slot's samples visit accumulators in a fixed order;
  * distributed.tile_owner_mask (Python) is the same deal."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NVCC = "/usr/local/cuda/bin/nvcc"

SRC = r'''
#include <cstdio>
#include <vector>
#include "lj_wave_common.cuh"
int main() {
    const int films[][2] = {{768, 575}, {512, 512}, {683, 512}, {1366, 1024}, {17, 5}, {16, 16}, {48, 33}};
    const int worlds[] = {0, 1, 2, 3, 5, 8};
    long long checked = 0;
    for (auto &f : films) {
        const int w = f[0], h = f[1], ntx = (w + 15) / 16, nty = (h + 15) / 16;
        // deal: bijection on the tile grid
        std::vector<int> seen(ntx * nty, 0);
        for (int ty = 0; ty < nty; ty++) for (int tx = 0; tx < ntx; tx++) {
            const int v = tile_dealt_index(tx, ty, ntx);
            if (v < 0 || v >= ntx * nty || seen[v]++) { printf("FAIL deal not a bijection %dx%d\n", w, h); return 1; }
            int bx, by; tile_from_dealt_index(v, ntx, bx, by);
            if (bx != tx || by != ty) { printf("FAIL deal inverse %dx%d\n", w, h); return 1; }
        }
        for (int nsamp : {1, 64, 150, 1024}) {
            int k0_ref = -1;
            for (int world : worlds) {
                std::vector<int> owner((size_t)w * h, -1);
                for (int rank = 0; rank < (world > 0 ? world : 1); rank++) {
                    for (long long target : {1LL << 16, 12LL << 20, 64LL << 20}) {
                        SlotMap m = lj_make_slot_map(w, h, 3, 3 + nsamp, rank, world, target);
                        if (k0_ref < 0) k0_ref = m.K0;
                        if (m.K0 != k0_ref) { printf("FAIL K0 depends on the partition / pool: %d vs %d\n", m.K0, k0_ref); return 1; }
                        if (m.K < 1 || m.K0 % m.K != 0 || (m.K & (m.K - 1)) || (m.K0 & (m.K0 - 1))) { printf("FAIL K %d K0 %d\n", m.K, m.K0); return 1; }
                        if (target != (12LL << 20)) continue;
                        std::vector<bool> cells((size_t)m.K0 * m.npix_local, false);
                        for (int pl = 0; pl < m.npix_local; pl++) {
                            int pix, x, y;
                            if (!slot_pixel(m, pl, pix, x, y)) continue;
                            if (pix != y * w + x || x < 0 || x >= w || y < 0 || y >= h) { printf("FAIL pixel mapping\n"); return 1; }
                            if (owner[pix] != -1) { printf("FAIL pixel %d owned twice (%dx%d world %d)\n", pix, w, h, world); return 1; }
                            owner[pix] = rank;
                            if (slot_stream(m, pl + m.npix_local * (m.K - 1), 7) != 7ULL * (unsigned long long)((long long)w * h) + (unsigned long long)pix) { printf("FAIL stream id\n"); return 1; }
                            // sample s of a slot lane k goes to accumulator (s - s_begin) mod K0; lanes k, k+K, ... of one slot
                            for (int j = 0; j < m.K0 && j < 8; j++) {
                                const int slot = (j % m.K) * m.npix_local + pl;
                                const size_t off = accum_offset(m, slot, 3 + j);
                                if (off % 3 || off / 3 >= cells.size() || cells[off / 3]) { printf("FAIL accumulator cell reused\n"); return 1; }
                                cells[off / 3] = true;
                                if (off != accum_offset(m, slot, 3 + j + m.K0)) { printf("FAIL accumulator not periodic in K0\n"); return 1; }
                            }
                            checked++;
                        }
                    }
                }
                for (size_t i = 0; i < owner.size(); i++) if (owner[i] < 0) { printf("FAIL pixel %zu unowned (%dx%d world %d)\n", i, w, h, world); return 1; }
            }
        }
    }
    printf("OK %lld\n", checked);
    return 0;
}
'''


@pytest.fixture(scope="module")
def slotmap_binary(tmp_path_factory):
    if not os.path.exists(NVCC):
        pytest.skip("nvcc not present")
    d = tmp_path_factory.mktemp("slotmap")
    src = d / "slotmap_host.cu"
    src.write_text(SRC)
    exe = d / "slotmap_host"
    r = subprocess.run([NVCC, "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a", "-I", os.path.join(ROOT, "lajolla_public_b200", "csrc"), "-o", str(exe), str(src)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout[-3000:]
    return str(exe)


def test_slot_map_and_tile_deal(slotmap_binary):
    r = subprocess.run([slotmap_binary], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    print(r.stdout[-500:])
    assert r.returncode == 0 and r.stdout.strip().startswith("OK"), r.stdout[-2000:]


def test_python_mask_restates_the_library_deal():
    from lajolla_public_b200.distributed import tile_owner_mask
    for (w, h) in ((768, 575), (48, 33), (17, 5)):
        ntx = (w + 15) // 16
        for world in (2, 3, 8):
            total = np.zeros((h, w), dtype=np.int32)
            for r in range(world):
                m = tile_owner_mask(w, h, r, world)
                total += m
                ys, xs = np.nonzero(m)
                ty, tx = ys // 16, xs // 16
                assert np.all((ty * ntx + (tx + 3 * ty) % ntx) % world == r)
            assert np.all(total == 1)
