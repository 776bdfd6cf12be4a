// Pieces shared by the wavefront engines (lj_wavefront.cu: path integrator, lj_volwave.cu: volumetric
// path integrator): queue records that cross the trace stage, warp-aggregated queue pushes, the
// work-fetch of the persistent trace kernels and their tuning knobs.
#pragma once
#include "lj_traverse8.cuh"

struct __align__(16) HitRec { float t, u, v; int geom, prim; float ngx, ngy, ngz; };         // 32 B
struct __align__(16) ExtRec { RayF ray; };                                                   // 32 B

// Warp-aggregated queue push: lanes of the calling (possibly partial) warp that push to the
// same queue elect one leader, which reserves a contiguous range with a single atomicAdd.
LJ_D int queue_push_pos(int *counter) {
    unsigned mask = __activemask();
    unsigned peers = __match_any_sync(mask, (unsigned long long)counter);
    int leader = __ffs(peers) - 1;
    int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(peers));
    base = __shfl_sync(peers, base, leader);
    return base + __popc(peers & ((1u << lane) - 1));
}

// Same, for call sites where every lane pushes to the SAME queue: the converged lanes are the
// group (no match_any, and the shuffle mask is what the hardware already has).
LJ_D int queue_push_uniform(int *counter) {
    const unsigned mask = __activemask();
    const int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(mask, base, leader);
    return base + __popc(mask & ((1u << lane) - 1));
}

template <bool ANY>
LJ_D bool warp_fetch(int *work_counter, int n, int &ray_idx, bool &exhausted) {
    // called by all 32 lanes; lanes with ray_idx < 0 ask for work
    const unsigned idle = __ballot_sync(0xffffffffu, ray_idx < 0);
    if (!idle || exhausted) return false;
    const int lane = threadIdx.x & 31;
    const int n_idle = __popc(idle);
    const int leader = __ffs(idle) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(work_counter, n_idle);
    base = __shfl_sync(0xffffffffu, base, leader);
    if (base + n_idle >= n) exhausted = true;
    if (ray_idx < 0) {
        int i = base + __popc(idle & ((1u << lane) - 1u));
        if (i < n) { ray_idx = i; return true; }
    }
    return false;
}

// Phase vote of a trace warp (see lj_traverse8.cuh): lanes either want to open a node or have a
// primitive pending.  Primitive tests are postponed until at least `prim_thresh` lanes have one
// (or no lane has node work left), so both phases run with well-filled warps.
struct TraceTuning { int refill_below, prim_thresh; };


// ---- slot -> pixel mapping and radiance accumulators, shared by both wavefront engines ----
// A pool holds K sample lanes of `npix_local` pixels: slot = k * npix_local + pl renders samples
// s_begin + k, s_begin + k + K, ... of local pixel pl.  Without a tile partition pl is the pixel index; with one
// (tile t belongs to the rank t % world, render.cpp:75-82's 16x16 tiles) the OWNED tiles are numbered densely,
// pl = local_tile * 256 + (y % 16) * 16 + (x % 16), so that a rank's pool holds only its own pixels (slots of the
// few pixels of border tiles that fall outside the film stay empty).
// Accumulation is independent of K and of the partition: sample s of a pixel is added to accumulator
// j = (s - s_begin) mod K0 of that pixel, K0 a power of two that depends only on the film and the sample count,
// and K is a power of two <= K0, so accumulator j receives samples j, j + K0, j + 2 K0, ... all from the one slot
// lane j mod K, in increasing order; the final per-pixel sum adds the K0 accumulators in order.  A render split
// by tiles over N devices therefore produces bit-identical sums to the unpartitioned render.
struct SlotMap {
    int npix_local;             // pixels (slots per sample lane) of this pool
    long long npix_total;       // width * height: the pcg32 stream id is s * npix_total + pixel
    int width, height, ntx;
    int tile_rank, tile_world;  // tile_world 0: no partition (identity mapping)
    int K, K0, s_begin;
};

LJ_HD bool slot_pixel(const SlotMap &m, int slot, int &pix, int &x, int &y) {
    const int pl = slot % m.npix_local;
    if (m.tile_world <= 0) { pix = pl; x = pl % m.width; y = pl / m.width; return true; }
    int tx, ty;
    tile_from_dealt_index((pl >> 8) * m.tile_world + m.tile_rank, m.ntx, tx, ty);     // lj_scene.cuh
    x = tx * 16 + (pl & 15);
    y = ty * 16 + ((pl >> 4) & 15);
    pix = y * m.width + x;
    return x < m.width && y < m.height;
}
LJ_HD int slot_pixel_index(const SlotMap &m, int slot) { int pix, x, y; slot_pixel(m, slot, pix, x, y); return pix; }
LJ_HD unsigned long long slot_stream(const SlotMap &m, int slot, int sample) {
    return (unsigned long long)sample * (unsigned long long)m.npix_total + (unsigned long long)slot_pixel_index(m, slot);
}
// the three doubles sample `sample` of the slot's pixel is accumulated into
LJ_HD size_t accum_offset(const SlotMap &m, int slot, int sample) {
    const int j = (sample - m.s_begin) & (m.K0 - 1);
    return 3 * ((size_t)j * m.npix_local + (size_t)(slot % m.npix_local));
}

// per-pixel sums: adds the K0 accumulators of every owned pixel in order (pixels of other ranks stay as they are: the
// caller zeroes `sum` first when a partition is active)
static __global__ void k_slot_sum(SlotMap m, const double *accum, double *sum) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)m.npix_local * 3) return;
    const int pl = (int)(i / 3), c = (int)(i % 3);
    int pix, x, y;
    if (!slot_pixel(m, pl, pix, x, y)) return;
    double a = 0;
    for (int j = 0; j < m.K0; j++) a += accum[3 * ((size_t)j * m.npix_local + pl) + c];
    sum[3 * (size_t)pix + c] = a;
}

// Host side: pool geometry for a render (film, sample range, partition, slots wanted in flight).
static inline int lj_pow2_floor(long long v) { int p = 1; while ((long long)p * 2 <= v) p *= 2; return p; }
static inline int lj_pow2_ceil(long long v) { int p = 1; while (p < v) p *= 2; return p; }
static inline SlotMap lj_make_slot_map(int width, int height, int s_begin, int s_end, int tile_rank, int tile_world, long long target_slots) {
    SlotMap m;
    m.width = width; m.height = height; m.ntx = (width + 15) / 16;
    m.npix_total = (long long)width * height;
    m.tile_rank = tile_rank; m.tile_world = tile_world; m.s_begin = s_begin;
    if (tile_world > 0) {
        const int ntiles = m.ntx * ((height + 15) / 16);
        const int mine = ntiles > tile_rank ? (ntiles - tile_rank + tile_world - 1) / tile_world : 0;
        m.npix_local = (mine > 0 ? mine : 1) * 256;
    } else {
        m.npix_local = (int)m.npix_total;
    }
    const int nsamp = s_end - s_begin;
    // K0 from the film and the sample count only (never from the partition or the pool size): up to 256 Mi accumulator
    // cells (6 GB of f64 RGB) over the whole film, so that at high sample counts a tile share of the film can keep as many
    // paths in flight as a sample-slice share (sponza, 1024 spp on 8 GPUs: K0 = 512 instead of 64)
    long long by_mem = (256LL << 20) / (m.npix_total > 0 ? m.npix_total : 1);
    int k0 = lj_pow2_ceil(nsamp);
    if (k0 > 1024) k0 = 1024;
    if (k0 > lj_pow2_floor(by_mem < 1 ? 1 : by_mem)) k0 = lj_pow2_floor(by_mem < 1 ? 1 : by_mem);
    m.K0 = k0;
    long long want = target_slots / m.npix_local;
    int k = lj_pow2_floor(want < 1 ? 1 : want);
    if (k > k0) k = k0;
    if (k > lj_pow2_floor(nsamp < 1 ? 1 : nsamp)) k = lj_pow2_floor(nsamp < 1 ? 1 : nsamp);
    m.K = k;
    return m;
}
