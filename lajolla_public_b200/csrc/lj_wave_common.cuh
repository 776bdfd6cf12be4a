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

