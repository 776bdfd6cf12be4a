// The two persistent trace kernels of the wavefront path tracer (stages "extend" and "shadow" of lj_wavefront.cu).
#include "lj_internal.h"
#include "lj_wf_pool.cuh"

// resident blocks per SM the trace kernels are compiled for (register budget = 65536 / (128 * this)).  Measured on
// sponza (extend stage): 5 -> 111 ms, 6 -> 105, 8 -> 95.3, 9 -> 93.2, 10 (48 registers, 60-84 B of spills) -> 91.5, 11 / 12 -> 100.5.
#ifndef LJ_PRIMS_PER_PHASE
#define LJ_PRIMS_PER_PHASE 2
#endif
#ifndef LJ_PRIMS_PER_PHASE_SHADOW
#define LJ_PRIMS_PER_PHASE_SHADOW 1
#endif
#ifndef LJ_NODES_PER_PHASE
#define LJ_NODES_PER_PHASE 1
#endif
#ifndef LJ_REFILL_TOP
#define LJ_REFILL_TOP 0
#endif
#ifndef LJ_TRACE_MIN_BLOCKS
#define LJ_TRACE_MIN_BLOCKS 10
#endif

namespace {

// ---- persistent trace kernels ----
// Rays differ a lot in traversal length, so a warp that just walks its 32 rays to the end runs
// mostly empty.  Instead every warp pulls rays from the queue through one atomic work counter
// and, whenever fewer than REFILL_BELOW lanes are still traversing, hands the idle lanes new
// rays (warp-aggregated fetch) before stepping on.

// ---- stage: extend = rtcIntersect1 (intersection.cpp:32) + classification by material kind ----
template <bool COUNT>
__global__ void __launch_bounds__(128, LJ_TRACE_MIN_BLOCKS) k_wf_extend(DScene sc, WfPool wp, int n_upper, unsigned long long *trav_counters, TraceTuning tune) {
    const int n = min(n_upper, wp.counts[Q_EXT]);
    if (blockIdx.x == 0 && threadIdx.x == 0) wp.counts[Q_REGEN] = 0;    // drained by the generate stage (see k_wf_generate)
    int ray_idx = -1; bool exhausted = false;
    Trav8 tv;
    tv.ng = make_uint2(0u, 0u); tv.tg = make_uint2(0u, 0u); tv.sp = 0; tv.top = make_uint2(0u, 0u);
    unsigned nn = 0, np = 0;
    while (true) {
        // retire finished lanes together (the hit reconstruction runs with as many lanes as possible)
        if (ray_idx >= 0 && !tv.active()) {
            const int slot = wp.ext_slot[ray_idx];
            Hit h; HitRec hr; int q;
            if (tv.finish(sc, &h)) {
                hr.t = h.t; hr.u = h.u; hr.v = h.v; hr.geom = h.geom; hr.prim = h.prim; hr.ngx = h.ngx; hr.ngy = h.ngy; hr.ngz = h.ngz;
                q = Q_KIND0 + (wp.single_kind >= 0 ? wp.single_kind : sc.materials[sc.shapes[h.geom].material_id].kind);
            } else {
                hr.t = 0; hr.u = hr.v = 0; hr.geom = -1; hr.prim = -1; hr.ngx = hr.ngy = hr.ngz = 0;
                q = Q_MISS;
            }
            wp.hit[slot] = hr;
            int *queue = (q == Q_MISS) ? wp.q_miss : wp.q_kind[q - Q_KIND0];
            queue[queue_push_pos(&wp.counts[q])] = slot;
            if (COUNT) { nn += tv.n_nodes; np += tv.n_prims; }
            ray_idx = -1;
        }
        if (warp_fetch<false>(&wp.counts[W_EXT], n, ray_idx, exhausted)) tv.init(wp.ext_ray[ray_idx].ray);
        if (!__ballot_sync(0xffffffffu, ray_idx >= 0)) break;
        const int stop_below = exhausted ? 1 : tune.refill_below;
        while (true) {
            const bool live = ray_idx >= 0;
#if LJ_REFILL_TOP
            if (live) tv.refill_from_stack();
#endif
            const bool wn = live && tv.wants_node(), wpr = live && tv.wants_prim();
            const unsigned mn = __ballot_sync(0xffffffffu, wn), mp = __ballot_sync(0xffffffffu, wpr);
            if (__popc(mn | mp) < stop_below) break;          // too few lanes still traversing: retire + refill first
            if (mp != 0u && (__popc(mp) >= tune.prim_thresh || mn == 0u)) {
                bool found = false;
                if (wpr) {
                    tv.prim_step<false, COUNT, !LJ_REFILL_TOP>(sc, found);
#if LJ_PRIMS_PER_PHASE > 1
#pragma unroll 1
                    for (int k = 1; k < LJ_PRIMS_PER_PHASE && tv.wants_prim(); k++) tv.prim_step<false, COUNT, !LJ_REFILL_TOP>(sc, found);
#endif
                }
            } else {
                if (wn) {
                    tv.node_step<false, COUNT, !LJ_REFILL_TOP>(sc);
#if LJ_NODES_PER_PHASE > 1
#pragma unroll 1
                    for (int k = 1; k < LJ_NODES_PER_PHASE && tv.wants_node(); k++) tv.node_step<false, COUNT, !LJ_REFILL_TOP>(sc);
#endif
                }
            }
        }
    }
    if (COUNT) { atomicAdd(&trav_counters[0], (unsigned long long)nn); atomicAdd(&trav_counters[1], (unsigned long long)np); }
}

// ---- stage: shadow = rtcOccluded1 (intersection.cpp:83) + accumulate the NEE term ----
template <bool COUNT>
__global__ void __launch_bounds__(128, LJ_TRACE_MIN_BLOCKS) k_wf_shadow(DScene sc, WfPool wp, int n_upper, unsigned long long *trav_counters, TraceTuning tune) {
    const int n = min(n_upper, wp.counts[Q_SHADOW]);
    int ray_idx = -1; bool exhausted = false;
    Trav8 tv;
    tv.ng = make_uint2(0u, 0u); tv.tg = make_uint2(0u, 0u); tv.sp = 0; tv.top = make_uint2(0u, 0u);
    bool found = false;
    unsigned nn = 0, np = 0;
    while (true) {
        if (ray_idx >= 0 && (found || !tv.active())) {
            if (!found) {
                const ShadowRec *s = &wp.shadow[ray_idx];
                RadRec *ra = &wp.rad[s->slot];
                ra->rx += s->cx; ra->ry += s->cy; ra->rz += s->cz;        // path_tracing.h:207
            }
            if (COUNT) { nn += tv.n_nodes; np += tv.n_prims; }
            ray_idx = -1;
        }
        if (warp_fetch<true>(&wp.counts[W_SHADOW], n, ray_idx, exhausted)) { tv.init(wp.shadow[ray_idx].ray); found = false; }
        if (!__ballot_sync(0xffffffffu, ray_idx >= 0)) break;
        const int stop_below = exhausted ? 1 : tune.refill_below;
        while (true) {
            const bool live = ray_idx >= 0 && !found;
#if LJ_REFILL_TOP
            if (live) tv.refill_from_stack();
#endif
            const bool wn = live && tv.wants_node(), wpr = live && tv.wants_prim();
            const unsigned mn = __ballot_sync(0xffffffffu, wn), mp = __ballot_sync(0xffffffffu, wpr);
            if (__popc(mn | mp) < stop_below) break;
            if (mp != 0u && (__popc(mp) >= tune.prim_thresh || mn == 0u)) {
                if (wpr) {
                    tv.prim_step<true, COUNT, !LJ_REFILL_TOP>(sc, found);
#if LJ_PRIMS_PER_PHASE_SHADOW > 1
#pragma unroll 1
                    for (int k = 1; k < LJ_PRIMS_PER_PHASE_SHADOW && !found && tv.wants_prim(); k++) tv.prim_step<true, COUNT, !LJ_REFILL_TOP>(sc, found);
#endif
                }
            } else {
                if (wn) { if (tv.node_step<true, COUNT, !LJ_REFILL_TOP>(sc)) found = true; }
            }
        }
    }
    if (COUNT) { atomicAdd(&trav_counters[2], (unsigned long long)nn); atomicAdd(&trav_counters[3], (unsigned long long)np); }
}

}  // namespace

void lj_wf_launch_extend(const DScene &sc, const WfPool &wp, int n_upper, unsigned long long *trav_counters, TraceTuning tune, bool counting, unsigned grid) {
    if (counting) k_wf_extend<true><<<grid, 128>>>(sc, wp, n_upper, trav_counters, tune);
    else k_wf_extend<false><<<grid, 128>>>(sc, wp, n_upper, trav_counters, tune);
}
void lj_wf_launch_shadow(const DScene &sc, const WfPool &wp, int n_upper, unsigned long long *trav_counters, TraceTuning tune, bool counting, unsigned grid, cudaStream_t stream) {
    if (counting) k_wf_shadow<true><<<grid, 128, 0, stream>>>(sc, wp, n_upper, trav_counters, tune);
    else k_wf_shadow<false><<<grid, 128, 0, stream>>>(sc, wp, n_upper, trav_counters, tune);
}
