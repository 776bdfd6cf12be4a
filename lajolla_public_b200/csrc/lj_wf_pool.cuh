// Path-slot pool of the wavefront path tracer (lj_wavefront.cu) and the launchers of its two persistent trace
// kernels, which live in their own translation unit (lj_wf_trace.cu) so that traversal experiments do not
// recompile the nine shade kernels.
#pragma once
#include "lj_wave_common.cuh"
#include "../../include/lajolla_gpu.h"

enum { Q_EXT = 0, Q_SHADOW = 1, Q_REGEN = 2, Q_FINISH = 3, Q_MISS = 4, Q_KIND0 = 5, Q_COUNT = Q_KIND0 + LJ_MAT_KIND_COUNT };

struct __align__(16) RayRec { double ox, oy, oz, spread, dx, dy, dz, eta_scale; };          // 64 B
struct __align__(16) ThrRec { double tx, ty, tz, p2, fx, fy, fz; int nv, sample; };          // 64 B
struct __align__(16) RadRec { double rx, ry, rz; unsigned long long rng; };                  // 32 B
struct __align__(16) ShadowRec { RayF ray; double cx, cy, cz; int slot, _pad; };             // 64 B

struct WfPool {
    int capacity;               // = map.K * map.npix_local slots
    SlotMap map;                // slot -> pixel / sample lane / accumulator (lj_wave_common.cuh)
    RayRec *ray; ThrRec *thr; RadRec *rad; HitRec *hit;
    double *accum;              // 3 * K0 per local pixel, see SlotMap
    ExtRec *ext_ray; int *ext_slot;
    ShadowRec *shadow;
    int *q_regen, *q_miss, *q_finish, *q_kind[LJ_MAT_KIND_COUNT];
    int *counts;                // Q_COUNT live queue sizes
    int single_kind;            // the scene's only material kind, or -1: the extend stage then classifies a hit without
                                // chasing shapes[geom] -> materials[id].kind (two dependent loads with few lanes active)
};

enum { W_EXT = 24, W_SHADOW = 25 };     // work counters of the trace kernels inside wp.counts

// extend = rtcIntersect1 + classification by material kind; shadow = rtcOccluded1 + NEE accumulation
void lj_wf_launch_extend(const DScene &sc, const WfPool &wp, int n_upper, unsigned long long *trav_counters, TraceTuning tune, bool counting, unsigned grid);
void lj_wf_launch_shadow(const DScene &sc, const WfPool &wp, int n_upper, unsigned long long *trav_counters, TraceTuning tune, bool counting, unsigned grid, cudaStream_t stream);
