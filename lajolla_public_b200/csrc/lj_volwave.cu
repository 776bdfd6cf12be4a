// Wavefront engine for the volumetric path integrator (vol_path_tracing.h:503-869, the integrator
// versions 3-6 forward to; mode B streams).  The reference's loop body is cut at its two ray casts:
//
//   generate -> extend (closest hit; main rays AND next-event segments: the reference's NEE walks a
//               chain of intersect() calls through index-matched surfaces, vol_path_tracing.h:343)
//            -> main   : free-flight sampling through the current medium (delta tracking with the
//                        null-collision loop kept inside the stage: it is RNG-serial), emitter MIS,
//                        index-matched pass-through, depth limit, light sample -> first NEE segment
//            -> nee    : ratio tracking along one NEE segment, blocked / pass-through / done
//            -> finish : NEE contribution + spectral MIS, phase-function or BSDF sampling (one kernel
//                        per material kind), Russian roulette -> next main ray
//
// Path slots are regenerated as in lj_wavefront.cu: slot (pixel, k) renders samples s_begin+k,
// s_begin+k+K, ... and owns its accumulator.  Same RNG draw order, early exits and operation order as
// lj_volpath.cuh (the one-lane-per-path form used by the parity engines), which this file reuses for
// media, phase functions and the tracking arithmetic.
#include "lj_internal.h"
#include "lj_volpath.cuh"
#include "lj_wave_common.cuh"
#include <cstdio>
#include <cstdlib>
#include <algorithm>

// resident blocks per SM the per-vertex and tracking kernels are compiled for (experiment knobs; 1 = no register cap).
// Coloured tracker: 3 (168 registers, 40 B of spills) -- 2 blocks (212 registers): hetvol_colored 24.8 Msamples/s, 3: 39.6, 4 (128 registers,
// 200 B of spills): 37.8.  Grey tracker: 5 (84-92 registers, no spills); 4 and 6 measure the same.
#ifndef LJ_VOL_TRACK_MIN_BLOCKS
#define LJ_VOL_TRACK_MIN_BLOCKS 3
#endif
#ifndef LJ_VOL_TRACK_GREY_MIN_BLOCKS
#define LJ_VOL_TRACK_GREY_MIN_BLOCKS 5
#endif
#ifndef LJ_VOL_MAIN_MIN_BLOCKS
#define LJ_VOL_MAIN_MIN_BLOCKS 1
#endif

namespace {

enum { VQ_EXT = 0, VQ_REGEN = 1, VQ_MAIN = 2, VQ_NEE = 3, VQ_FIN_MEDIUM = 4, VQ_FIN_KIND0 = 5,
       VQ_COUNT = VQ_FIN_KIND0 + LJ_MAT_KIND_COUNT,
       // grid media: segments whose tracking is complete (input of the main / nee stages), and the two
       // ping-pong queues of segments suspended at the step budget
       VT_MAIN_READY = VQ_COUNT, VT_NEE_READY, VT_CONT_MAIN0, VT_CONT_MAIN1, VT_CONT_NEE0, VT_CONT_NEE1, VT_END,
       VW_EXT = 24, VW_TRACK_MAIN = 25, VW_TRACK_NEE = 26 };
static_assert(VT_END <= VW_EXT, "queue counters overlap the work counters");
constexpr int NEE_RAY = (int)0x80000000;    // flag in ext_slot: the ray is an NEE segment

// Per-slot path state that survives between stages (f64, as the reference computes in double).
struct __align__(16) VolPath {
    double ox, oy, oz, spread;              // ray.org, ray_diff.spread
    double dx, dy, dz, eta_scale;           // ray.dir
    double tx, ty, tz, dir_pdf;             // current_path_throughput, pdf of the latest direction sampling
    double rx, ry, rz; unsigned long long rng;
    double mx, my, mz; int medium, bounces; // multi_trans_pdf
    double cx, cy, cz; int sample, first;   // nee_p_cache
};                                          // 192 B
// Next-event state while its segments are traced (vol_path_tracing.h:299-495 locals).
struct __align__(16) VolNee {
    double px, py, pz; int light_id, shadow_medium;     // segment origin p
    double lx, ly, lz; int shadow_bounces, kind;        // dir_light; kind: -1 = medium scatter, else material kind
    double qx, qy, qz; int blocked, pad0;               // p_prime (point on the light)
    double nx, ny, nz, pad1;                            // normal at p_prime
    double ax, ay, az, pad2;                            // transmittance_light
    double bx, by, bz, pad3;                            // p_trans_nee
    double ex, ey, ez, pad4;                            // p_trans_dir
    double gx, gy, gz, pad5;                            // ray.org at the time the main ray was cast (vertex reconstruction)
};                                                      // 256 B

// Result of the free-flight tracking of a main ray (written by k_vol_track<false>, read by k_vol_main).
struct __align__(16) VolTrack {
    double tx, ty, tz, ox;                  // transmittance, new ray.org.x
    double dx, dy, dz, oy;                  // trans_dir_pdf, new ray.org.y
    double nx, ny, nz, oz;                  // trans_nee_pdf, new ray.org.z
    double mjx, mjy, mjz, accum_t;          // suspended segment: majorant along the ray, distance covered
    int scatter, channel, iteration, pad0;  // suspended segment: sampled channel, collisions so far
};                                          // 144 B

struct VolPool {
    int capacity;               // = map.K * map.npix_local slots
    SlotMap map;                // slot -> pixel / sample lane / accumulator (lj_wave_common.cuh)
    VolPath *path; VolNee *nee; HitRec *hit, *nee_hit; VolTrack *track;
    double *accum;
    ExtRec *ext_ray; int *ext_slot;
    int *q_regen, *q_main, *q_nee, *q_fin_medium, *q_fin_kind[LJ_MAT_KIND_COUNT];
    int *q_main_ready, *q_nee_ready, *q_cont_main[2], *q_cont_nee[2];
    int *counts;
};

LJ_D void vol_terminate(const VolPool &vp, int slot, int sample, d3 L) {
    // render.cpp:140 with vector.h:289-292: a sample is dropped only if ALL channels are non-finite
    if (isfinite(L.x) || isfinite(L.y) || isfinite(L.z)) {
        double *a = vp.accum + accum_offset(vp.map, slot, sample);
        a[0] += L.x; a[1] += L.y; a[2] += L.z;
    }
    vp.q_regen[queue_push_uniform(&vp.counts[VQ_REGEN])] = slot;
}

LJ_D void vol_push_ray(const VolPool &vp, int slot_and_flag, const Ray &r) {
    const int pos = queue_push_uniform(&vp.counts[VQ_EXT]);
    ExtRec e; e.ray = to_rayf(r);
    vp.ext_ray[pos] = e;
    vp.ext_slot[pos] = slot_and_flag;
}

LJ_D Ray vol_main_ray(const DScene &sc, const VolPath &p) {
    return make_ray(mk3(p.ox, p.oy, p.oz), mk3(p.dx, p.dy, p.dz), sc.eps, INFINITY);   // tnear set once, :513
}

// ---- stage: generate (vol_path_tracing.h:509-532) ----
__global__ void __launch_bounds__(128) k_vol_generate(DScene sc, VolPool vp, int n_upper, int s_end) {
    const int n_regen = min(n_upper, vp.counts[VQ_REGEN]);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_regen; i += gridDim.x * blockDim.x) {
        const int slot = vp.q_regen[i];
        VolPath p = vp.path[slot];
        const int s = p.sample + vp.map.K;
        int pix, x, y;
        const bool on_film = slot_pixel(vp.map, slot, pix, x, y);
        if (s >= s_end || !on_film) continue;
        Pcg32 rng = pcg_init((unsigned long long)s * (unsigned long long)vp.map.npix_total + (unsigned long long)pix);
        Ray ray = volpath_primary(sc, x, y, rng);
        p.ox = ray.org.x; p.oy = ray.org.y; p.oz = ray.org.z; p.spread = 0;
        p.dx = ray.dir.x; p.dy = ray.dir.y; p.dz = ray.dir.z; p.eta_scale = 1;
        p.tx = p.ty = p.tz = 1; p.dir_pdf = 0;
        p.rx = p.ry = p.rz = 0; p.rng = rng.state;
        p.mx = p.my = p.mz = 1; p.medium = sc.camera.medium_id; p.bounces = 0;
        p.cx = p.cy = p.cz = 0; p.sample = s; p.first = 1;
        vp.path[slot] = p;
        vol_push_ray(vp, slot, vol_main_ray(sc, p));
    }
}

// ---- stage: extend = intersect() for main rays and NEE segments ----
template <bool COUNT>
__global__ void __launch_bounds__(128, 10) k_vol_extend(DScene sc, VolPool vp, int n_upper, unsigned long long *trav_counters, TraceTuning tune) {
    const int n = min(n_upper, vp.counts[VQ_EXT]);
    int ray_idx = -1; bool exhausted = false;
    Trav8 tv;
    tv.ng = make_uint2(0u, 0u); tv.tg = make_uint2(0u, 0u); tv.sp = 0; tv.top = make_uint2(0u, 0u);
    unsigned nn = 0, np = 0;
    while (true) {
        if (ray_idx >= 0 && !tv.active()) {
            const int tagged = vp.ext_slot[ray_idx];
            const int slot = tagged & ~NEE_RAY;
            Hit h; HitRec hr;
            if (tv.finish(sc, &h)) {
                hr.t = h.t; hr.u = h.u; hr.v = h.v; hr.geom = h.geom; hr.prim = h.prim; hr.ngx = h.ngx; hr.ngy = h.ngy; hr.ngz = h.ngz;
            } else {
                hr.t = 0; hr.u = hr.v = 0; hr.geom = -1; hr.prim = -1; hr.ngx = hr.ngy = hr.ngz = 0;
            }
            const bool is_nee = (tagged & NEE_RAY) != 0;
            (is_nee ? vp.nee_hit : vp.hit)[slot] = hr;
            const int q = is_nee ? VQ_NEE : VQ_MAIN;
            (is_nee ? vp.q_nee : vp.q_main)[queue_push_pos(&vp.counts[q])] = slot;
            if (COUNT) { nn += tv.n_nodes; np += tv.n_prims; }
            ray_idx = -1;
        }
        if (warp_fetch<false>(&vp.counts[VW_EXT], n, ray_idx, exhausted)) tv.init(vp.ext_ray[ray_idx].ray);
        if (!__ballot_sync(0xffffffffu, ray_idx >= 0)) break;
        const int stop_below = exhausted ? 1 : tune.refill_below;
        while (true) {
            const bool live = ray_idx >= 0;
            const bool wn = live && tv.wants_node(), wpr = live && tv.wants_prim();
            const unsigned mn = __ballot_sync(0xffffffffu, wn), mp = __ballot_sync(0xffffffffu, wpr);
            if (__popc(mn | mp) < stop_below) break;          // same loop as k_wf_extend (lj_wf_trace.cu)
            if (mp != 0u && (__popc(mp) >= tune.prim_thresh || mn == 0u)) {
                bool found = false;
                if (wpr) {
                    tv.prim_step<false, COUNT>(sc, found);
                    if (tv.wants_prim()) tv.prim_step<false, COUNT>(sc, found);
                }
            } else {
                if (wn) tv.node_step<false, COUNT>(sc);
            }
        }
    }
    if (COUNT) { atomicAdd(&trav_counters[0], (unsigned long long)nn); atomicAdd(&trav_counters[1], (unsigned long long)np); }
}

// Russian roulette + next main ray (vol_path_tracing.h:851-864); `p` holds the updated state.
LJ_D void vol_roulette_and_continue(const DScene &sc, const VolPool &vp, int slot, VolPath &p, Pcg32 &rng) {
    d3 thr = mk3(p.tx, p.ty, p.tz);
    if (p.bounces >= sc.options.rr_depth) {
        const double rr_prob = lmin(max3((1 / p.eta_scale) * thr), 0.95);
        const double u = pcg_real(rng);
        if (u > rr_prob) { vol_terminate(vp, slot, p.sample, mk3(p.rx, p.ry, p.rz)); return; }
        thr = thr / rr_prob;
    }
    p.tx = thr.x; p.ty = thr.y; p.tz = thr.z;
    p.bounces++;
    p.rng = rng.state;
    vp.path[slot] = p;
    vol_push_ray(vp, slot, vol_main_ray(sc, p));
}

LJ_D Pcg32 vol_rng(const VolPool &vp, int slot, const VolPath &p) {
    Pcg32 rng; rng.state = p.rng;
    rng.inc = (slot_stream(vp.map, slot, p.sample) << 1u) | 1u;
    return rng;
}
LJ_D Hit to_hit(const HitRec &hr) {
    Hit h; h.t = hr.t; h.u = hr.u; h.v = hr.v; h.geom = hr.geom; h.prim = hr.prim; h.ngx = hr.ngx; h.ngy = hr.ngy; h.ngz = hr.ngz;
    return h;
}

// ---- stage: track = the RNG-serial collision loops, one collision per lane per warp iteration ----
// NEE = false: free-flight sampling of the main rays (vol_path_tracing.h:556-626) -> VolTrack
// NEE = true : ratio tracking of one NEE segment (:354-410) -> transmittance / pdf products in VolNee
// Segments need between zero and max_null_collisions iterations, so a warp that walked its 32 segments
// to the end would idle most lanes; instead a lane whose loop is over is retired and handed the next
// segment from the queue (same scheme as the trace kernels).  The arithmetic is lj_volpath.cuh's.
// Grid media only: in a homogeneous medium the majorant is sigma_t itself, the loop ends at the first collision
// and stays inline in k_vol_main / k_vol_nee.
LJ_D int tracked_medium(const DScene &sc, int medium) {
    return (medium != -1 && sc.media[medium].kind == LJ_MEDIUM_HETEROGENEOUS) ? medium : -1;
}

// GREY: every medium of the scene has equal channels everywhere (lj_scene::media_grey), so every spectrum of the loops has
// three equal channels: they are carried and computed as ONE double (see lj_volpath.cuh) and written back replicated.
template <bool GREY> struct TrackSpectrum { typedef d3 type; };
template <> struct TrackSpectrum<true> { typedef double type; };
LJ_D d3 spread3(d3 a) { return a; }
LJ_D d3 spread3(double a) { return mk3(a); }
LJ_D void narrow3(d3 a, d3 &out) { out = a; }
LJ_D void narrow3(d3 a, double &out) {
    if (!(a.x == a.y && a.x == a.z)) __trap();      // a coloured spectrum in a scene flagged grey: never silently wrong
    out = a.x;
}

template <bool NEE, bool GREY>
__global__ void __launch_bounds__(128, GREY ? LJ_VOL_TRACK_GREY_MIN_BLOCKS : LJ_VOL_TRACK_MIN_BLOCKS) k_vol_track(DScene sc, VolPool vp, int n_upper, int refill_below, int budget, int min_items, int parity) {
    typedef typename TrackSpectrum<GREY>::type S;
    // work = the segments suspended by the previous launch, then the fresh ones from the extend stage
    const int *q_cont = NEE ? vp.q_cont_nee[parity] : vp.q_cont_main[parity];
    int *q_susp = NEE ? vp.q_cont_nee[parity ^ 1] : vp.q_cont_main[parity ^ 1];
    const int c_susp = (NEE ? VT_CONT_NEE0 : VT_CONT_MAIN0) + (parity ^ 1);
    const int n_cont = min(n_upper, vp.counts[(NEE ? VT_CONT_NEE0 : VT_CONT_MAIN0) + parity]);
    const int n = min(n_upper, n_cont + vp.counts[NEE ? VQ_NEE : VQ_MAIN]);
    const int *q_fresh = NEE ? vp.q_nee : vp.q_main;
    // A segment is suspended (and resumed by the next launch) only once the queue has run dry and it has done
    // `budget` more steps: while there is work to refill with, long segments cost nothing; afterwards they would
    // hold the whole iteration back.  With few segments in total nothing else could run anyway: no suspension.
    if (n < min_items) budget = 0x7fffffff;
    int *q_ready = NEE ? vp.q_nee_ready : vp.q_main_ready;
    int item = -1, slot = -1, medium = -1, steps = 0; bool exhausted = false;     // medium = -1: nothing to track for this item
    TrackStateT<S> ts; ts.active = false;
    Pcg32 rng; rng.state = 0; rng.inc = 1;
    Ray ray = make_ray(mk3(0.0), mk3(0.0), 0, 0);       // main ray, or the shadow segment
    double t_end = 0;
    bool reached = false;       // main rays: the free flight ended on the surface (its position is recomputed at retire time)
    S A, Bv, Cv;                                        // main: transmittance, dir pdf, nee pdf; NEE: T_light, p_trans_nee, p_trans_dir
    narrow3(mk3(1.0), A); narrow3(mk3(1.0), Bv); narrow3(mk3(1.0), Cv);
    bool scatter = false;
    while (true) {
        const bool running = item >= 0 && ts.active && steps < budget;
        if (item >= 0 && !running) {                    // retire: finished -> ready queue, out of budget -> suspended queue
            if (medium != -1) {
                VolTrack tr;
                const d3 A3 = spread3(A), B3 = spread3(Bv), C3 = spread3(Cv), mj3 = spread3(ts.majorant);
                if (NEE) {
                    VolNee *ne = &vp.nee[slot];
                    ne->ax = A3.x; ne->ay = A3.y; ne->az = A3.z; ne->bx = B3.x; ne->by = B3.y; ne->bz = B3.z; ne->ex = C3.x; ne->ey = C3.y; ne->ez = C3.z;
                    tr.tx = tr.ty = tr.tz = tr.dx = tr.dy = tr.dz = tr.nx = tr.ny = tr.nz = 0; tr.ox = tr.oy = tr.oz = 0;
                } else {
                    tr.tx = A3.x; tr.ty = A3.y; tr.tz = A3.z; tr.dx = B3.x; tr.dy = B3.y; tr.dz = B3.z; tr.nx = C3.x; tr.ny = C3.y; tr.nz = C3.z;
                    d3 org = ray.org;
                    if (reached) {              // ray.org = hit_position (vol_path_tracing.h:617), the expression of the fetch below
                        const VolPath *pp = &vp.path[slot];
                        org = mk3(pp->ox, pp->oy, pp->oz) + ray.dir * (double)vp.hit[slot].t;
                    }
                    tr.ox = org.x; tr.oy = org.y; tr.oz = org.z;
                }
                tr.mjx = mj3.x; tr.mjy = mj3.y; tr.mjz = mj3.z; tr.accum_t = ts.accum_t;
                tr.scatter = scatter ? 1 : 0; tr.channel = ts.channel; tr.iteration = ts.iteration; tr.pad0 = 0;
                if (!NEE || ts.active) vp.track[slot] = tr;
                vp.path[slot].rng = rng.state;
            }
            if (ts.active) q_susp[queue_push_uniform(&vp.counts[c_susp])] = slot;
            else q_ready[queue_push_uniform(&vp.counts[NEE ? VT_NEE_READY : VT_MAIN_READY])] = slot;
            item = -1;
        }
        if (warp_fetch<false>(&vp.counts[NEE ? VW_TRACK_NEE : VW_TRACK_MAIN], n, item, exhausted)) {
            const bool resumed = item < n_cont;
            slot = resumed ? q_cont[item] : q_fresh[item - n_cont];
            const VolPath *pp = &vp.path[slot];
            rng.state = pp->rng;
            rng.inc = (slot_stream(vp.map, slot, pp->sample) << 1u) | 1u;
            scatter = false; ts.active = false; steps = 0;
            if (NEE) {
                const VolNee ne = vp.nee[slot];
                const HitRec hr = vp.nee_hit[slot];
                medium = tracked_medium(sc, ne.shadow_medium);
                const d3 pos = mk3(ne.px, ne.py, ne.pz), dir_light = mk3(ne.lx, ne.ly, ne.lz), p_prime = mk3(ne.qx, ne.qy, ne.qz);
                ray = make_ray(pos, dir_light, sc.eps, (1 - sc.eps) * distance(pos, p_prime));
                t_end = distance(pos, p_prime);
                if (hr.geom >= 0) t_end = distance(pos, ray.org + ray.dir * (double)hr.t);      // |p - shadow_vertex.position|
                narrow3(mk3(ne.ax, ne.ay, ne.az), A); narrow3(mk3(ne.bx, ne.by, ne.bz), Bv); narrow3(mk3(ne.ex, ne.ey, ne.ez), Cv);
            } else {
                const HitRec hr = vp.hit[slot];
                medium = tracked_medium(sc, pp->medium);
                ray = make_ray(mk3(pp->ox, pp->oy, pp->oz), mk3(pp->dx, pp->dy, pp->dz), sc.eps, INFINITY);
                t_end = INFINITY; reached = false;
                if (hr.geom >= 0) { const d3 hit_position = ray.org + ray.dir * (double)hr.t; t_end = distance(hit_position, ray.org); }
                narrow3(mk3(1.0), A); narrow3(mk3(1.0), Bv); narrow3(mk3(1.0), Cv);
            }
            if (resumed) {
                const VolTrack tr = vp.track[slot];
                if (!NEE) { narrow3(mk3(tr.tx, tr.ty, tr.tz), A); narrow3(mk3(tr.dx, tr.dy, tr.dz), Bv); narrow3(mk3(tr.nx, tr.ny, tr.nz), Cv); }
                narrow3(mk3(tr.mjx, tr.mjy, tr.mjz), ts.majorant); ts.accum_t = tr.accum_t; ts.channel = tr.channel; ts.iteration = tr.iteration;
                ts.maj_c = sp_comp(ts.majorant, ts.channel); ts.inv_mx = 1.0 / sp_max(ts.majorant); ts.active = true;
            } else if (medium != -1) track_begin(sc, sc.media[medium], ray, rng, ts);
        }
        if (!__ballot_sync(0xffffffffu, item >= 0)) break;
        const int stop_below = exhausted ? 1 : refill_below;
        while (true) {
            if (item >= 0 && ts.active && steps < budget) {
                if (NEE) ratio_track_step(sc, sc.media[medium], ray.org, ray.dir, t_end, rng, ts, A, Bv, Cv);
                else free_flight_step(sc, sc.media[medium], ray, t_end, ray.org, rng, ts, A, Bv, Cv, scatter, &reached);
                steps += exhausted ? 1 : 0;
            }
            if (__popc(__ballot_sync(0xffffffffu, item >= 0 && ts.active && steps < budget)) < stop_below) break;
        }
    }
}

// ---- stage: main (vol_path_tracing.h:543-760 up to the light sample of the NEE, :311-320) ----
__global__ void __launch_bounds__(128, LJ_VOL_MAIN_MIN_BLOCKS) k_vol_main(DScene sc, VolPool vp, int n_upper, const int *queue, int count_idx) {
    const int n = min(n_upper, vp.counts[count_idx]);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned together = __activemask();
        const int slot = queue[i];
        VolPath p = vp.path[slot];
        const HitRec hr = vp.hit[slot];
        Pcg32 rng = vol_rng(vp, slot, p);
        Ray ray = vol_main_ray(sc, p);
        const d3 org0 = ray.org;
        const bool hit = hr.geom >= 0;
        PathVertex vertex;
        vertex.position = mk3(0.0); vertex.material_id = 0; vertex.shape_id = 0;
        double t_hit = INFINITY;
        if (hit) {
            RayDiff rd; rd.radius = 0; rd.spread = p.spread;
            vertex = make_vertex(sc, ray, to_hit(hr), rd);
            t_hit = distance(vertex.position, ray.org);
        }
        __syncwarp(together);
        d3 transmittance = mk3(1.0), trans_dir_pdf = mk3(1.0), trans_nee_pdf = mk3(1.0);
        bool scatter = false;
        d3 thr = mk3(p.tx, p.ty, p.tz), rad = mk3(p.rx, p.ry, p.rz);
        d3 multi = mk3(p.mx, p.my, p.mz);
        bool dead = false;
        if (p.medium != -1 && sc.media[p.medium].kind != LJ_MEDIUM_HETEROGENEOUS) {
            scatter = free_flight(sc, sc.media[p.medium], ray, t_hit, vertex.position, rng, transmittance, trans_dir_pdf, trans_nee_pdf);
            multi = multi * trans_dir_pdf;
        } else if (p.medium != -1) {
            // free-flight sampling was done by k_vol_track<false> (it also advanced p.rng)
            const VolTrack tr = vp.track[slot];
            transmittance = mk3(tr.tx, tr.ty, tr.tz); trans_dir_pdf = mk3(tr.dx, tr.dy, tr.dz); trans_nee_pdf = mk3(tr.nx, tr.ny, tr.nz);
            ray.org = mk3(tr.ox, tr.oy, tr.oz); scatter = tr.scatter != 0;
            multi = multi * trans_dir_pdf;
        } else {
            if (hit) ray.org = vertex.position;
            else { vol_terminate(vp, slot, p.sample, mk3(0.0)); dead = true; }          // :637 returns zero, not the radiance so far
        }
        __syncwarp(together);
        if (dead) continue;
        thr = thr * (transmittance / average(trans_dir_pdf));

        if (!scatter && hit && sc.shapes[vertex.shape_id].area_light_id >= 0) {
            d3 Le = vertex_emission(sc, vertex, -ray.dir);
            if (p.bounces == 0) {
                rad += thr * Le;
                vol_terminate(vp, slot, p.sample, rad);
                continue;
            }
            PointAndNormal light_point; light_point.position = vertex.position; light_point.normal = vertex.geometry_normal;
            const int light_id = sc.shapes[vertex.shape_id].area_light_id;
            const lj_light &light = sc.lights[light_id];
            const d3 cache = mk3(p.cx, p.cy, p.cz);
            d3 pdf_nee = (__ldg(sc.light_pmf + light_id) * pdf_point_on_light(sc, light, light_point, cache)) * trans_nee_pdf;
            double jacobian = lmax(-dot(-ray.dir, light_point.normal), 0.0) / distance_squared(cache, light_point.position);
            d3 pdf_phase = (p.dir_pdf * multi) * jacobian;
            d3 w = (pdf_phase * pdf_phase) / (pdf_phase * pdf_phase + pdf_nee * pdf_nee);
            rad += thr * Le * w;
        }

        p.tx = thr.x; p.ty = thr.y; p.tz = thr.z; p.rx = rad.x; p.ry = rad.y; p.rz = rad.z;
        p.mx = multi.x; p.my = multi.y; p.mz = multi.z;
        p.ox = ray.org.x; p.oy = ray.org.y; p.oz = ray.org.z;
        p.first = 0;

        if (!scatter && hit && vertex.material_id == -1) {             // index-matched surface: pass through
            p.medium = update_medium(vertex, ray.dir, p.medium);
            p.ox = vertex.position.x; p.oy = vertex.position.y; p.oz = vertex.position.z;
            p.bounces++;
            p.rng = rng.state;
            vp.path[slot] = p;
            vol_push_ray(vp, slot, vol_main_ray(sc, p));
            continue;
        }
        const int max_depth = sc.options.max_depth;
        if (p.bounces >= max_depth - 1 && max_depth != -1) { vol_terminate(vp, slot, p.sample, rad); continue; }

        const bool nee_medium = scatter && p.medium != -1;
        if (!nee_medium && !hit) {                                     // nothing to scatter on: straight to the roulette
            vol_roulette_and_continue(sc, vp, slot, p, rng);
            continue;
        }
        // next_event_estimation_final, :311-332: light sample and the first segment
        d2 light_uv; light_uv.x = pcg_real(rng); light_uv.y = pcg_real(rng);
        const double light_w = pcg_real(rng);
        const double shape_w = pcg_real(rng);
        const int light_id = sample_table1d(sc.light_cdf, sc.num_lights, light_w);
        const lj_light &light = sc.lights[light_id];
        PointAndNormal pl = sample_point_on_light(sc, light, ray.org, light_uv, shape_w);
        const d3 dir_light = normalize(pl.position - ray.org);
        VolNee ne;
        ne.px = ray.org.x; ne.py = ray.org.y; ne.pz = ray.org.z; ne.light_id = light_id; ne.shadow_medium = p.medium;
        ne.lx = dir_light.x; ne.ly = dir_light.y; ne.lz = dir_light.z; ne.shadow_bounces = 0;
        ne.kind = nee_medium ? -1 : sc.materials[vertex.material_id].kind;
        ne.qx = pl.position.x; ne.qy = pl.position.y; ne.qz = pl.position.z; ne.blocked = 0; ne.pad0 = 0;
        ne.nx = pl.normal.x; ne.ny = pl.normal.y; ne.nz = pl.normal.z; ne.pad1 = 0;
        ne.ax = ne.ay = ne.az = 1; ne.bx = ne.by = ne.bz = 1; ne.ex = ne.ey = ne.ez = 1; ne.pad2 = ne.pad3 = ne.pad4 = ne.pad5 = 0;
        ne.gx = org0.x; ne.gy = org0.y; ne.gz = org0.z;
        p.rng = rng.state;
        vp.path[slot] = p;
        vp.nee[slot] = ne;
        const double eps = sc.eps;
        vol_push_ray(vp, slot | NEE_RAY, make_ray(ray.org, dir_light, eps, (1 - eps) * distance(ray.org, pl.position)));
    }
}

// ---- stage: nee = one segment of next_event_estimation_final's loop (:335-439) ----
__global__ void __launch_bounds__(128, LJ_VOL_MAIN_MIN_BLOCKS) k_vol_nee(DScene sc, VolPool vp, int n_upper, const int *queue, int count_idx) {
    const int n = min(n_upper, vp.counts[count_idx]);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned together = __activemask();
        const int slot = queue[i];
        VolNee ne = vp.nee[slot];
        const HitRec hr = vp.nee_hit[slot];
        VolPath p;                           // only the bounce count is needed here
        p.bounces = vp.path[slot].bounces;
        d3 pos = mk3(ne.px, ne.py, ne.pz);
        const d3 dir_light = mk3(ne.lx, ne.ly, ne.lz), p_prime = mk3(ne.qx, ne.qy, ne.qz);
        const double eps = sc.eps;
        const Ray shadow_ray = make_ray(pos, dir_light, eps, (1 - eps) * distance(pos, p_prime));
        const bool hit = hr.geom >= 0;
        PathVertex sv;
        sv.material_id = 0;
        double next_t = distance(pos, p_prime);
        if (hit) {
            RayDiff rd; rd.radius = 0; rd.spread = 0;
            sv = make_vertex(sc, shadow_ray, to_hit(hr), rd);
            next_t = distance(pos, sv.position);
        }
        __syncwarp(together);
        // grid media: the ratio tracking of this segment was done by k_vol_track<true> (ne.a/b/e and the rng are up to date)
        if (ne.shadow_medium != -1 && sc.media[ne.shadow_medium].kind != LJ_MEDIUM_HETEROGENEOUS) {
            VolPath *pp = &vp.path[slot];
            p.rng = pp->rng; p.sample = pp->sample;
            Pcg32 rng = vol_rng(vp, slot, p);
            d3 T = mk3(ne.ax, ne.ay, ne.az), ptn = mk3(ne.bx, ne.by, ne.bz), ptd = mk3(ne.ex, ne.ey, ne.ez);
            ratio_track_segment(sc, sc.media[ne.shadow_medium], shadow_ray, pos, dir_light, next_t, rng, T, ptn, ptd);
            ne.ax = T.x; ne.ay = T.y; ne.az = T.z; ne.bx = ptn.x; ne.by = ptn.y; ne.bz = ptn.z; ne.ex = ptd.x; ne.ey = ptd.y; ne.ez = ptd.z;
            pp->rng = rng.state;
        }
        __syncwarp(together);
        bool done = false;
        if (!hit) done = true;
        else if (sv.material_id >= 0) { ne.blocked = 1; done = true; }
        else {
            ne.shadow_bounces++;
            const int max_depth = sc.options.max_depth;
            if (max_depth != -1 && p.bounces + ne.shadow_bounces >= max_depth) { ne.blocked = 1; done = true; }
            else {
                ne.shadow_medium = update_medium(sv, dir_light, ne.shadow_medium);
                pos = pos + next_t * dir_light;
                ne.px = pos.x; ne.py = pos.y; ne.pz = pos.z;
            }
        }
        vp.nee[slot] = ne;
        if (done) {
            const int q = ne.kind < 0 ? VQ_FIN_MEDIUM : VQ_FIN_KIND0 + ne.kind;
            int *queue = ne.kind < 0 ? vp.q_fin_medium : vp.q_fin_kind[ne.kind];
            queue[queue_push_pos(&vp.counts[q])] = slot;
        } else {
            vol_push_ray(vp, slot | NEE_RAY, make_ray(pos, dir_light, eps, (1 - eps) * distance(pos, p_prime)));
        }
    }
}

// NEE contribution once all segments are through (vol_path_tracing.h:443-491).  `evaluated_f` and
// `pdf_scatter` (phase or BSDF pdf of dir_light) come from the caller; returns contrib * w.
LJ_D d3 vol_nee_value(const DScene &sc, const VolNee &ne, d3 p_origin, d3 evaluated_f, double pdf_scatter) {
    const lj_light &light = sc.lights[ne.light_id];
    PointAndNormal pl; pl.position = mk3(ne.qx, ne.qy, ne.qz); pl.normal = mk3(ne.nx, ne.ny, ne.nz);
    const d3 dir_light = mk3(ne.lx, ne.ly, ne.lz);
    const d3 T = mk3(ne.ax, ne.ay, ne.az), ptn = mk3(ne.bx, ne.by, ne.bz), ptd = mk3(ne.ex, ne.ey, ne.ez);
    d3 Le = light_emission(sc, light, -dir_light, 0.0, pl);
    double jacobian = lmax(-dot(dir_light, pl.normal), 0.0) / distance_squared(p_origin, pl.position);
    d3 pdf_nee = (__ldg(sc.light_pmf + ne.light_id) * pdf_point_on_light(sc, light, pl, p_origin)) * ptn;
    d3 pdf_dir = (pdf_scatter * jacobian) * ptd;
    d3 contrib = T * evaluated_f * Le * jacobian / average(pdf_nee);
    d3 w = (pdf_nee * pdf_nee) / (pdf_nee * pdf_nee + pdf_dir * pdf_dir);
    return contrib * w;
}

// ---- stage: finish after a scattering event inside a medium (:736-784) ----
__global__ void __launch_bounds__(128) k_vol_finish_medium(DScene sc, VolPool vp, int n_upper) {
    const int n = min(n_upper, vp.counts[VQ_FIN_MEDIUM]);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int slot = vp.q_fin_medium[i];
        VolPath p = vp.path[slot];
        const VolNee ne = vp.nee[slot];
        Pcg32 rng = vol_rng(vp, slot, p);
        const d3 org = mk3(p.ox, p.oy, p.oz), dir = mk3(p.dx, p.dy, p.dz);
        const d3 dir_view = -dir, dir_light = mk3(ne.lx, ne.ly, ne.lz);
        const lj_medium &media = sc.media[p.medium];
        d3 nee = mk3(0.0);
        if (!ne.blocked && max3(mk3(ne.ax, ne.ay, ne.az)) > 0) {
            pcg_real(rng); pcg_real(rng);                    // the two unused phase_uv draws, :479
            nee = vol_nee_value(sc, ne, org, phase_eval(media, dir_view, dir_light), phase_pdf(media, dir_view, dir_light));
        }
        const d3 sigma_s = medium_sigma_s(sc, media, org);
        d3 thr = mk3(p.tx, p.ty, p.tz), rad = mk3(p.rx, p.ry, p.rz);
        rad += thr * sigma_s * nee;
        if (max3(nee) > 0) { p.cx = org.x; p.cy = org.y; p.cz = org.z; }
        d2 phase_uv; phase_uv.x = pcg_real(rng); phase_uv.y = pcg_real(rng);
        const d3 next_dir = phase_sample(media, dir_view, phase_uv);
        const double ppdf = phase_pdf(media, dir_view, next_dir);
        thr = thr * (phase_eval(media, dir_view, next_dir) / ppdf * sigma_s);
        p.dx = next_dir.x; p.dy = next_dir.y; p.dz = next_dir.z;
        p.dir_pdf = ppdf;
        p.mx = p.my = p.mz = 1;
        p.tx = thr.x; p.ty = thr.y; p.tz = thr.z; p.rx = rad.x; p.ry = rad.y; p.rz = rad.z;
        vol_roulette_and_continue(sc, vp, slot, p, rng);
    }
}

// ---- stage: finish at a surface, one launch per material kind (:785-848) ----
#ifndef LJ_VOL_FINISH_MIN_BLOCKS
#define LJ_VOL_FINISH_MIN_BLOCKS 4
#endif
template <int KIND>
__global__ void __launch_bounds__(128, LJ_VOL_FINISH_MIN_BLOCKS) k_vol_finish_surface(DScene sc, VolPool vp, int n_upper) {
    const int n = min(n_upper, vp.counts[VQ_FIN_KIND0 + KIND]);
    const int *queue = vp.q_fin_kind[KIND];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned together = __activemask();
        const int slot = queue[i];
        VolPath p = vp.path[slot];
        const VolNee ne = vp.nee[slot];
        const HitRec hr = vp.hit[slot];
        Pcg32 rng = vol_rng(vp, slot, p);
        const d3 org = mk3(p.ox, p.oy, p.oz), dir = mk3(p.dx, p.dy, p.dz);
        const d3 dir_view = -dir, dir_light = mk3(ne.lx, ne.ly, ne.lz);
        RayDiff rd; rd.radius = 0; rd.spread = p.spread;
        const PathVertex vertex = make_vertex(sc, make_ray(mk3(ne.gx, ne.gy, ne.gz), dir, sc.eps, INFINITY), to_hit(hr), rd);
        MatParams mat = load_mat_params(sc, vertex);
        mat.kind = KIND;
        __syncwarp(together);
        d3 nee = mk3(0.0);
        if (!ne.blocked && max3(mk3(ne.ax, ne.ay, ne.az)) > 0) {
            const d3 f = bsdf_eval(mat, vertex, dir_view, dir_light);
            const double pdf_bsdf = bsdf_pdf(mat, vertex, dir_view, dir_light);
            if (pdf_bsdf > 0) nee = vol_nee_value(sc, ne, org, f, pdf_bsdf);       // :466-469: pdf <= 0 returns zero
        }
        __syncwarp(together);
        d3 thr = mk3(p.tx, p.ty, p.tz), rad = mk3(p.rx, p.ry, p.rz);
        rad += thr * nee;
        if (max3(nee) > 0) { p.cx = org.x; p.cy = org.y; p.cz = org.z; }
        p.rx = rad.x; p.ry = rad.y; p.rz = rad.z;
        d2 bsdf_uv; bsdf_uv.x = pcg_real(rng); bsdf_uv.y = pcg_real(rng);
        const double bsdf_w = pcg_real(rng);
        BSDFSample bs;
        if (!bsdf_sample(mat, vertex, dir_view, bsdf_uv, bsdf_w, bs)) { vol_terminate(vp, slot, p.sample, rad); continue; }
        const d3 dir_bsdf = bs.dir_out;
        p.dx = dir_bsdf.x; p.dy = dir_bsdf.y; p.dz = dir_bsdf.z;
        if (bs.eta == 0) {
            p.spread = rd_reflect(rd, vertex.mean_curvature, bs.roughness);
        } else {
            p.spread = rd_refract(rd, vertex.mean_curvature, bs.eta, bs.roughness);
            p.eta_scale /= (bs.eta * bs.eta);
            p.medium = update_medium(vertex, dir_bsdf, p.medium);
        }
        const d3 f = bsdf_eval(mat, vertex, dir_view, dir_bsdf);
        const double pdf_bsdf = bsdf_pdf(mat, vertex, dir_view, dir_bsdf);
        thr = thr * (f / pdf_bsdf);
        p.tx = thr.x; p.ty = thr.y; p.tz = thr.z;
        vol_roulette_and_continue(sc, vp, slot, p, rng);
    }
}

__global__ void k_vol_reset(int *counts, int a, int b) {
    if (threadIdx.x == 0) { if (a >= 0) counts[a] = 0; if (b >= 0) counts[b] = 0; }
}
__global__ void k_vol_reset_stage_queues(int *counts, int parity) {
    if (threadIdx.x < VQ_COUNT - VQ_MAIN) counts[VQ_MAIN + threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        counts[VW_EXT] = 0; counts[VW_TRACK_MAIN] = 0; counts[VW_TRACK_NEE] = 0;
        counts[VT_MAIN_READY] = 0; counts[VT_NEE_READY] = 0;
        counts[VT_CONT_MAIN0 + parity] = 0; counts[VT_CONT_NEE0 + parity] = 0;     // consumed by this iteration
    }
}
__global__ void k_vol_init(VolPool vp, int s_begin) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < vp.capacity; i += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(i / vp.map.npix_local);
        VolPath p; memset(&p, 0, sizeof p);
        p.sample = s_begin + k - vp.map.K;
        vp.path[i] = p;
        vp.q_regen[i] = (int)i;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) vp.counts[VQ_REGEN] = vp.capacity;
}
template <int KIND>
void launch_finish(const DScene &d, const VolPool &vp, int n_upper, int grid_cap) {
    int grid = (n_upper + 127) / 128;
    if (grid > grid_cap) grid = grid_cap;
    if (grid < 1) grid = 1;
    k_vol_finish_surface<KIND><<<grid, 128>>>(d, vp, n_upper);
}

} // namespace

static int ensure_vol_pool(lj_scene *s, const SlotMap &map, VolPool &vp) {
    const size_t cap = (size_t)map.K * map.npix_local;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t o_path = take(cap * sizeof(VolPath)), o_nee = take(cap * sizeof(VolNee));
    const size_t o_hit = take(cap * sizeof(HitRec)), o_nhit = take(cap * sizeof(HitRec)), o_acc = take((size_t)map.K0 * map.npix_local * 3 * sizeof(double));
    const size_t o_trk = take(cap * sizeof(VolTrack));
    const size_t o_er = take(cap * sizeof(ExtRec)), o_es = take(cap * sizeof(int));
    const size_t o_qr = take(cap * sizeof(int)), o_qm = take(cap * sizeof(int)), o_qn = take(cap * sizeof(int)), o_qfm = take(cap * sizeof(int));
    size_t o_qk[LJ_MAT_KIND_COUNT];
    for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) o_qk[k] = take(cap * sizeof(int));
    size_t o_tq[6];
    for (int k = 0; k < 6; k++) o_tq[k] = take(cap * sizeof(int));
    const size_t o_cnt = take(64 * sizeof(int));
    void *arena = nullptr;
    if (int rc = lj_ctx_arena(s->ctx, off, &arena)) return rc;
    char *b = (char *)arena;
    vp.capacity = (int)cap; vp.map = map;
    vp.path = (VolPath *)(b + o_path); vp.nee = (VolNee *)(b + o_nee); vp.hit = (HitRec *)(b + o_hit); vp.nee_hit = (HitRec *)(b + o_nhit); vp.track = (VolTrack *)(b + o_trk);
    vp.accum = (double *)(b + o_acc); vp.ext_ray = (ExtRec *)(b + o_er); vp.ext_slot = (int *)(b + o_es);
    vp.q_regen = (int *)(b + o_qr); vp.q_main = (int *)(b + o_qm); vp.q_nee = (int *)(b + o_qn); vp.q_fin_medium = (int *)(b + o_qfm);
    for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) vp.q_fin_kind[k] = (int *)(b + o_qk[k]);
    vp.q_main_ready = (int *)(b + o_tq[0]); vp.q_nee_ready = (int *)(b + o_tq[1]);
    vp.q_cont_main[0] = (int *)(b + o_tq[2]); vp.q_cont_main[1] = (int *)(b + o_tq[3]);
    vp.q_cont_nee[0] = (int *)(b + o_tq[4]); vp.q_cont_nee[1] = (int *)(b + o_tq[5]);
    vp.counts = (int *)(b + o_cnt);
    return LJ_OK;
}

int lj_render_volwave(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    const DScene &d = s->d;
    const long long npix = (long long)rl.width * rl.height;
    const int nsamp = rl.s_end - rl.s_begin;
    const bool grid_media = s->grid_media;      // any heterogeneous medium -> run the tracking kernels
    const bool media_grey = s->media_grey && !getenv("LJ_NO_GREY_TRACK");      // every medium has equal channels: one-channel tracking
    // slots in flight: grid media want many (the iteration count, not the arithmetic, bounds them: each iteration
    // waits for its slowest tracking chain), homogeneous scenes run best when the records stay close to L2 size
    long long target = rl.p.paths_in_flight > 0 ? rl.p.paths_in_flight : (grid_media ? (16LL << 20) : (1LL << 20));
    // (a target just below a power of two of sample lanes is rounded UP: K is a power of two, see SlotMap)
    const SlotMap map = lj_make_slot_map(rl.width, rl.height, rl.s_begin, rl.s_end, rl.p.tile_rank, rl.p.tile_world, target + target / 2);
    if ((long long)map.npix_local * map.K > 0x3fffffffLL / 4) { lj_set_error("volumetric wavefront: pool too large"); return LJ_EINVAL; }
    VolPool vp;
    int rc = ensure_vol_pool(s, map, vp);
    if (rc) return rc;
    const int capacity = vp.capacity;
    const bool *kind_present = s->kind_present;
    auto env_int = [](const char *name, int dflt) { const char *v = getenv(name); return v ? atoi(v) : dflt; };
    TraceTuning tune; tune.refill_below = env_int("LJ_REFILL_BELOW", 24); tune.prim_thresh = env_int("LJ_PRIM_THRESH", 8);
    const int track_refill = env_int("LJ_TRACK_REFILL_BELOW", 24);
    const int track_budget = std::max(1, env_int("LJ_TRACK_BUDGET", 16));     // collision steps a segment may take after the queue ran dry
    const int track_min_items = env_int("LJ_TRACK_MIN_ITEMS", 2048);
    const int sm = s->ctx->sm_count;
    const int trace_grid_cap = sm * 16, stage_grid_cap = sm * 16;
    const bool counting = (rl.p.flags & 1) != 0;
    const bool stage_timing = (rl.p.flags & 2) != 0;
    unsigned long long *trav = (unsigned long long *)(vp.counts + 32);

    LJ_CUDA(cudaMemsetAsync(vp.counts, 0, 64 * sizeof(int)));
    LJ_CUDA(cudaMemsetAsync(vp.accum, 0, (size_t)map.K0 * map.npix_local * 3 * sizeof(double)));
    if (map.tile_world > 0) LJ_CUDA(cudaMemsetAsync(dev_sum, 0, sizeof(double) * 3 * (size_t)npix));
    k_vol_init<<<sm * 8, 256>>>(vp, rl.s_begin);
    LJ_CUDA(cudaGetLastError());
    // Queue sizes live on the device and every kernel clamps its range to them, so the host does not have to
    // know them to launch the next iteration: it sizes grids from the last sizes it has seen (or the pool
    // capacity) and reads the counts back through a ring of pinned buffers, `depth` iterations behind.  The
    // iterations launched after the one that emptied the pool find empty queues and do nothing.
    const int depth = stage_timing ? 1 : std::max(1, std::min(8, env_int("LJ_WF_PIPELINE", 4)));
    static_assert(VT_END <= LJ_RING_INTS, "ring slot too small for the queue counters");
    if ((rc = lj_ctx_ring(s->ctx)) != LJ_OK) return rc;
    int *h_ring = s->ctx->h_ring;
    cudaEvent_t *done = s->ctx->ring_done;
    cudaEvent_t *ev = s->ctx->ring_ev[0];           // stage timing runs with depth 1: ring slot 0's timing events
    int n_regen = capacity, n_ext = 0, n_cont = 0;  // sizes after the last retired iteration (n_cont: suspended segments)
    unsigned long long launches = 2, rays_primary = 0, rays_ext = 0;
    unsigned long long launched = 0, retired = 0;
    auto grid_for = [](int n, int cap) { int g = (n + 127) / 128; return g < 1 ? 1 : (g > cap ? cap : g); };
    bool finished = false;
    int loop_rc = LJ_OK;
    while (!finished) {
        const bool exact = launched == retired;     // nothing in flight: the sizes are the true ones
        const int regen_upper = exact ? n_regen : capacity;
        const int upper = exact ? n_ext + n_regen : capacity;
        if (stage_timing) cudaEventRecord(ev[0]);
        if (regen_upper > 0) {
            k_vol_generate<<<grid_for(regen_upper, stage_grid_cap), 128>>>(d, vp, regen_upper, rl.s_end);
            launches++;
        }
        k_vol_reset<<<1, 32>>>(vp.counts, VQ_REGEN, -1);
        if (stage_timing) cudaEventRecord(ev[1]);
        if (counting) k_vol_extend<true><<<grid_for(upper, trace_grid_cap), 128>>>(d, vp, upper, trav, tune);
        else k_vol_extend<false><<<grid_for(upper, trace_grid_cap), 128>>>(d, vp, upper, trav, tune);
        k_vol_reset<<<1, 32>>>(vp.counts, VQ_EXT, -1);
        if (stage_timing) cudaEventRecord(ev[2]);
        const int parity = (int)(launched & 1);
        // with segments suspended from earlier launches more than `upper` slots may be in the tracking queues
        const int track_upper = exact ? upper + n_cont : capacity;
        if (grid_media) {
            if (media_grey) {
                k_vol_track<false, true><<<grid_for(track_upper, trace_grid_cap), 128>>>(d, vp, track_upper, track_refill, track_budget, track_min_items, parity);
                k_vol_track<true, true><<<grid_for(track_upper, trace_grid_cap), 128>>>(d, vp, track_upper, track_refill, track_budget, track_min_items, parity);
            } else {
                k_vol_track<false, false><<<grid_for(track_upper, trace_grid_cap), 128>>>(d, vp, track_upper, track_refill, track_budget, track_min_items, parity);
                k_vol_track<true, false><<<grid_for(track_upper, trace_grid_cap), 128>>>(d, vp, track_upper, track_refill, track_budget, track_min_items, parity);
            }
            k_vol_main<<<grid_for(track_upper, stage_grid_cap), 128>>>(d, vp, track_upper, vp.q_main_ready, VT_MAIN_READY);
            k_vol_nee<<<grid_for(track_upper, stage_grid_cap), 128>>>(d, vp, track_upper, vp.q_nee_ready, VT_NEE_READY);
        } else {
            k_vol_main<<<grid_for(upper, stage_grid_cap), 128>>>(d, vp, upper, vp.q_main, VQ_MAIN);
            k_vol_nee<<<grid_for(upper, stage_grid_cap), 128>>>(d, vp, upper, vp.q_nee, VQ_NEE);
        }
        k_vol_finish_medium<<<grid_for(track_upper, stage_grid_cap), 128>>>(d, vp, track_upper);
        if (kind_present[0]) launch_finish<0>(d, vp, track_upper, stage_grid_cap);
        if (kind_present[1]) launch_finish<1>(d, vp, track_upper, stage_grid_cap);
        if (kind_present[2]) launch_finish<2>(d, vp, track_upper, stage_grid_cap);
        if (kind_present[3]) launch_finish<3>(d, vp, track_upper, stage_grid_cap);
        if (kind_present[4]) launch_finish<4>(d, vp, track_upper, stage_grid_cap);
        if (kind_present[5]) launch_finish<5>(d, vp, track_upper, stage_grid_cap);
        if (kind_present[6]) launch_finish<6>(d, vp, track_upper, stage_grid_cap);
        if (kind_present[7]) launch_finish<7>(d, vp, track_upper, stage_grid_cap);
        if (kind_present[8]) launch_finish<8>(d, vp, track_upper, stage_grid_cap);
        if (stage_timing) cudaEventRecord(ev[3]);
        const int ring_slot = (int)(launched % depth);
        cudaMemcpyAsync(h_ring + ring_slot * VT_END, vp.counts, sizeof(int) * VT_END, cudaMemcpyDeviceToHost);
        cudaEventRecord(done[ring_slot]);
        k_vol_reset_stage_queues<<<1, 32>>>(vp.counts, parity);
        launches += grid_media ? 10 : 8;
        for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) launches += kind_present[k] ? 1 : 0;
        launched++;
        // retire what has completed; block only when the ring is full
        while (retired < launched) {
            const int rs = (int)(retired % depth);
            if (launched - retired < (unsigned long long)depth && cudaEventQuery(done[rs]) == cudaErrorNotReady) break;
            const cudaError_t e = cudaEventSynchronize(done[rs]);
            if (e != cudaSuccess) { lj_set_error("volumetric wavefront iteration failed: %s", cudaGetErrorString(e)); loop_rc = LJ_ECUDA; finished = true; break; }
            const int *h_counts = h_ring + rs * VT_END;
            const int par_out = (int)((retired & 1) ^ 1);       // the queues this iteration suspended segments into
            const int traced = h_counts[VQ_MAIN] + h_counts[VQ_NEE];
            // the generate stage pushed n_regen' <= n_regen primaries this iteration: traced - n_ext of them
            const int prim_now = traced - n_ext > 0 ? traced - n_ext : 0;
            rays_primary += prim_now; rays_ext += traced - prim_now;
            if (stage_timing) {
                float ms;
                cudaEventElapsedTime(&ms, ev[0], ev[1]); st->generate_ms += ms;
                cudaEventElapsedTime(&ms, ev[1], ev[2]); st->extend_ms += ms;
                cudaEventElapsedTime(&ms, ev[2], ev[3]); st->shade_ms += ms;
            }
            n_regen = h_counts[VQ_REGEN];
            n_ext = h_counts[VQ_EXT];
            n_cont = h_counts[VT_CONT_MAIN0 + par_out] + h_counts[VT_CONT_NEE0 + par_out];
            st->iterations++;
            retired++;
            if (n_regen == 0 && n_ext == 0 && n_cont == 0) { finished = true; break; }
        }
    }
    cudaStreamSynchronize(0);           // drains the (empty) iterations still in flight before the ring is freed
    if (loop_rc != LJ_OK) return loop_rc;
    k_slot_sum<<<(unsigned)(((long long)map.npix_local * 3 + 255) / 256), 256>>>(map, vp.accum, dev_sum);
    LJ_CUDA(cudaGetLastError());
    if (counting) {
        unsigned long long h_trav[2];
        LJ_CUDA(cudaMemcpy(h_trav, trav, sizeof h_trav, cudaMemcpyDeviceToHost));
        st->nodes_visited = h_trav[0]; st->prims_tested = h_trav[1];
    }
    launches++;
    st->kernel_launches += launches;
    st->rays_primary += rays_primary; st->rays_extension += rays_ext;
    st->samples = (uint64_t)npix * (uint64_t)nsamp;
    return LJ_OK;
}
