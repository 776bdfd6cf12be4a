// Wavefront path tracer (mode B streams): path_tracing.h:7-325 split into stages
//   generate -> extend (closest hit) -> shade[material kind] / miss -> shadow (any hit) -> accumulate
// connected by queues compacted with warp-aggregated atomics.  A pool of path
// slots is kept full by regeneration: slot (pixel, k) renders samples
// s_begin+k, s_begin+k+K, ... of its pixel one after the other and owns its own
// accumulator, so no atomics touch radiance and the per-pixel sum order is fixed.
//
// Per-slot state lives in HBM as 32/64-byte records (gathered by slot id after
// compaction, so records are whole 32-byte sectors); queues and shadow / extend
// ray records are written and read by queue position (coalesced).
#include "lj_internal.h"
#include "lj_materials.cuh"
#include "lj_traverse8.cuh"
#include "lj_wave_common.cuh"
#include "lj_wf_pool.cuh"
#include <cstdio>
#include <cstdlib>

namespace {

LJ_D void terminate_path(const WfPool &wp, int slot, int sample, d3 radiance) {
    double *a = wp.accum + accum_offset(wp.map, slot, sample);
    a[0] += radiance.x; a[1] += radiance.y; a[2] += radiance.z;      // render.cpp:92: radiance += path_tracing(...)
    wp.q_regen[queue_push_uniform(&wp.counts[Q_REGEN])] = slot;
}

LJ_D void push_extend(const WfPool &wp, int slot, const Ray &r) {
    int pos = queue_push_uniform(&wp.counts[Q_EXT]);
    ExtRec e; e.ray = to_rayf(r);
    wp.ext_ray[pos] = e;
    wp.ext_slot[pos] = slot;
}

// Where a stage's outputs go.  QueueSink: the wavefront queues (warp-aggregated pushes; the stage kernels).  LaneSink: the
// lane's own registers (k_wf_tail, which runs a path to its end in one lane).  The per-slot arithmetic is the same code.
struct QueueSink {
    const WfPool &wp;
    LJ_D QueueSink(const WfPool &w) : wp(w) {}
    LJ_D void extend(int slot, const Ray &r) { push_extend(wp, slot, r); }
    LJ_D void shadow(int slot, const RayF &ray, d3 contrib) {
        ShadowRec s; s.ray = ray; s.cx = contrib.x; s.cy = contrib.y; s.cz = contrib.z; s.slot = slot; s._pad = 0;
        wp.shadow[queue_push_uniform(&wp.counts[Q_SHADOW])] = s;
    }
    LJ_D void terminate(int slot, int sample, d3 radiance) { terminate_path(wp, slot, sample, radiance); }
    LJ_D void finish(int slot) { wp.q_finish[queue_push_uniform(&wp.counts[Q_FINISH])] = slot; }   // last NEE term still in the shadow queue
};
struct LaneSink {
    const WfPool &wp;
    int what;                  // 0: path ended and was accumulated, 1: `next` is its extension ray, 2: ended, accumulate after the shadow ray
    bool has_shadow;
    RayF next, shadow_ray; double cx, cy, cz;
    LJ_D LaneSink(const WfPool &w) : wp(w), what(0), has_shadow(false) {}
    LJ_D void extend(int, const Ray &r) { next = to_rayf(r); what = 1; }
    LJ_D void shadow(int, const RayF &ray, d3 contrib) { shadow_ray = ray; cx = contrib.x; cy = contrib.y; cz = contrib.z; has_shadow = true; }
    LJ_D void terminate(int slot, int sample, d3 radiance) {
        double *a = wp.accum + accum_offset(wp.map, slot, sample);
        a[0] += radiance.x; a[1] += radiance.y; a[2] += radiance.z;
        what = 0;
    }
    LJ_D void finish(int) { what = 2; }
};

// ---- stage: generate (path_tracing.h:9-14, render.cpp:88-93 loop head) ----
// The next sample of a slot: false when the slot has none left (or lies outside the film under a tile partition).
LJ_D bool generate_slot(const DScene &sc, const WfPool &wp, int slot, int s_end, Ray &ray) {
    ThrRec tr = wp.thr[slot];
    int s = tr.sample + wp.map.K;
    int pix, x, y;
    const bool on_film = slot_pixel(wp.map, slot, pix, x, y);
    if (s >= s_end || !on_film) { wp.hit[slot].geom = -2; return false; }   // slot exhausted: never traced again (see k_wf_shade, slot order)
    int w = sc.camera.width, h = sc.camera.height;
    Pcg32 rng = pcg_init((unsigned long long)s * (unsigned long long)wp.map.npix_total + (unsigned long long)pix);
    double u_first = pcg_real(rng), u_second = pcg_real(rng);   // right-to-left argument evaluation, see lj_path.cuh
    ray = sample_primary(sc.camera, mk2((x + u_second) / w, (y + u_first) / h));
    RayDiff rd = init_ray_differential(w, h);
    RayRec rr; rr.ox = ray.org.x; rr.oy = ray.org.y; rr.oz = ray.org.z; rr.spread = rd.spread;
    rr.dx = ray.dir.x; rr.dy = ray.dir.y; rr.dz = ray.dir.z; rr.eta_scale = 1;
    tr.tx = tr.ty = tr.tz = 1; tr.p2 = 0; tr.fx = tr.fy = tr.fz = 0; tr.nv = 2; tr.sample = s;
    RadRec ra; ra.rx = ra.ry = ra.rz = 0; ra.rng = rng.state;
    wp.ray[slot] = rr; wp.thr[slot] = tr; wp.rad[slot] = ra;
    return true;
}

// Counter discipline of an iteration (a kernel may zero any counter that neither it nor a kernel that can still be running
// uses): generate zeroes the hit / miss queues the previous iteration's miss and shade stages consumed, extend zeroes the
// regeneration queue generate has just drained, miss (launched after the shadow + finish stages of the previous iteration
// have been joined, before the shade kernels) zeroes the extension queue extend has just drained and the shadow / finish
// queues of the previous iteration.  No host-side reset launches.
__global__ void __launch_bounds__(128) k_wf_generate(DScene sc, WfPool wp, int n_upper, int s_end) {
    const int n_regen = min(n_upper, wp.counts[Q_REGEN]);
    if (blockIdx.x == 0 && threadIdx.x < Q_COUNT) {
        const int q = threadIdx.x;
        if (q == Q_MISS || q >= Q_KIND0) wp.counts[q] = 0;
        if (q == 0) wp.counts[W_EXT] = 0;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_regen; i += gridDim.x * blockDim.x) {
        const int slot = wp.q_regen[i];
        Ray ray;
        if (generate_slot(sc, wp, slot, s_end, ray)) push_extend(wp, slot, ray);
    }
}

// ---- stage: finish = accumulate paths whose last NEE term had to wait for the shadow stage ----
// `no_regen`: no slot of this render has a further sample (K = samples per pixel), so a finished slot is marked exhausted here
// instead of going through the regeneration queue -- which lets this stage run beside the next iteration's generate / extend.
__global__ void __launch_bounds__(128) k_wf_finish(WfPool wp, int n_upper, int no_regen) {
    int n = min(n_upper, wp.counts[Q_FINISH]);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int slot = wp.q_finish[i];
        RadRec ra = wp.rad[slot];
        if (no_regen) {
            LaneSink out(wp);
            out.terminate(slot, wp.thr[slot].sample, mk3(ra.rx, ra.ry, ra.rz));
            wp.hit[slot].geom = -2;
        } else {
            terminate_path(wp, slot, wp.thr[slot].sample, mk3(ra.rx, ra.ry, ra.rz));
        }
    }
}

__global__ void k_wf_init(WfPool wp, int s_begin) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < wp.capacity; i += (long long)gridDim.x * blockDim.x) {
        int k = (int)(i / wp.map.npix_local);
        ThrRec tr; tr.tx = tr.ty = tr.tz = 0; tr.p2 = 0; tr.fx = tr.fy = tr.fz = 0; tr.nv = 0;
        tr.sample = s_begin + k - wp.map.K;
        wp.thr[i] = tr;
        wp.q_regen[i] = (int)i;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) wp.counts[Q_REGEN] = wp.capacity;
}

} // namespace

namespace {

// One bounce of path_tracing.h's loop body (:96-232) at a fresh vertex: NEE (shadow record),
// BSDF sampling (next extension ray + the f / p2 the next stage needs).  Returns false when the
// path ends here.  KIND is the compile-time material kind of the queue being processed.
template <int KIND, typename Sink>
LJ_D bool bounce_at_vertex(const DScene &sc, Sink &out, int slot, const PathVertex &vertex, d3 dir_view,
                           Pcg32 &rng, d3 throughput, RayRec &rr, ThrRec &tr, bool &pushed_shadow) {
    const unsigned grp = __activemask();     // lanes whose path is alive; see `together` in k_wf_shade
    MatParams mat = load_mat_params(sc, vertex);
    mat.kind = KIND;
    __syncwarp(grp);                         // texture kinds / mip levels differ per lane
    const double eps = sc.eps;
    d2 light_uv; light_uv.x = pcg_real(rng); light_uv.y = pcg_real(rng);
    double light_w = pcg_real(rng);
    double shape_w = pcg_real(rng);
    int light_id = sample_table1d(sc.light_cdf, sc.num_lights, light_w);
    const lj_light &light = sc.lights[light_id];
    PointAndNormal pl = sample_point_on_light(sc, light, vertex.position, light_uv, shape_w);
    __syncwarp(grp);
    {
        // G as if unoccluded; the shadow stage zeroes the term when the ray is blocked
        double G; d3 dir_light; double tfar;
        if (light.kind != LJ_LIGHT_ENVMAP) {
            dir_light = normalize(pl.position - vertex.position);
            tfar = (1 - eps) * distance(pl.position, vertex.position);
            G = lmax(-dot(dir_light, pl.normal), 0.0) / distance_squared(pl.position, vertex.position);
        } else {
            dir_light = -pl.normal; tfar = INFINITY; G = 1;
        }
        double p1 = __ldg(sc.light_pmf + light_id) * pdf_point_on_light(sc, light, pl, vertex.position);
        if (G > 0 && p1 > 0) {
            d3 f = bsdf_eval(mat, vertex, dir_view, dir_light);
            d3 L = light_emission(sc, light, -dir_light, 0.0, pl);
            d3 C1 = G * f * L;
            double p2 = bsdf_pdf(mat, vertex, dir_view, dir_light);
            p2 *= G;
            double w1 = (p1 * p1) / (p1 * p1 + p2 * p2);
            C1 = C1 / p1;
            d3 contrib = throughput * C1 * w1;
            if (contrib.x != 0 || contrib.y != 0 || contrib.z != 0) {
                out.shadow(slot, to_rayf(make_ray(vertex.position, dir_light, eps, tfar)), contrib);
                pushed_shadow = true;
            }
        }
    }
    __syncwarp(grp);
    d2 bsdf_uv; bsdf_uv.x = pcg_real(rng); bsdf_uv.y = pcg_real(rng);
    double bsdf_w = pcg_real(rng);
    BSDFSample bs;
    if (!bsdf_sample(mat, vertex, dir_view, bsdf_uv, bsdf_w, bs)) return false;
    RayDiff rd; rd.radius = 0; rd.spread = rr.spread;
    if (bs.eta == 0) {
        rr.spread = rd_reflect(rd, vertex.mean_curvature, bs.roughness);
    } else {
        rr.spread = rd_refract(rd, vertex.mean_curvature, bs.eta, bs.roughness);
        rr.eta_scale /= (bs.eta * bs.eta);
    }
    d3 f = bsdf_eval(mat, vertex, dir_view, bs.dir_out);
    double p2 = bsdf_pdf(mat, vertex, dir_view, bs.dir_out);
    if (p2 <= 0) return false;                     // path_tracing.h:253-256 (the reference traces the ray first and discards it)
    rr.ox = vertex.position.x; rr.oy = vertex.position.y; rr.oz = vertex.position.z;
    rr.dx = bs.dir_out.x; rr.dy = bs.dir_out.y; rr.dz = bs.dir_out.z;
    tr.fx = f.x; tr.fy = f.y; tr.fz = f.z; tr.p2 = p2;
    return true;
}

// ---- stage: shade, one launch per material kind present in the scene ----
#ifndef LJ_SHADE_MIN_BLOCKS
#define LJ_SHADE_MIN_BLOCKS 4
#endif
#ifndef LJ_SHADE_MIN_BLOCKS_LAMBERTIAN
#define LJ_SHADE_MIN_BLOCKS_LAMBERTIAN LJ_SHADE_MIN_BLOCKS
#endif
constexpr int shade_min_blocks(int kind) { return kind == LJ_MAT_LAMBERTIAN ? LJ_SHADE_MIN_BLOCKS_LAMBERTIAN : LJ_SHADE_MIN_BLOCKS; }
// One path vertex: finish the previous bounce with the new hit (emission + MIS, Russian roulette), then NEE + BSDF
// sampling at the new vertex (path_tracing.h:58-61, :96-322).  Called by the lanes of a warp whose slot has a hit.
template <int KIND, typename Sink>
LJ_D void shade_slot(const DScene &sc, const WfPool &wp, Sink &out, const int slot, const HitRec &hr) {
    // lanes that entered this iteration together; __syncwarp(together) below forces them to
    // reconverge after short divergent blocks (the compiler otherwise runs the long common
    // tail once per diverged group: 1.7x the instructions, seen in profiles/r01c)
    const unsigned together = __activemask();
    RayRec rr = wp.ray[slot];
    ThrRec tr = wp.thr[slot];
    RadRec ra = wp.rad[slot];
    Pcg32 rng; rng.state = ra.rng;
    rng.inc = (slot_stream(wp.map, slot, tr.sample) << 1u) | 1u;
    Ray ray = make_ray(mk3(rr.ox, rr.oy, rr.oz), mk3(rr.dx, rr.dy, rr.dz), 0, 0);
    d3 throughput = mk3(tr.tx, tr.ty, tr.tz), radiance = mk3(ra.rx, ra.ry, ra.rz);
    Hit h; h.t = hr.t; h.u = hr.u; h.v = hr.v; h.geom = hr.geom; h.prim = hr.prim; h.ngx = hr.ngx; h.ngy = hr.ngy; h.ngz = hr.ngz;
    bool primary = tr.nv == 2;
    RayDiff rd; rd.radius = 0; rd.spread = primary ? rr.spread : 0.0;   // bounce rays: intersect(scene, bsdf_ray) default differential
    PathVertex vertex = make_vertex(sc, ray, h, rd);
    bool alive = true;
    bool is_light = sc.shapes[vertex.shape_id].area_light_id >= 0;
    if (primary) {
        if (is_light) radiance += throughput * vertex_emission(sc, vertex, -ray.dir);   // path_tracing.h:58-61
        tr.nv = 3;
    } else {
        // finish the previous bounce with the new vertex (path_tracing.h:242-322)
        d3 f = mk3(tr.fx, tr.fy, tr.fz);
        double G = fabs(dot(ray.dir, vertex.geometry_normal)) / distance_squared(vertex.position, ray.org);
        double p2 = tr.p2 * G;
        if (is_light) {
            d3 L = vertex_emission(sc, vertex, -ray.dir);
            d3 C2 = G * f * L;
            int lid = sc.shapes[vertex.shape_id].area_light_id;
            PointAndNormal lp; lp.position = vertex.position; lp.normal = vertex.geometry_normal;
            double p1 = __ldg(sc.light_pmf + lid) * pdf_point_on_light(sc, sc.lights[lid], lp, ray.org);
            double w2 = (p2 * p2) / (p1 * p1 + p2 * p2);
            C2 = C2 / p2;
            radiance += throughput * C2 * w2;
        }
        double rr_prob = 1;
        if (tr.nv - 1 >= sc.options.rr_depth) {
            rr_prob = lmin(max3((1 / rr.eta_scale) * throughput), 0.95);
            if (pcg_real(rng) > rr_prob) alive = false;
        }
        if (alive) {
            throughput = throughput * (G * f) / (p2 * rr_prob);
            tr.nv += 1;
        }
    }
    __syncwarp(together);
    const int max_depth = sc.options.max_depth;
    if (alive && !(max_depth == -1 || tr.nv <= max_depth + 1)) alive = false;
    bool pushed_shadow = false;
    if (alive) alive = bounce_at_vertex<KIND>(sc, out, slot, vertex, -ray.dir, rng, throughput, rr, tr, pushed_shadow);
    if (!alive) {
        if (!pushed_shadow) { out.terminate(slot, tr.sample, radiance); return; }
        // the path ends here but its last NEE term is still in the shadow queue: park the
        // radiance and let the finish stage (after the shadow stage) do the accumulation
        ra.rx = radiance.x; ra.ry = radiance.y; ra.rz = radiance.z; ra.rng = rng.state;
        wp.rad[slot] = ra;
        out.finish(slot);
        return;
    }
    tr.tx = throughput.x; tr.ty = throughput.y; tr.tz = throughput.z;
    ra.rx = radiance.x; ra.ry = radiance.y; ra.rz = radiance.z; ra.rng = rng.state;
    wp.ray[slot] = rr; wp.thr[slot] = tr; wp.rad[slot] = ra;
    out.extend(slot, make_ray(mk3(rr.ox, rr.oy, rr.oz), mk3(rr.dx, rr.dy, rr.dz), sc.eps, INFINITY));
}

// Two ways to walk the work.  Queue order (SLOTS = false): the slots the extend stage put into this kind's queue.
// Slot order (SLOTS = true, scenes with ONE material kind): every live slot is traced in every iteration, so the
// kernel walks the pool itself and skips slots without a hit (missed: geom -1, exhausted: geom -2) -- the 64-byte
// slot records are then read and written fully coalesced (2 KB contiguous per warp) instead of gathered through the
// queue, and the queue read drops out of the dependent-load chain.  Same per-slot arithmetic either way.
template <int KIND, bool SLOTS>
__global__ void __launch_bounds__(128, shade_min_blocks(KIND)) k_wf_shade(DScene sc, WfPool wp, int n_upper) {
    const int n = SLOTS ? wp.capacity : min(n_upper, wp.counts[Q_KIND0 + KIND]);
    const int *queue = wp.q_kind[KIND];
    const int lane = threadIdx.x & 31;
    for (int base = blockIdx.x * blockDim.x + threadIdx.x - lane; base < n; base += gridDim.x * blockDim.x) {
        const int i = base + lane;
        int slot = -1; HitRec hr; hr.geom = -1;
        if (i < n) { slot = SLOTS ? i : queue[i]; hr = wp.hit[slot]; }
        if (hr.geom >= 0) { QueueSink out(wp); shade_slot<KIND>(sc, wp, out, slot, hr); }
        __syncwarp();       // the warp walks the range in lockstep: lanes without work wait here
    }
}

// ---- stage: miss (path_tracing.h:17-28 and :283-307) ----
template <typename Sink>
LJ_D void miss_slot(const DScene &sc, const WfPool &wp, Sink &out, int slot) {
    RayRec rr = wp.ray[slot];
    ThrRec tr = wp.thr[slot];
    RadRec ra = wp.rad[slot];
    d3 radiance = mk3(ra.rx, ra.ry, ra.rz);
    if (sc.envmap_light_id != -1) {
        const lj_light &env = sc.lights[sc.envmap_light_id];
        d3 dir = mk3(rr.dx, rr.dy, rr.dz);
        PointAndNormal dummy; dummy.position = mk3(0.0); dummy.normal = mk3(0.0);
        d3 L = light_emission(sc, env, -dir, rr.spread, dummy);
        if (tr.nv == 2) {
            radiance = L;
        } else {
            d3 throughput = mk3(tr.tx, tr.ty, tr.tz), f = mk3(tr.fx, tr.fy, tr.fz);
            double G = 1;
            double p2 = tr.p2 * G;
            d3 C2 = G * f * L;
            PointAndNormal lp; lp.position = mk3(0.0); lp.normal = -dir;
            double p1 = __ldg(sc.light_pmf + sc.envmap_light_id) * pdf_point_on_light(sc, env, lp, mk3(rr.ox, rr.oy, rr.oz));
            double w2 = (p2 * p2) / (p1 * p1 + p2 * p2);
            C2 = C2 / p2;
            radiance += throughput * C2 * w2;
        }
    }
    out.terminate(slot, tr.sample, radiance);
}
__global__ void __launch_bounds__(128) k_wf_miss(DScene sc, WfPool wp, int n_upper) {
    int n = min(n_upper, wp.counts[Q_MISS]);
    if (blockIdx.x == 0 && threadIdx.x == 0) {      // drained by the extend stage / by the previous iteration's shadow and finish stages; the shade kernels refill them
        wp.counts[Q_EXT] = 0; wp.counts[Q_SHADOW] = 0; wp.counts[Q_FINISH] = 0; wp.counts[W_SHADOW] = 0;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        QueueSink out(wp);
        miss_slot(sc, wp, out, wp.q_miss[i]);
    }
}

// ---- tail: every remaining path to its end, one lane each ----
// Once few paths are left, an iteration costs its launch and latency floor (six launches, ~0.17 ms on sponza) however
// few rays it carries, and a render's last 30-odd iterations carry almost none: an 8-spp sponza render spent 5.5 of
// 25.7 ms there, which is also what limits fixed-work multi-GPU scaling at low sample counts.  This kernel takes over
// when the live count falls below a threshold: one lane per remaining slot runs trace -> miss / shade -> shadow ray ->
// accumulate / regenerate in a loop until the slot has no samples left, with the SAME per-slot functions and slot
// records as the stage kernels (their outputs go to the lane through LaneSink instead of the queues) and the same
// traversal code run to completion per ray (trace_closest / trace_any), so results stay bit-identical.
template <int KIND>
__device__ __noinline__ void tail_shade(const DScene &sc, const WfPool &wp, LaneSink &out, int slot, const HitRec &hr) {
    shade_slot<KIND>(sc, wp, out, slot, hr);
}
enum { T_PRIMARY = 0, T_EXTENSION = 1, T_SHADOW = 2 };     // ray counters of the tail kernel (u64, after the traversal counters)
__global__ void __launch_bounds__(128, 2) k_wf_tail(DScene sc, WfPool wp, int s_end, unsigned long long *ray_counters) {
    const int n_ext = wp.counts[Q_EXT], n_regen = wp.counts[Q_REGEN];
    const int n = n_ext + n_regen;
    unsigned c_primary = 0, c_extension = 0, c_shadow = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int slot; bool have_ray, primary = false; RayF ray;
        if (i < n_ext) { slot = wp.ext_slot[i]; ray = wp.ext_ray[i].ray; have_ray = true; }
        else { slot = wp.q_regen[i - n_ext]; have_ray = false; }
        while (true) {
            if (!have_ray) {
                Ray r;
                if (!generate_slot(sc, wp, slot, s_end, r)) break;
                ray = to_rayf(r); primary = true;
            }
            Hit h;
            const bool hit = trace_closest(sc, ray, &h);
            if (primary) c_primary++; else c_extension++;
            primary = false;
            LaneSink out(wp);
            if (!hit) { miss_slot(sc, wp, out, slot); have_ray = false; continue; }
            HitRec hr; hr.t = h.t; hr.u = h.u; hr.v = h.v; hr.geom = h.geom; hr.prim = h.prim; hr.ngx = h.ngx; hr.ngy = h.ngy; hr.ngz = h.ngz;
            switch (sc.materials[sc.shapes[h.geom].material_id].kind) {
                case 0: tail_shade<0>(sc, wp, out, slot, hr); break;
                case 1: tail_shade<1>(sc, wp, out, slot, hr); break;
                case 2: tail_shade<2>(sc, wp, out, slot, hr); break;
                case 3: tail_shade<3>(sc, wp, out, slot, hr); break;
                case 4: tail_shade<4>(sc, wp, out, slot, hr); break;
                case 5: tail_shade<5>(sc, wp, out, slot, hr); break;
                case 6: tail_shade<6>(sc, wp, out, slot, hr); break;
                case 7: tail_shade<7>(sc, wp, out, slot, hr); break;
                default: tail_shade<8>(sc, wp, out, slot, hr); break;
            }
            if (out.has_shadow) {                               // k_wf_shadow: the NEE term counts unless the ray is blocked
                c_shadow++;
                if (!trace_any(sc, out.shadow_ray)) { RadRec *ra = &wp.rad[slot]; ra->rx += out.cx; ra->ry += out.cy; ra->rz += out.cz; }
            }
            if (out.what == 2) {                                // k_wf_finish
                const RadRec ra = wp.rad[slot];
                out.terminate(slot, wp.thr[slot].sample, mk3(ra.rx, ra.ry, ra.rz));
                have_ray = false;
            } else if (out.what == 1) { ray = out.next; have_ray = true; }
            else have_ray = false;
        }
    }
    if (c_primary) atomicAdd(&ray_counters[T_PRIMARY], (unsigned long long)c_primary);
    if (c_extension) atomicAdd(&ray_counters[T_EXTENSION], (unsigned long long)c_extension);
    if (c_shadow) atomicAdd(&ray_counters[T_SHADOW], (unsigned long long)c_shadow);
}

template <int KIND>
void launch_shade(const DScene &d, const WfPool &wp, int n_upper, int grid_cap, bool slot_order) {
    int grid = ((slot_order ? wp.capacity : n_upper) + 127) / 128;
    if (grid > grid_cap) grid = grid_cap;
    if (grid < 1) grid = 1;
    if (slot_order) k_wf_shade<KIND, true><<<grid, 128>>>(d, wp, n_upper);
    else k_wf_shade<KIND, false><<<grid, 128>>>(d, wp, n_upper);
}

} // namespace

static int ensure_pool(lj_scene *s, const SlotMap &map, WfPool &wp) {
    const size_t cap = (size_t)map.K * map.npix_local;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    size_t o_ray = take(cap * sizeof(RayRec)), o_thr = take(cap * sizeof(ThrRec)), o_rad = take(cap * sizeof(RadRec));
    size_t o_hit = take(cap * sizeof(HitRec)), o_acc = take((size_t)map.K0 * map.npix_local * 3 * sizeof(double));
    size_t o_er = take(cap * sizeof(ExtRec)), o_es = take(cap * sizeof(int)), o_sh = take(cap * sizeof(ShadowRec));
    size_t o_qr = take(cap * sizeof(int)), o_qm = take(cap * sizeof(int)), o_qf = take(cap * sizeof(int));
    size_t o_qk[LJ_MAT_KIND_COUNT];
    for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) o_qk[k] = take(cap * sizeof(int));
    size_t o_cnt = take(64 * sizeof(int));
    void *arena = nullptr;
    if (int rc = lj_ctx_arena(s->ctx, off, &arena)) return rc;
    char *b = (char *)arena;
    wp.capacity = (int)cap; wp.map = map;
    wp.ray = (RayRec *)(b + o_ray); wp.thr = (ThrRec *)(b + o_thr); wp.rad = (RadRec *)(b + o_rad); wp.hit = (HitRec *)(b + o_hit);
    wp.accum = (double *)(b + o_acc); wp.ext_ray = (ExtRec *)(b + o_er); wp.ext_slot = (int *)(b + o_es);
    wp.shadow = (ShadowRec *)(b + o_sh); wp.q_regen = (int *)(b + o_qr); wp.q_miss = (int *)(b + o_qm); wp.q_finish = (int *)(b + o_qf);
    for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) wp.q_kind[k] = (int *)(b + o_qk[k]);
    wp.counts = (int *)(b + o_cnt);
    return LJ_OK;
}

int lj_render_wavefront(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    const DScene &d = s->d;
    if (!d.bvh8) return lj_render_megakernel(s, rl, dev_sum, st);   // LJ_BVH=2 validation mode: binary tree only
    const long long npix = (long long)rl.width * rl.height;
    const int nsamp = rl.s_end - rl.s_begin;
    // pool sizing: K sample lanes per owned pixel (a power of two, see SlotMap).  Default 32 Mi slots (~15 GB of the context's
    // arena): with shade in slot order and the tail kernel, fewer and larger iterations win on every scene measured (sponza
    // 64 spp 166.5 / 161.2 / 156.4 / 152.1 ms at 4 / 8 / 16 / 32 Mi; matpreview 156.4 / 132.5 / 120.5 / 112.8; veach_mi 23.7 ms at
    // 4 Mi, 19.8 at 32 Mi) except with the full Disney BSDF, whose shade kernel spills 1.4 kB per thread and wants its slot
    // records near L2 (disney_bsdf 131-133 ms at 2-4 Mi, 154 at 16 Mi): 4 Mi there.  profiles/r02_experiments.md.
    const long long dflt_slots = s->kind_present[LJ_MAT_DISNEY_BSDF] ? (4LL << 20) : (32LL << 20);
    const long long target = rl.p.paths_in_flight > 0 ? rl.p.paths_in_flight : dflt_slots;
    const SlotMap map = lj_make_slot_map(rl.width, rl.height, rl.s_begin, rl.s_end, rl.p.tile_rank, rl.p.tile_world, target);
    if ((long long)map.npix_local * map.K > 0x7fffffffLL / 4) { lj_set_error("wavefront: pool too large"); return LJ_EINVAL; }
    WfPool wp;
    int rc = ensure_pool(s, map, wp);
    if (rc) return rc;
    const int capacity = wp.capacity;
    const bool *kind_present = s->kind_present;

    const int sm = s->ctx->sm_count;
    const int shade_grid_cap = sm * 16;
    auto env_int = [](const char *name, int dflt) { const char *v = getenv(name); return v ? atoi(v) : dflt; };
    TraceTuning tune_ext, tune_sh;     // tuning knobs (defaults measured on B200, see profiles/)
    tune_ext.refill_below = env_int("LJ_REFILL_BELOW", 24); tune_ext.prim_thresh = env_int("LJ_PRIM_THRESH", 12);
    tune_sh.refill_below = env_int("LJ_REFILL_BELOW_SHADOW", 24); tune_sh.prim_thresh = env_int("LJ_PRIM_THRESH_SHADOW", 6);
    const int trace_grid_cap = sm * env_int("LJ_TRACE_GRID_MULT", 16);
    const bool trace_iters = getenv("LJ_WF_TRACE") != nullptr;
    const bool stage_timing = (rl.p.flags & 2) != 0 || trace_iters;
    const bool counting = (rl.p.flags & 1) != 0 && d.bvh8 != nullptr;
    unsigned long long *trav = (unsigned long long *)(wp.counts + 32);   // 4 x u64 after the queue counters

    // Queue sizes live on the device and every kernel clamps its range to them, so the host does not need them to
    // launch the next iteration: it sizes grids from the last sizes it has seen (or the pool capacity) and reads the
    // counters back through a ring of pinned buffers, up to `depth` iterations behind.  Iterations launched after the
    // one that emptied the pool find empty queues and do nothing (LJ_WF_PIPELINE=1 restores one read-back per iteration).
    const int depth = trace_iters ? 1 : std::max(1, std::min(8, env_int("LJ_WF_PIPELINE", 4)));
    static_assert(Q_COUNT <= LJ_RING_INTS, "ring slot too small for the queue counters");
    if ((rc = lj_ctx_ring(s->ctx)) != LJ_OK) return rc;
    int *h_ring = s->ctx->h_ring;
    cudaEvent_t *done = s->ctx->ring_done;
    cudaEvent_t (*ev)[5] = s->ctx->ring_ev;
    auto cleanup = [&]() {};
    // Shadow + finish of iteration i beside generate + extend of iteration i + 1 (second stream): possible when no slot can
    // regenerate (K sample lanes = samples per pixel), because then nothing the finish stage produces is needed before the
    // next shade stage -- a finished slot is marked exhausted on the spot.  The extend kernel leaves about a third of the
    // issue slots idle and the shadow kernel's warps fit beside its 40 per SM.  LJ_WF_OVERLAP=0 turns it off.
    const bool no_regen = map.K >= nsamp;
    // (not with per-stage timing, lj_render_params.flags bit1: a stage's CUDA-event interval is only that stage's time when the
    //  stages run one after the other -- bench.py's roofline divides by the extend kernel's own duration)
    const bool overlap = no_regen && env_int("LJ_WF_OVERLAP", 1) != 0 && !stage_timing;
    cudaStream_t sb = s->ctx->stream_b;
    cudaEvent_t *ring_a = s->ctx->ring_a, *ring_b = s->ctx->ring_b;

    cudaMemsetAsync(wp.counts, 0, 64 * sizeof(int));
    cudaMemsetAsync(wp.accum, 0, (size_t)map.K0 * map.npix_local * 3 * sizeof(double));
    if (map.tile_world > 0) cudaMemsetAsync(dev_sum, 0, sizeof(double) * 3 * (size_t)npix);
    k_wf_init<<<sm * 8, 256>>>(wp, rl.s_begin);
    if (cudaGetLastError() != cudaSuccess) { cleanup(); lj_set_error("wavefront: init launch failed"); return LJ_ECUDA; }
    int n_regen = capacity, n_ext = 0;      // queue sizes after the last retired iteration
    int n_kinds = 0;
    for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) n_kinds += kind_present[k] ? 1 : 0;
    const bool slots_ok = n_kinds == 1 && !getenv("LJ_SHADE_QUEUE");
    wp.single_kind = -1;
    if (n_kinds == 1 && !getenv("LJ_NO_SINGLE_KIND")) for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) if (kind_present[k]) wp.single_kind = k;
    // slot order while at least this share of the pool's slots had a hit in the last iteration the host has seen (below it
    // the idle lanes of missed / exhausted slots cost more than the coalesced records save: open scenes such as cbox, renders
    // without regeneration, and the tail of every render)
    const int slots_min_live_pct = env_int("LJ_SHADE_SLOTS_MIN_LIVE_PCT", 95);
    int n_hits = capacity;
    unsigned long long launches = 2, rays_primary = 0, rays_ext = 0, rays_shadow = 0;
    unsigned long long launched = 0, retired = 0;
    auto grid_for = [](int n, int cap) { int g = (n + 127) / 128; return g < 1 ? 1 : (g > cap ? cap : g); };
    bool finished = false;
    int loop_rc = LJ_OK;
    // retire what has completed (reads the counters an iteration left in the ring); blocks only when the host is as far
    // ahead as it may go -- or, with `drain`, until nothing is in flight
    auto retire = [&](bool drain) {
        while (retired < launched) {
            const int rs = (int)(retired % depth);

            // how far the host runs ahead: one speculative iteration while iterations are large (it hides the read-back
            // and costs at most one empty iteration at the end), the full ring once they are small and launch-bound
            const unsigned long long ahead = (n_ext + n_regen) >= (1 << 19) ? (depth < 2 ? depth : 2) : depth;
            if (!drain && launched - retired < ahead && cudaEventQuery(done[rs]) == cudaErrorNotReady) break;
            const cudaError_t ce = cudaEventSynchronize(done[rs]);
            if (ce != cudaSuccess) { lj_set_error("wavefront iteration failed: %s", cudaGetErrorString(ce)); loop_rc = LJ_ECUDA; finished = true; break; }
            const int *h_counts = h_ring + rs * Q_COUNT;
            int traced = h_counts[Q_MISS];
            for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) traced += h_counts[Q_KIND0 + k];
            // rays traced this iteration = hits + misses; all but the n_ext carried over are primaries
            const int prim_now = traced - n_ext > 0 ? traced - n_ext : 0;
            rays_primary += prim_now; rays_ext += traced - prim_now;
            rays_shadow += h_counts[Q_SHADOW];
            float ms_g = 0, ms_e = 0, ms_s = 0, ms_h = 0;
            if (stage_timing) {
                cudaEventElapsedTime(&ms_g, ev[rs][0], ev[rs][1]); cudaEventElapsedTime(&ms_e, ev[rs][1], ev[rs][2]);
                cudaEventElapsedTime(&ms_s, ev[rs][2], ev[rs][3]); cudaEventElapsedTime(&ms_h, ev[rs][3], ev[rs][4]);
                st->generate_ms += ms_g; st->extend_ms += ms_e; st->shade_ms += ms_s; st->shadow_ms += ms_h;
            }
            if (trace_iters)
                fprintf(stderr, "[wf] iter %llu: traced %d (regen %d) shadow %d -> next ext %d regen %d | extend %.3f shade %.3f shadow %.3f ms\n",
                        (unsigned long long)st->iterations, traced, n_regen, h_counts[Q_SHADOW], h_counts[Q_EXT], h_counts[Q_REGEN], ms_e, ms_s, ms_h);
            n_regen = h_counts[Q_REGEN];
            n_ext = h_counts[Q_EXT];
            n_hits = traced - h_counts[Q_MISS];
            st->iterations++;
            retired++;
            // (slots whose samples are used up sit in the regeneration queue once more: that iteration generates
            //  nothing, extend zeroes the queue, and the counters come back empty)
            //  Without regeneration the slots in the regeneration queue are finished ones waiting to be marked exhausted.)
            if (n_ext == 0 && (no_regen || n_regen == 0)) { finished = true; break; }
        }
    };
    // Tail (k_wf_tail): once at most this many slots are live, one launch runs them all to completion
    const int tail_threshold = counting ? 0 : env_int("LJ_WF_TAIL", 32768);     // (the counting pass keeps every ray in the counting kernels)
    unsigned long long *tail_rays = (unsigned long long *)(wp.counts + 40);      // 3 x u64, zeroed with the counters
    while (!finished) {
        if (tail_threshold > 0 && retired > 0 && n_ext + (no_regen ? 0 : n_regen) <= tail_threshold) {     // (sizes are real once an iteration has been retired)
            retire(true);                               // exact sizes, nothing in flight
            if (overlap && launched > 0) cudaStreamWaitEvent(0, ring_b[(launched - 1) % depth], 0);
            if (finished) break;
            const int live_now = n_ext + n_regen;          // (regeneration entries included: the tail kernel marks them exhausted)
            k_wf_tail<<<grid_for(live_now, shade_grid_cap), 128>>>(d, wp, rl.s_end, tail_rays);
            unsigned long long h_tail[3] = {0, 0, 0};
            const cudaError_t ce = cudaMemcpy(h_tail, tail_rays, sizeof h_tail, cudaMemcpyDeviceToHost);
            if (ce != cudaSuccess) { lj_set_error("wavefront tail failed: %s", cudaGetErrorString(ce)); loop_rc = LJ_ECUDA; break; }
            rays_primary += h_tail[T_PRIMARY]; rays_ext += h_tail[T_EXTENSION]; rays_shadow += h_tail[T_SHADOW];
            if (trace_iters) fprintf(stderr, "[wf] tail: %d live slots -> %llu + %llu + %llu rays\n", live_now, h_tail[0], h_tail[1], h_tail[2]);
            launches++; st->iterations++;
            finished = true;
            break;
        }
        // Slots only ever leave the pool (exhausted ones at the generate stage), so the live count of the last retired
        // iteration bounds every queue of every later iteration: grids are sized from it even while iterations are in
        // flight (all stage kernels are grid-stride / work-counter loops clamped to the device counts, so a bound that
        // is a few iterations stale costs nothing but a few idle blocks).
        const bool exact = launched == retired;          // nothing in flight: the sizes are the true ones
        const int live = n_ext + n_regen > capacity ? capacity : n_ext + n_regen;
        const int regen_upper = (exact && !overlap) ? n_regen : live;
        const int upper = live;
        const int ring_slot = (int)(launched % depth);
        cudaEvent_t *e = ev[ring_slot];
        if (stage_timing) cudaEventRecord(e[0]);
        k_wf_generate<<<grid_for(regen_upper, shade_grid_cap), 128>>>(d, wp, regen_upper, rl.s_end);
        if (stage_timing) cudaEventRecord(e[1]);
        lj_wf_launch_extend(d, wp, upper, trav, tune_ext, counting, grid_for(upper, trace_grid_cap));
        if (stage_timing) cudaEventRecord(e[2]);
        // slot order while most of the pool is live (the last sizes the host has seen), queue order in the tail
        const bool slot_order = slots_ok && (long long)n_hits * 100 >= (long long)capacity * slots_min_live_pct;
        // the shadow + finish stages of the previous iteration (on the second stream when overlapped) are joined here: miss
        // resets their queues, shade reads the radiance they updated
        if (overlap && launched > 0) cudaStreamWaitEvent(0, ring_b[(launched - 1) % depth], 0);
        k_wf_miss<<<grid_for(upper, shade_grid_cap), 128>>>(d, wp, upper);
        if (kind_present[0]) launch_shade<0>(d, wp, upper, shade_grid_cap, slot_order);
        if (kind_present[1]) launch_shade<1>(d, wp, upper, shade_grid_cap, slot_order);
        if (kind_present[2]) launch_shade<2>(d, wp, upper, shade_grid_cap, slot_order);
        if (kind_present[3]) launch_shade<3>(d, wp, upper, shade_grid_cap, slot_order);
        if (kind_present[4]) launch_shade<4>(d, wp, upper, shade_grid_cap, slot_order);
        if (kind_present[5]) launch_shade<5>(d, wp, upper, shade_grid_cap, slot_order);
        if (kind_present[6]) launch_shade<6>(d, wp, upper, shade_grid_cap, slot_order);
        if (kind_present[7]) launch_shade<7>(d, wp, upper, shade_grid_cap, slot_order);
        if (kind_present[8]) launch_shade<8>(d, wp, upper, shade_grid_cap, slot_order);
        if (overlap) {
            // every counter the host needs is final once the shade stage is done (the finish stage pushes nothing here)
            cudaMemcpyAsync(h_ring + ring_slot * Q_COUNT, wp.counts, sizeof(int) * Q_COUNT, cudaMemcpyDeviceToHost);
            cudaEventRecord(done[ring_slot]);
            cudaEventRecord(ring_a[ring_slot]);
            cudaStreamWaitEvent(sb, ring_a[ring_slot], 0);
            if (stage_timing) cudaEventRecord(e[3], sb);
            lj_wf_launch_shadow(d, wp, upper, trav, tune_sh, counting, grid_for(upper, trace_grid_cap), sb);
            k_wf_finish<<<grid_for(upper, shade_grid_cap), 128, 0, sb>>>(wp, upper, 1);
            if (stage_timing) cudaEventRecord(e[4], sb);
            cudaEventRecord(ring_b[ring_slot], sb);
        } else {
            if (stage_timing) cudaEventRecord(e[3]);
            lj_wf_launch_shadow(d, wp, upper, trav, tune_sh, counting, grid_for(upper, trace_grid_cap), 0);
            k_wf_finish<<<grid_for(upper, shade_grid_cap), 128>>>(wp, upper, 0);
            if (stage_timing) cudaEventRecord(e[4]);
            cudaMemcpyAsync(h_ring + ring_slot * Q_COUNT, wp.counts, sizeof(int) * Q_COUNT, cudaMemcpyDeviceToHost);
            cudaEventRecord(done[ring_slot]);
        }
        launches += 5;
        for (int k = 0; k < LJ_MAT_KIND_COUNT; k++) launches += kind_present[k] ? 1 : 0;
        launched++;
        retire(false);
    }
    cudaStreamSynchronize(0);           // drains the (empty) iterations still in flight
    if (overlap) cudaStreamSynchronize(sb);
    cleanup();
    if (loop_rc != LJ_OK) return loop_rc;
    k_slot_sum<<<(unsigned)(((long long)map.npix_local * 3 + 255) / 256), 256>>>(map, wp.accum, dev_sum);
    LJ_CUDA(cudaGetLastError());
    if (counting) {
        unsigned long long h_trav[4];
        LJ_CUDA(cudaMemcpy(h_trav, trav, sizeof h_trav, cudaMemcpyDeviceToHost));
        st->nodes_visited = h_trav[0]; st->prims_tested = h_trav[1];          // closest-hit rays
        st->shadow_nodes_visited = h_trav[2]; st->shadow_prims_tested = h_trav[3];
    }
    launches++;
    st->kernel_launches += launches;
    st->rays_primary += rays_primary; st->rays_extension += rays_ext; st->rays_shadow += rays_shadow;
    st->samples = (uint64_t)npix * (uint64_t)nsamp;
    return LJ_OK;
}
