// Volumetric path tracing: media, grid volumes, phase functions and the integrators of
// /root/reference/src/vol_path_tracing.h, restated for the device with the same RNG draw
// order, the same early exits and the same operation order (f64, no FMA contraction).
//
//   vol_path_tracing_1        vol_path_tracing.h:6-41    absorption only
//   vol_path_tracing_2        vol_path_tracing.h:46-147  single scattering, monochrome distance sampling
//   update_medium             vol_path_tracing.h:149-163
//   next_event_estimation_final  vol_path_tracing.h:299-495  ratio tracking through media + index-matched surfaces
//   vol_path_tracing (final)  vol_path_tracing.h:503-869  delta tracking, spectral MIS, NEE/phase MIS, surfaces, RR
// Versions 3, 4, 5 and 6 of the reference all forward to the final integrator
// (vol_path_tracing.h:880, :1052, :1297; render.cpp:111-124).
//
// Quirks kept on purpose (they change pixel values): ray.org is NOT advanced when the
// majorant of the sampled channel is zero or the null-collision cap is hit; the NEE sigma
// lookups of the final version use positions along the shadow ray; `current_path_throughput
// *= f / pdf` has no pdf <= 0 guard (vol_path_tracing.h:843-846); a missing hit leaves the
// reference's PathVertex uninitialised -- every use of it is dead in those branches, here
// the vertex is simply left untouched.
#pragma once
#include "lj_path.cuh"

LJ_D d3 exp3(d3 a) { return mk3(exp(a.x), exp(a.y), exp(a.z)); }   // spectrum.h:28-30
LJ_D d3 load3f(const float *p, size_t i) { return mk3((double)__ldg(p + 3 * i), (double)__ldg(p + 3 * i + 1), (double)__ldg(p + 3 * i + 2)); }
LJ_D d3 load3f4(const float4 *g4, size_t i) { const float4 t = __ldg(g4 + 2 * i); return mk3((double)t.x, (double)t.y, (double)t.z); }   // paired records: stride 2 float4
// volume.h:68-80: the trilinear blend, terms added in the reference's order
LJ_D d3 volume_blend(double scale, double dx, double dy, double dz, d3 v000, d3 v001, d3 v010, d3 v011, d3 v100, d3 v101, d3 v110, d3 v111) {
    return scale * (v000 * ((1 - dx) * (1 - dy) * (1 - dz)) +
                    v001 * (dx * (1 - dy) * (1 - dz)) +
                    v010 * ((1 - dx) * dy * (1 - dz)) +
                    v011 * (dx * dy * (1 - dz)) +
                    v100 * ((1 - dx) * (1 - dy) * dz) +
                    v101 * (dx * (1 - dy) * dz) +
                    v110 * ((1 - dx) * dy * dz) +
                    v111 * (dx * dy * dz));
}

// the blend above for a one-channel f32 grid
LJ_D double volume_blend_grey(const float *gf, double scale, double dx, double dy, double dz, size_t i000, size_t i001, size_t i010, size_t i011,
                              size_t i100, size_t i101, size_t i110, size_t i111) {
    const double a000 = (double)__ldg(gf + i000), a001 = (double)__ldg(gf + i001), a010 = (double)__ldg(gf + i010), a011 = (double)__ldg(gf + i011);
    const double a100 = (double)__ldg(gf + i100), a101 = (double)__ldg(gf + i101), a110 = (double)__ldg(gf + i110), a111 = (double)__ldg(gf + i111);
    return scale * (a000 * ((1 - dx) * (1 - dy) * (1 - dz)) +
                    a001 * (dx * (1 - dy) * (1 - dz)) +
                    a010 * ((1 - dx) * dy * (1 - dz)) +
                    a011 * (dx * dy * (1 - dz)) +
                    a100 * ((1 - dx) * (1 - dy) * dz) +
                    a101 * (dx * (1 - dy) * dz) +
                    a110 * ((1 - dx) * dy * dz) +
                    a111 * (dx * dy * dz));
}

// ---- volume.h:45-81: trilinear lookup in double, zero outside the grid's box ----
// the cell of a point: corner indices and fractional offsets (volume.h:46-66); false outside the box
struct GridCell { size_t i000, i001, i010, i011, i100, i101, i110, i111; double dx, dy, dz; };
LJ_D bool grid_cell(const lj_volume &v, d3 p, GridCell &c) {
    d3 pmin = mk3(v.p_min[0], v.p_min[1], v.p_min[2]), pmax = mk3(v.p_max[0], v.p_max[1], v.p_max[2]);
    d3 pn = (p - pmin) / (pmax - pmin);
    if (pn.x < 0 || pn.x > 1 || pn.y < 0 || pn.y > 1 || pn.z < 0 || pn.z > 1) return false;
    const int rx = v.resolution[0], ry = v.resolution[1], rz = v.resolution[2];
    pn.x *= (double)(rx - 1); pn.y *= (double)(ry - 1); pn.z *= (double)(rz - 1);
    int x0 = clampi((int)pn.x, 0, rx - 1), y0 = clampi((int)pn.y, 0, ry - 1), z0 = clampi((int)pn.z, 0, rz - 1);
    int x1 = clampi(x0 + 1, 0, rx - 1), y1 = clampi(y0 + 1, 0, ry - 1), z1 = clampi(z0 + 1, 0, rz - 1);
    c.dx = pn.x - x0; c.dy = pn.y - y0; c.dz = pn.z - z0;
    c.i000 = (size_t)(z0 * ry + y0) * rx + x0; c.i001 = (size_t)(z0 * ry + y0) * rx + x1;
    c.i010 = (size_t)(z0 * ry + y1) * rx + x0; c.i011 = (size_t)(z0 * ry + y1) * rx + x1;
    c.i100 = (size_t)(z1 * ry + y0) * rx + x0; c.i101 = (size_t)(z1 * ry + y0) * rx + x1;
    c.i110 = (size_t)(z1 * ry + y1) * rx + x0; c.i111 = (size_t)(z1 * ry + y1) * rx + x1;
    return true;
}
// A volume whose three channels are equal everywhere (a constant with equal components, or a grid stored as one f32
// channel): one channel of volume_lookup's result, which is all of them.
LJ_D bool volume_is_grey(const lj_volume &v) {
    return v.kind == LJ_VOLUME_CONSTANT ? (v.value[0] == v.value[1] && v.value[0] == v.value[2]) : v._pad == 1;
}
LJ_D double volume_lookup_grey(const DScene &sc, const lj_volume &v, d3 p) {
    if (v.kind == LJ_VOLUME_CONSTANT) return v.value[0];
    GridCell c;
    if (!grid_cell(v, p, c)) return 0.0;
    return volume_blend_grey(sc.grid_f32 + v.data_offset, v.scale, c.dx, c.dy, c.dz, c.i000, c.i001, c.i010, c.i011, c.i100, c.i101, c.i110, c.i111);
}
LJ_D d3 volume_lookup(const DScene &sc, const lj_volume &v, d3 p) {
    if (v.kind == LJ_VOLUME_CONSTANT) return mk3(v.value[0], v.value[1], v.value[2]);
    GridCell c;
    if (!grid_cell(v, p, c)) return mk3(0.0);
    const double dx = c.dx, dy = c.dy, dz = c.dz;
    const size_t i000 = c.i000, i001 = c.i001, i010 = c.i010, i011 = c.i011, i100 = c.i100, i101 = c.i101, i110 = c.i110, i111 = c.i111;
    if (v._pad == 1) {
        // one f32 channel (the three channels of the Spectrum grid are equal): the same expression once; every
        // channel of the reference's result is this value
        return mk3(volume_blend_grey(sc.grid_f32 + v.data_offset, v.scale, dx, dy, dz, i000, i001, i010, i011, i100, i101, i110, i111));
    }
    d3 v000, v001, v010, v011, v100, v101, v110, v111;
    if (v._pad == 8) {
        // one half of a paired record (8 floats per voxel: density rgb, 0, albedo rgb, 0; _pad2 picks the half)
        const float4 *g4 = (const float4 *)(sc.grid_f32 + v.data_offset) + v._pad2;
        v000 = load3f4(g4, i000); v001 = load3f4(g4, i001); v010 = load3f4(g4, i010); v011 = load3f4(g4, i011);
        v100 = load3f4(g4, i100); v101 = load3f4(g4, i101); v110 = load3f4(g4, i110); v111 = load3f4(g4, i111);
    } else if (v._pad == 3) {
        const float *gf = sc.grid_f32 + v.data_offset;
        v000 = load3f(gf, i000); v001 = load3f(gf, i001); v010 = load3f(gf, i010); v011 = load3f(gf, i011);
        v100 = load3f(gf, i100); v101 = load3f(gf, i101); v110 = load3f(gf, i110); v111 = load3f(gf, i111);
    } else {
        const double *g = sc.grid_data + v.data_offset;
        v000 = load3(g, i000); v001 = load3(g, i001); v010 = load3(g, i010); v011 = load3(g, i011);
        v100 = load3(g, i100); v101 = load3(g, i101); v110 = load3(g, i110); v111 = load3(g, i111);
    }
    return volume_blend(v.scale, dx, dy, dz, v000, v001, v010, v011, v100, v101, v110, v111);
}
// volume.h:122-141 (pbrt slab test from t=0 to ray.tfar; tnear is ignored as in the reference)
LJ_D bool volume_intersect(const lj_volume &v, const Ray &ray) {
    if (v.kind == LJ_VOLUME_CONSTANT) return true;
    double t0 = 0, t1 = ray.tfar;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        double o = comp(ray.org, i), d = comp(ray.dir, i);
        double tn = (v.p_min[i] - o) / d, tf = (v.p_max[i] - o) / d;
        if (tn > tf) { double s = tn; tn = tf; tf = s; }
        t0 = tn > t0 ? tn : t0;
        t1 = tf < t1 ? tf : t1;
        if (t0 > t1) return false;
    }
    return true;
}

// ---- medium.cpp:27-37, media/homogeneous.inl, media/heterogeneous.inl ----
LJ_D d3 medium_majorant(const DScene &sc, const lj_medium &m, const Ray &ray) {
    if (m.kind == LJ_MEDIUM_HOMOGENEOUS)
        return mk3(m.sigma_a[0], m.sigma_a[1], m.sigma_a[2]) + mk3(m.sigma_s[0], m.sigma_s[1], m.sigma_s[2]);
    if (!volume_intersect(m.density, ray)) return mk3(0.0);
    if (m.density.kind == LJ_VOLUME_CONSTANT) return mk3(m.density.value[0], m.density.value[1], m.density.value[2]);
    return m.density.scale * mk3(m.density.max_data[0], m.density.max_data[1], m.density.max_data[2]);   // volume.h:97-99
}
LJ_D d3 medium_sigma_s(const DScene &sc, const lj_medium &m, d3 p) {
    if (m.kind == LJ_MEDIUM_HOMOGENEOUS) return mk3(m.sigma_s[0], m.sigma_s[1], m.sigma_s[2]);
    d3 density = volume_lookup(sc, m.density, p);
    d3 albedo = volume_lookup(sc, m.albedo, p);
    return density * albedo;
}
LJ_D d3 medium_sigma_a(const DScene &sc, const lj_medium &m, d3 p) {
    if (m.kind == LJ_MEDIUM_HOMOGENEOUS) return mk3(m.sigma_a[0], m.sigma_a[1], m.sigma_a[2]);
    d3 density = volume_lookup(sc, m.density, p);
    d3 albedo = volume_lookup(sc, m.albedo, p);
    return density * (1.0 - albedo);
}

// sigma_s and sigma_a at one point: both are functions of the same two grid lookups (density, albedo),
// which the reference performs twice (media/heterogeneous.inl:11-21); identical values, half the gathers.
LJ_D void medium_sigma_sa(const DScene &sc, const lj_medium &m, d3 p, d3 &sigma_s, d3 &sigma_a) {
    if (m.kind == LJ_MEDIUM_HOMOGENEOUS) {
        sigma_s = mk3(m.sigma_s[0], m.sigma_s[1], m.sigma_s[2]);
        sigma_a = mk3(m.sigma_a[0], m.sigma_a[1], m.sigma_a[2]);
        return;
    }
    d3 density, albedo;
    if (m.density._pad == 8) {
        // density and albedo share their cells (paired records, lj_api.cu): one cell computation, 8 x 32 contiguous bytes;
        // each grid's value is volume_lookup's (same corners, same weights, same blend, its own scale)
        GridCell c;
        if (!grid_cell(m.density, p, c)) { density = mk3(0.0); albedo = mk3(0.0); }
        else {
            const float4 *g4 = (const float4 *)(sc.grid_f32 + m.density.data_offset);
            density = volume_blend(m.density.scale, c.dx, c.dy, c.dz, load3f4(g4, c.i000), load3f4(g4, c.i001), load3f4(g4, c.i010), load3f4(g4, c.i011),
                                   load3f4(g4, c.i100), load3f4(g4, c.i101), load3f4(g4, c.i110), load3f4(g4, c.i111));
            g4 += 1;
            albedo = volume_blend(m.albedo.scale, c.dx, c.dy, c.dz, load3f4(g4, c.i000), load3f4(g4, c.i001), load3f4(g4, c.i010), load3f4(g4, c.i011),
                                  load3f4(g4, c.i100), load3f4(g4, c.i101), load3f4(g4, c.i110), load3f4(g4, c.i111));
        }
    } else {
        density = volume_lookup(sc, m.density, p);
        albedo = volume_lookup(sc, m.albedo, p);
    }
    sigma_s = density * albedo;
    sigma_a = density * (1.0 - albedo);
}

// ---- phase functions (phase_functions/isotropic.inl, henyeygreenstein.inl) ----
LJ_D double phase_value(const lj_medium &m, d3 dir_in, d3 dir_out) {
    if (m.phase_kind == LJ_PHASE_ISOTROPIC) return LJ_INVFOURPI;
    const double g = m.g;
    return LJ_INVFOURPI * (1 - g * g) / (pow((1 + g * g + 2 * g * dot(dir_in, dir_out)), 3.0 / 2.0));
}
LJ_D d3 phase_eval(const lj_medium &m, d3 dir_in, d3 dir_out) { return mk3(phase_value(m, dir_in, dir_out)); }
LJ_D double phase_pdf(const lj_medium &m, d3 dir_in, d3 dir_out) { return phase_value(m, dir_in, dir_out); }
LJ_D d3 phase_sample(const lj_medium &m, d3 dir_in, d2 rnd) {
    const double g = m.g;
    if (m.phase_kind == LJ_PHASE_ISOTROPIC || fabs(g) < 1e-3) {
        double z = 1 - 2 * rnd.x;
        double r = sqrt(fmax(0.0, 1 - z * z));
        double phi = 2 * LJ_PI * rnd.y;
        return mk3(r * cos(phi), r * sin(phi), z);
    }
    double tmp = (g * g - 1) / (2 * rnd.x * g - (g + 1));
    double cos_elevation = (tmp * tmp - (1 + g * g)) / (2 * g);
    double sin_elevation = sqrt(lmax(1 - cos_elevation * cos_elevation, 0.0));
    double azimuth = 2 * LJ_PI * rnd.y;
    Frame frame = make_frame(dir_in);
    return to_world(frame, mk3(sin_elevation * cos(azimuth), sin_elevation * sin(azimuth), cos_elevation));
}

// vol_path_tracing.h:149-163
LJ_D int update_medium(const PathVertex &vertex, d3 ray_dir, int medium) {
    if (vertex.interior_medium_id != vertex.exterior_medium_id) {
        if (dot(ray_dir, vertex.geometry_normal) > 0) medium = vertex.exterior_medium_id;
        else medium = vertex.interior_medium_id;
    }
    return medium;
}

LJ_D Ray volpath_primary(const DScene &sc, int x, int y, Pcg32 &rng) {
    const int w = sc.camera.width, h = sc.camera.height;
    // Vector2 screen_pos((x + next()) / w, (y + next()) / h): constructor arguments, evaluated right to left by g++
    double u_first = pcg_real(rng);
    double u_second = pcg_real(rng);
    return sample_primary(sc.camera, mk2((x + u_second) / w, (y + u_first) / h));
}

// ---- vol_path_tracing.h:6-41 ----
LJ_D d3 vol_path_tracing_1(const DScene &sc, int x, int y, Pcg32 &rng, RayCounters &rc) {
    Ray ray = volpath_primary(sc, x, y, rng);
    RayDiff rd; rd.radius = 0; rd.spread = 0;
    PathVertex vertex;
    rc.primary++;
    if (scene_intersect(sc, ray, rd, vertex)) {
        if (vertex.exterior_medium_id != -1) {
            const lj_medium &m = sc.media[vertex.exterior_medium_id];
            double t_hit = distance(vertex.position, ray.org);
            d3 sigma_a = medium_sigma_a(sc, m, vertex.position);
            d3 transmittance = exp3(-sigma_a * t_hit);
            d3 Le = mk3(0.0);
            if (sc.shapes[vertex.shape_id].area_light_id >= 0) Le = vertex_emission(sc, vertex, -ray.dir);
            return transmittance * Le;
        }
    }
    return mk3(0.0);
}

// ---- vol_path_tracing.h:46-147 ----
LJ_D d3 vol_path_tracing_2(const DScene &sc, int x, int y, Pcg32 &rng, RayCounters &rc) {
    Ray ray = volpath_primary(sc, x, y, rng);
    RayDiff rd; rd.radius = 0; rd.spread = 0;
    PathVertex vertex;
    vertex.position = mk3(0.0);
    rc.primary++;
    bool hit = scene_intersect(sc, ray, rd, vertex);
    int medium_id; double t_hit;
    if (hit) { medium_id = vertex.exterior_medium_id; t_hit = distance(vertex.position, ray.org); }
    else { medium_id = sc.camera.medium_id; t_hit = INFINITY; }
    const lj_medium &m = sc.media[medium_id];
    d3 sigma_s = medium_sigma_s(sc, m, vertex.position);
    d3 sigma_a = medium_sigma_a(sc, m, vertex.position);
    d3 sigma_t = sigma_s + sigma_a;
    double u = pcg_real(rng);
    double t = -log(1 - u) / sigma_t.x;
    if (t < t_hit) {
        d3 trans_pdf = exp3(-sigma_t * t) * sigma_t;
        d3 transmittance = exp3(-sigma_t * t);
        d3 p = ray.org + t * ray.dir;
        d2 light_uv; light_uv.x = pcg_real(rng); light_uv.y = pcg_real(rng);
        double light_w = pcg_real(rng);
        double shape_w = pcg_real(rng);
        int light_id = sample_table1d(sc.light_cdf, sc.num_lights, light_w);
        const lj_light &light = sc.lights[light_id];
        PointAndNormal pl = sample_point_on_light(sc, light, p, light_uv, shape_w);
        d3 dir_view = -ray.dir;
        d3 dir_light = normalize(pl.position - p);
        d3 rho = phase_eval(m, dir_view, dir_light);
        d3 Le = light_emission(sc, light, -dir_light, 0.0, pl);
        d3 exp_term = exp3(-sigma_t * (distance(p, pl.position)));
        double visibility = 1;
        Ray shadow_ray = make_ray(p, dir_light, sc.eps, (1 - sc.eps) * distance(pl.position, p));
        rc.shadow++;
        if (scene_occluded(sc, shadow_ray)) visibility = 0;
        double jacobian = fabs(dot(dir_light, pl.normal)) / distance_squared(p, pl.position) * visibility;
        d3 L_s1_estimate = rho * Le * exp_term * jacobian;
        double L_s1_pdf = __ldg(sc.light_pmf + light_id) * pdf_point_on_light(sc, light, pl, p);
        return (transmittance / trans_pdf) * sigma_s * (L_s1_estimate / L_s1_pdf);
    } else {
        d3 trans_pdf = exp3(-sigma_t * t_hit);
        d3 transmittance = exp3(-sigma_t * t_hit);
        d3 Le = mk3(0.0);
        if (sc.shapes[vertex.shape_id].area_light_id >= 0) Le = vertex_emission(sc, vertex, -ray.dir);
        return transmittance / trans_pdf * Le;
    }
}

// Tracking loops in begin / step form, so that the per-lane integrator (run to completion) and the
// wavefront tracking kernel (lj_volwave.cu: one collision per lane per warp iteration, finished lanes
// refilled) execute the very same arithmetic.
//
// The steps are written over a spectrum type S: d3 (the reference's Spectrum), or double for a GREY scene — every
// medium has equal channels at every point (media_all_grey).  Then the three channels of every quantity of the loops
// are one number produced by one sequence of operations, and computing a single channel gives the reference's bits
// at a third of the arithmetic (one exp instead of three, no per-channel divisions).
LJ_D d3 sp_exp(d3 a) { return exp3(a); }
LJ_D double sp_exp(double a) { return exp(a); }
LJ_D double sp_max(d3 a) { return max3(a); }
LJ_D double sp_max(double a) { return a; }
LJ_D double sp_comp(d3 a, int c) { return comp(a, c); }
LJ_D double sp_comp(double a, int) { return a; }
LJ_D d3 sp_quot(d3 a, d3 b) { return a / b; }                                        // d3 / d3: true division per channel
LJ_D double sp_quot(double a, double b) { return div_z(a, b); }

LJ_D bool medium_is_grey(const lj_medium &m) {
    if (m.kind == LJ_MEDIUM_HOMOGENEOUS)
        return m.sigma_a[0] == m.sigma_a[1] && m.sigma_a[0] == m.sigma_a[2] && m.sigma_s[0] == m.sigma_s[1] && m.sigma_s[0] == m.sigma_s[2];
    return volume_is_grey(m.density) && volume_is_grey(m.albedo) &&
           (m.density.kind == LJ_VOLUME_CONSTANT || (m.density.max_data[0] == m.density.max_data[1] && m.density.max_data[0] == m.density.max_data[2]));
}
template <typename S> LJ_D S medium_majorant_t(const DScene &sc, const lj_medium &m, const Ray &ray) {
    if constexpr (sizeof(S) == sizeof(d3)) return medium_majorant(sc, m, ray);
    else {      // channel 0 of medium_majorant
        if (m.kind == LJ_MEDIUM_HOMOGENEOUS) return m.sigma_a[0] + m.sigma_s[0];
        if (!volume_intersect(m.density, ray)) return 0.0;
        if (m.density.kind == LJ_VOLUME_CONSTANT) return m.density.value[0];
        return m.density.scale * m.density.max_data[0];
    }
}
LJ_D void medium_sigma_sa(const DScene &sc, const lj_medium &m, d3 p, double &sigma_s, double &sigma_a) {    // channel 0 of the d3 overload
    if (m.kind == LJ_MEDIUM_HOMOGENEOUS) { sigma_s = m.sigma_s[0]; sigma_a = m.sigma_a[0]; return; }
    const double density = volume_lookup_grey(sc, m.density, p);
    const double albedo = volume_lookup_grey(sc, m.albedo, p);
    sigma_s = density * albedo;
    sigma_a = density * (1.0 - albedo);
}

// inv_mx = 1 / max(majorant): the reciprocal every `spectrum / max(majorant)` of a step multiplies by (vector.h:193-197), computed once
// per segment instead of once per step -- same operands, same value
template <typename S> struct TrackStateT { S majorant; double maj_c, accum_t, inv_mx; int channel, iteration; bool active; };
typedef TrackStateT<d3> TrackState;

template <typename S>
LJ_D void track_begin(const DScene &sc, const lj_medium &media, const Ray &ray, Pcg32 &rng, TrackStateT<S> &ts) {
    ts.majorant = medium_majorant_t<S>(sc, media, ray);
    double u = pcg_real(rng);
    ts.channel = clampi((int)(u * 3), 0, 2);
    ts.accum_t = 0;
    ts.iteration = 0;
    ts.maj_c = sp_comp(ts.majorant, ts.channel);
    ts.inv_mx = 1.0 / sp_max(ts.majorant);
    ts.active = !(ts.maj_c <= 0) && !(ts.iteration >= sc.options.max_null_collisions);
}
// `a / s` of a spectrum by a scalar whose reciprocal is at hand: the multiplications of operator/(d3, double)
LJ_D d3 sp_mul_inv(d3 a, double inv) { return mk3(a.x * inv, a.y * inv, a.z * inv); }
LJ_D double sp_mul_inv(double a, double inv) { return a * inv; }

// One iteration of the NEE ratio-tracking loop (vol_path_tracing.h:364-408).
template <typename S>
LJ_D void ratio_track_step(const DScene &sc, const lj_medium &media, d3 p, d3 dir_light, double next_t, Pcg32 &rng, TrackStateT<S> &ts,
                           S &transmittance_light, S &p_trans_nee, S &p_trans_dir) {
    const S majorant = ts.majorant;
    double t = -log(1 - pcg_real(rng)) / ts.maj_c;
    double dt = next_t - ts.accum_t;
    { const double s_ = ts.accum_t + t; ts.accum_t = (next_t < s_) ? next_t : s_; }   // std::min(accum_t + t, next_t)
    if (t < dt) {
        d3 pos = p + dir_light * ts.accum_t;
        S sigma_s, sigma_a;
        medium_sigma_sa(sc, media, pos, sigma_s, sigma_a);
        S sigma_t = sigma_s + sigma_a;
        S real_prob = sp_quot(sigma_t, majorant);       // sigma_t / majorant; zero in empty voxels: div_z
        S sigma_n = majorant * (1.0 - real_prob);
        S e = sp_exp(-majorant * t);
        const double inv_mx = ts.inv_mx;               // `x / max(majorant)` = x * (1 / max(majorant)), vector.h:193-197
        transmittance_light = transmittance_light * sp_mul_inv(e * sigma_n, inv_mx);
        p_trans_nee = p_trans_nee * sp_mul_inv(e * majorant, inv_mx);
        p_trans_dir = p_trans_dir * sp_mul_inv(e * majorant * (1.0 - real_prob), inv_mx);
        if (sp_max(transmittance_light) <= 0) { ts.active = false; return; }
    } else {
        S e = sp_exp(-majorant * dt);
        transmittance_light = transmittance_light * e;
        p_trans_nee = p_trans_nee * e;
        p_trans_dir = p_trans_dir * e;
        ts.active = false; return;
    }
    ts.iteration++;
    if (ts.iteration >= sc.options.max_null_collisions) ts.active = false;
}

// One NEE segment through a medium: ratio tracking with per-channel spectral MIS (vol_path_tracing.h:354-410).
LJ_D void ratio_track_segment(const DScene &sc, const lj_medium &media, const Ray &shadow_ray, d3 p, d3 dir_light, double next_t,
                              Pcg32 &rng, d3 &transmittance_light, d3 &p_trans_nee, d3 &p_trans_dir) {
    TrackState ts;
    track_begin(sc, media, shadow_ray, rng, ts);
    while (ts.active) ratio_track_step(sc, media, p, dir_light, next_t, rng, ts, transmittance_light, p_trans_nee, p_trans_dir);
}

// One iteration of the free-flight (delta tracking) loop (vol_path_tracing.h:565-621).  Sets `scatter`
// and moves ray.org on a real collision or when the surface is reached.
template <typename S>
LJ_D void free_flight_step(const DScene &sc, const lj_medium &media, Ray &ray, double t_hit, d3 hit_position, Pcg32 &rng, TrackStateT<S> &ts,
                           S &transmittance, S &trans_dir_pdf, S &trans_nee_pdf, bool &scatter, bool *reached_surface = nullptr) {
    const S majorant = ts.majorant;
    double t = -log(1 - pcg_real(rng)) / ts.maj_c;
    double dt = t_hit - ts.accum_t;
    { const double s_ = ts.accum_t + t; ts.accum_t = (t_hit < s_) ? t_hit : s_; }
    if (t < dt) {
        d3 p = ray.org + ray.dir * ts.accum_t;
        S sigma_s, sigma_a;
        medium_sigma_sa(sc, media, p, sigma_s, sigma_a);
        S sigma_t = sigma_s + sigma_a;
        S real_prob = sp_quot(sigma_t, majorant);       // sigma_t / majorant; zero in empty voxels: div_z
        S sigma_n = majorant * (1.0 - real_prob);
        S e = sp_exp(-majorant * t);
        const double inv_mx = ts.inv_mx;               // `x / max(majorant)` = x * (1 / max(majorant)), vector.h:193-197
        if (pcg_real(rng) < sp_comp(real_prob, ts.channel)) {
            transmittance = transmittance * sp_mul_inv(e, inv_mx);
            trans_dir_pdf = trans_dir_pdf * sp_mul_inv(e * majorant * real_prob, inv_mx);
            ray.org = p;
            scatter = true;
            ts.active = false; return;
        } else {
            transmittance = transmittance * sp_mul_inv(e * sigma_n, inv_mx);
            trans_dir_pdf = trans_dir_pdf * sp_mul_inv(e * majorant * (1.0 - real_prob), inv_mx);
            trans_nee_pdf = trans_nee_pdf * sp_mul_inv(e * majorant, inv_mx);
        }
    } else {
        S e = sp_exp(-majorant * dt);
        transmittance = transmittance * e;
        trans_dir_pdf = trans_dir_pdf * e;
        trans_nee_pdf = trans_nee_pdf * e;
        // dt finite => a hit exists.  A caller that does not want to carry hit_position through its loop (the tracking
        // kernel: six registers) passes a flag instead and moves the origin itself afterwards.
        if (reached_surface) *reached_surface = true; else ray.org = hit_position;
        ts.active = false; return;
    }
    ts.iteration++;
    if (ts.iteration >= sc.options.max_null_collisions) ts.active = false;
}

// Free-flight sampling along the main ray through a medium: delta tracking with null collisions and
// per-channel spectral MIS (vol_path_tracing.h:556-621).  Moves ray.org to the scattering point (returns
// true) or to the surface hit; leaves it where it is when the sampled channel's majorant is zero or the
// null-collision cap is reached (as the reference does).
LJ_D bool free_flight(const DScene &sc, const lj_medium &media, Ray &ray, double t_hit, d3 hit_position, Pcg32 &rng,
                      d3 &transmittance, d3 &trans_dir_pdf, d3 &trans_nee_pdf) {
    TrackState ts;
    track_begin(sc, media, ray, rng, ts);
    bool scatter = false;
    while (ts.active) free_flight_step(sc, media, ray, t_hit, hit_position, rng, ts, transmittance, trans_dir_pdf, trans_nee_pdf, scatter);
    return scatter;
}

// ---- vol_path_tracing.h:299-495 ----
// Light sample + transmittance to it by ratio tracking through media and index-matched
// surfaces (closest-hit chain: the reference calls intersect(), not occluded(), :343).
LJ_D d3 next_event_estimation_final(const DScene &sc, Pcg32 &rng, d3 p, int current_medium, int bounces, d3 dir_view,
                                    bool is_surface, const PathVertex &vertex, RayCounters &rc) {
    d2 light_uv; light_uv.x = pcg_real(rng); light_uv.y = pcg_real(rng);
    double light_w = pcg_real(rng);
    double shape_w = pcg_real(rng);
    int light_id = sample_table1d(sc.light_cdf, sc.num_lights, light_w);
    const lj_light &light = sc.lights[light_id];
    PointAndNormal pl = sample_point_on_light(sc, light, p, light_uv, shape_w);
    d3 dir_light = normalize(pl.position - p);
    d3 p_prime = pl.position;
    d3 p_origin = p;

    int shadow_medium = current_medium;
    int shadow_bounces = 0;
    d3 transmittance_light = mk3(1.0);
    d3 p_trans_nee = mk3(1.0);
    d3 p_trans_dir = mk3(1.0);
    const double eps = sc.eps;
    const int max_depth = sc.options.max_depth;

    while (true) {
        Ray shadow_ray = make_ray(p, dir_light, eps, (1 - eps) * distance(p, p_prime));
        RayDiff rd; rd.radius = 0; rd.spread = 0;
        PathVertex shadow_vertex;
        rc.extension++;
        bool hit = scene_intersect(sc, shadow_ray, rd, shadow_vertex);
        double next_t = distance(p, p_prime);
        if (hit) next_t = distance(p, shadow_vertex.position);

        if (shadow_medium != -1)
            ratio_track_segment(sc, sc.media[shadow_medium], shadow_ray, p, dir_light, next_t, rng, transmittance_light, p_trans_nee, p_trans_dir);

        if (!hit) break;
        if (shadow_vertex.material_id >= 0) return mk3(0.0);        // blocked by an opaque surface
        shadow_bounces++;
        if (max_depth != -1 && bounces + shadow_bounces >= max_depth) return mk3(0.0);
        shadow_medium = update_medium(shadow_vertex, shadow_ray.dir, shadow_medium);
        p = p + next_t * dir_light;
    }

    if (max3(transmittance_light) > 0) {
        d3 pdf_dir = mk3(0.0);
        d3 evaluated_f = mk3(0.0);
        d3 Le = light_emission(sc, light, -dir_light, 0.0, pl);
        double jacobian = lmax(-dot(dir_light, pl.normal), 0.0) / distance_squared(p_origin, p_prime);
        d3 pdf_nee = (__ldg(sc.light_pmf + light_id) * pdf_point_on_light(sc, light, pl, p_origin)) * p_trans_nee;
        if (is_surface) {
            MatParams mat = load_mat_params(sc, vertex);
            evaluated_f = bsdf_eval(mat, vertex, dir_view, dir_light);
            double pdf_bsdf = bsdf_pdf(mat, vertex, dir_view, dir_light);
            if (pdf_bsdf <= 0) return mk3(0.0);
            pdf_dir = (pdf_bsdf * jacobian) * p_trans_dir;
        } else {
            const lj_medium &media = sc.media[current_medium];
            pcg_real(rng); pcg_real(rng);                            // the two unused phase_uv draws, :479
            evaluated_f = phase_eval(media, dir_view, dir_light);
            double pdf_phase = phase_pdf(media, dir_view, dir_light);
            pdf_dir = (pdf_phase * jacobian) * p_trans_dir;
        }
        d3 contrib = transmittance_light * evaluated_f * Le * jacobian / average(pdf_nee);
        d3 w = (pdf_nee * pdf_nee) / (pdf_nee * pdf_nee + pdf_dir * pdf_dir);
        return contrib * w;
    }
    return mk3(0.0);
}

// ---- vol_path_tracing.h:503-869 ----
LJ_D d3 vol_path_tracing_final(const DScene &sc, int x, int y, Pcg32 &rng, RayCounters &rc) {
    Ray ray = volpath_primary(sc, x, y, rng);
    ray.tnear = sc.eps;                                   // get_intersection_epsilon, :513
    RayDiff ray_diff; ray_diff.radius = 0; ray_diff.spread = 0;

    int current_medium = sc.camera.medium_id;
    d3 current_path_throughput = mk3(1.0);
    d3 radiance = mk3(0.0);
    int bounces = 0;
    double dir_pdf = 0;
    d3 nee_p_cache = mk3(0.0);
    d3 multi_trans_pdf = mk3(1.0);
    double eta_scale = 1;
    const int max_depth = sc.options.max_depth;
    bool first = true;

    while (true) {
        bool scatter = false;
        PathVertex vertex;
        if (first) { rc.primary++; first = false; } else rc.extension++;
        const bool hit = scene_intersect(sc, ray, ray_diff, vertex);
        double t_hit = INFINITY;
        if (hit) t_hit = distance(vertex.position, ray.org);

        d3 transmittance = mk3(1.0);
        d3 trans_dir_pdf = mk3(1.0);
        d3 trans_nee_pdf = mk3(1.0);

        if (current_medium != -1) {
            scatter = free_flight(sc, sc.media[current_medium], ray, t_hit, vertex.position, rng, transmittance, trans_dir_pdf, trans_nee_pdf);
            multi_trans_pdf = multi_trans_pdf * trans_dir_pdf;
        } else {
            if (hit) ray.org = vertex.position;
            else return mk3(0.0);                          // :637 (drops the radiance gathered so far, as the reference does)
        }

        current_path_throughput = current_path_throughput * (transmittance / average(trans_dir_pdf));

        if (!scatter && hit && sc.shapes[vertex.shape_id].area_light_id >= 0) {
            d3 Le = vertex_emission(sc, vertex, -ray.dir);
            if (bounces == 0) {
                radiance += current_path_throughput * Le;
                return radiance;
            } else {
                PointAndNormal light_point; light_point.position = vertex.position; light_point.normal = vertex.geometry_normal;
                int light_id = sc.shapes[vertex.shape_id].area_light_id;
                const lj_light &light = sc.lights[light_id];
                d3 pdf_nee = (__ldg(sc.light_pmf + light_id) * pdf_point_on_light(sc, light, light_point, nee_p_cache)) * trans_nee_pdf;
                double jacobian = lmax(-dot(-ray.dir, light_point.normal), 0.0) / distance_squared(nee_p_cache, light_point.position);
                d3 pdf_phase = (dir_pdf * multi_trans_pdf) * jacobian;
                d3 w = (pdf_phase * pdf_phase) / (pdf_phase * pdf_phase + pdf_nee * pdf_nee);
                radiance += current_path_throughput * Le * w;
            }
        }

        if (!scatter && hit && vertex.material_id == -1) {
            current_medium = update_medium(vertex, ray.dir, current_medium);
            ray.org = vertex.position;
            bounces++;
            continue;
        }

        if (bounces >= max_depth - 1 && max_depth != -1) break;

        if (scatter && current_medium != -1) {
            const lj_medium &media = sc.media[current_medium];
            d3 sigma_s = medium_sigma_s(sc, media, ray.org);
            d3 nee = next_event_estimation_final(sc, rng, ray.org, current_medium, bounces, -ray.dir, false, vertex, rc);
            radiance += current_path_throughput * sigma_s * nee;
            if (max3(nee) > 0) nee_p_cache = ray.org;
            d2 phase_uv; phase_uv.x = pcg_real(rng); phase_uv.y = pcg_real(rng);
            d3 next_dir = phase_sample(media, -ray.dir, phase_uv);
            double ppdf = phase_pdf(media, -ray.dir, next_dir);
            current_path_throughput = current_path_throughput * (phase_eval(media, -ray.dir, next_dir) / ppdf * sigma_s);
            ray.dir = next_dir;
            dir_pdf = ppdf;
            multi_trans_pdf = mk3(1.0);
        } else if (hit) {
            d3 nee = next_event_estimation_final(sc, rng, ray.org, current_medium, bounces, -ray.dir, true, vertex, rc);
            radiance += current_path_throughput * nee;
            if (max3(nee) > 0) nee_p_cache = ray.org;
            MatParams mat = load_mat_params(sc, vertex);
            d3 dir_view = -ray.dir;
            d2 bsdf_uv; bsdf_uv.x = pcg_real(rng); bsdf_uv.y = pcg_real(rng);
            double bsdf_w = pcg_real(rng);
            BSDFSample bs;
            if (!bsdf_sample(mat, vertex, dir_view, bsdf_uv, bsdf_w, bs)) break;
            d3 dir_bsdf = bs.dir_out;
            ray.dir = dir_bsdf;
            if (bs.eta == 0) {
                ray_diff.spread = rd_reflect(ray_diff, vertex.mean_curvature, bs.roughness);
            } else {
                ray_diff.spread = rd_refract(ray_diff, vertex.mean_curvature, bs.eta, bs.roughness);
                eta_scale /= (bs.eta * bs.eta);
                current_medium = update_medium(vertex, ray.dir, current_medium);
            }
            d3 f = bsdf_eval(mat, vertex, dir_view, dir_bsdf);
            double pdf_bsdf = bsdf_pdf(mat, vertex, dir_view, dir_bsdf);
            current_path_throughput = current_path_throughput * (f / pdf_bsdf);
        }

        if (bounces >= sc.options.rr_depth) {
            double rr_prob = lmin(max3((1 / eta_scale) * current_path_throughput), 0.95);
            double u = pcg_real(rng);
            if (u > rr_prob) break;
            current_path_throughput = current_path_throughput / rr_prob;
        }
        bounces++;
    }
    return radiance;
}

// render.cpp:111-124: version -> integrator
LJ_D d3 vol_path_tracing_dispatch(const DScene &sc, int x, int y, Pcg32 &rng, RayCounters &rc) {
    const int v = sc.options.vol_path_version;
    if (v == 1) return vol_path_tracing_1(sc, x, y, rng, rc);
    if (v == 2) return vol_path_tracing_2(sc, x, y, rng, rc);
    return vol_path_tracing_final(sc, x, y, rng, rc);
}
