// Unidirectional path tracer with NEE + MIS, one radiance sample per call.
// Restates /root/reference/src/path_tracing.h:7-325 (same RNG draw order, same
// early exits, same operation order).  Used directly by the tile-serial parity
// kernel and the megakernel engine; the wavefront engine (lj_wavefront.cu)
// runs the same steps split across stages.
#pragma once
#include "lj_materials.cuh"
#include "lj_traverse8.cuh"

struct RayCounters { unsigned long long primary, extension, shadow; };

LJ_D bool scene_intersect(const DScene &sc, const Ray &ray, const RayDiff &rd, PathVertex &v) {
    RayF rf = to_rayf(ray);
    Hit h;
    if (!trace_closest(sc, rf, &h)) return false;
    v = make_vertex(sc, ray, h, rd);
    return true;
}
LJ_D bool scene_occluded(const DScene &sc, const Ray &ray) {
    RayF rf = to_rayf(ray);
    return trace_any(sc, rf);
}


LJ_D d3 path_tracing_sample(const DScene &sc, int x, int y, Pcg32 &rng, RayCounters &rc) {
    const int w = sc.camera.width, h = sc.camera.height;
    // path_tracing.h:11-12: the two draws are constructor arguments; g++ evaluates
    // them right to left, so the FIRST draw feeds y and the second feeds x.
    double u_first = pcg_real(rng);
    double u_second = pcg_real(rng);
    d2 screen_pos = mk2((x + u_second) / w, (y + u_first) / h);
    Ray ray = sample_primary(sc.camera, screen_pos);
    RayDiff ray_diff = init_ray_differential(w, h);

    PathVertex vertex;
    rc.primary++;
    if (!scene_intersect(sc, ray, ray_diff, vertex)) {
        if (sc.envmap_light_id != -1) {
            PointAndNormal dummy; dummy.position = mk3(0.0); dummy.normal = mk3(0.0);
            return light_emission(sc, sc.lights[sc.envmap_light_id], -ray.dir, ray_diff.spread, dummy);
        }
        return mk3(0.0);
    }

    d3 radiance = mk3(0.0);
    d3 throughput = mk3(1.0);
    double eta_scale = 1;
    if (sc.shapes[vertex.shape_id].area_light_id >= 0) {
        radiance += throughput * vertex_emission(sc, vertex, -ray.dir);
    }

    const int max_depth = sc.options.max_depth;
    const double eps = sc.eps;
    for (int num_vertices = 3; max_depth == -1 || num_vertices <= max_depth + 1; num_vertices++) {
        MatParams mat = load_mat_params(sc, vertex);
        d2 light_uv; light_uv.x = pcg_real(rng); light_uv.y = pcg_real(rng);
        double light_w = pcg_real(rng);
        double shape_w = pcg_real(rng);
        int light_id = sample_table1d(sc.light_cdf, sc.num_lights, light_w);
        const lj_light &light = sc.lights[light_id];
        PointAndNormal pl = sample_point_on_light(sc, light, vertex.position, light_uv, shape_w);

        d3 C1 = mk3(0.0);
        double w1 = 0;
        {
            double G = 0;
            d3 dir_light;
            if (light.kind != LJ_LIGHT_ENVMAP) {
                dir_light = normalize(pl.position - vertex.position);
                Ray shadow_ray = make_ray(vertex.position, dir_light, eps, (1 - eps) * distance(pl.position, vertex.position));
                rc.shadow++;
                if (!scene_occluded(sc, shadow_ray)) {
                    G = lmax(-dot(dir_light, pl.normal), 0.0) / distance_squared(pl.position, vertex.position);
                }
            } else {
                dir_light = -pl.normal;
                Ray shadow_ray = make_ray(vertex.position, dir_light, eps, INFINITY);
                rc.shadow++;
                if (!scene_occluded(sc, shadow_ray)) G = 1;
            }
            double p1 = __ldg(sc.light_pmf + light_id) * pdf_point_on_light(sc, light, pl, vertex.position);
            if (G > 0 && p1 > 0) {
                d3 dir_view = -ray.dir;
                d3 f = bsdf_eval(mat, vertex, dir_view, dir_light);
                d3 L = light_emission(sc, light, -dir_light, 0.0, pl);
                C1 = G * f * L;
                double p2 = bsdf_pdf(mat, vertex, dir_view, dir_light);
                p2 *= G;
                w1 = (p1 * p1) / (p1 * p1 + p2 * p2);
                C1 = C1 / p1;
            }
        }
        radiance += throughput * C1 * w1;

        d3 dir_view = -ray.dir;
        d2 bsdf_uv; bsdf_uv.x = pcg_real(rng); bsdf_uv.y = pcg_real(rng);
        double bsdf_w = pcg_real(rng);
        BSDFSample bs;
        if (!bsdf_sample(mat, vertex, dir_view, bsdf_uv, bsdf_w, bs)) break;
        d3 dir_bsdf = bs.dir_out;
        if (bs.eta == 0) {
            ray_diff.spread = rd_reflect(ray_diff, vertex.mean_curvature, bs.roughness);
        } else {
            ray_diff.spread = rd_refract(ray_diff, vertex.mean_curvature, bs.eta, bs.roughness);
            eta_scale /= (bs.eta * bs.eta);
        }
        Ray bsdf_ray = make_ray(vertex.position, dir_bsdf, eps, INFINITY);
        PathVertex bsdf_vertex;
        RayDiff zero_rd; zero_rd.radius = 0; zero_rd.spread = 0;   // intersect(scene, bsdf_ray) uses the default RayDifferential{}
        rc.extension++;
        bool hit = scene_intersect(sc, bsdf_ray, zero_rd, bsdf_vertex);
        double G;
        if (hit) G = fabs(dot(dir_bsdf, bsdf_vertex.geometry_normal)) / distance_squared(bsdf_vertex.position, vertex.position);
        else G = 1;
        d3 f = bsdf_eval(mat, vertex, dir_view, dir_bsdf);
        double p2 = bsdf_pdf(mat, vertex, dir_view, dir_bsdf);
        if (p2 <= 0) break;
        p2 *= G;
        if (hit && sc.shapes[bsdf_vertex.shape_id].area_light_id >= 0) {
            d3 L = vertex_emission(sc, bsdf_vertex, -dir_bsdf);
            d3 C2 = G * f * L;
            int lid = sc.shapes[bsdf_vertex.shape_id].area_light_id;
            PointAndNormal lp; lp.position = bsdf_vertex.position; lp.normal = bsdf_vertex.geometry_normal;
            double p1 = __ldg(sc.light_pmf + lid) * pdf_point_on_light(sc, sc.lights[lid], lp, vertex.position);
            double w2 = (p2 * p2) / (p1 * p1 + p2 * p2);
            C2 = C2 / p2;
            radiance += throughput * C2 * w2;
        } else if (!hit && sc.envmap_light_id != -1) {
            const lj_light &env = sc.lights[sc.envmap_light_id];
            PointAndNormal dummy; dummy.position = mk3(0.0); dummy.normal = mk3(0.0);
            d3 L = light_emission(sc, env, -dir_bsdf, ray_diff.spread, dummy);
            d3 C2 = G * f * L;
            PointAndNormal lp; lp.position = mk3(0.0); lp.normal = -dir_bsdf;
            double p1 = __ldg(sc.light_pmf + sc.envmap_light_id) * pdf_point_on_light(sc, env, lp, vertex.position);
            double w2 = (p2 * p2) / (p1 * p1 + p2 * p2);
            C2 = C2 / p2;
            radiance += throughput * C2 * w2;
        }
        if (!hit) break;
        double rr_prob = 1;
        if (num_vertices - 1 >= sc.options.rr_depth) {
            rr_prob = lmin(max3((1 / eta_scale) * throughput), 0.95);
            if (pcg_real(rng) > rr_prob) break;
        }
        ray = bsdf_ray;
        vertex = bsdf_vertex;
        throughput = throughput * (G * f) / (p2 * rr_prob);
    }
    return radiance;
}
