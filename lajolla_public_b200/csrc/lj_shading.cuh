// Per-lane evaluation of textures, shape shading data, shape / light sampling.
// Each function restates the reference routine cited next to it (paths under
// /root/reference/src) with the same operation order, in f64.
#pragma once
#include "lj_scene.cuh"

// ---------------- textures (texture.h:94-154, mipmap.h:50-88) ----------------
template <int C> struct Texel;
template <> struct Texel<1> { typedef double T; };
template <> struct Texel<3> { typedef d3 T; };

LJ_D double texel_load(const double *p, size_t i, double *) { return __ldg(p + i); }
LJ_D d3 texel_load(const double *p, size_t i, d3 *) { return mk3(__ldg(p + 3 * i), __ldg(p + 3 * i + 1), __ldg(p + 3 * i + 2)); }

// level-0 texels kept as f32 (DScene::texels0_*): same values, widened
LJ_D unsigned tex0_offset(const DScene &sc, int image_id, double *) { return sc.tex0_off1 ? __ldg(sc.tex0_off1 + image_id) : 0xffffffffu; }
LJ_D unsigned tex0_offset(const DScene &sc, int image_id, d3 *) { return sc.tex0_off3 ? __ldg(sc.tex0_off3 + image_id) : 0xffffffffu; }
LJ_D double texel0_load(const DScene &sc, size_t i, double *) { return (double)__ldg(sc.texels0_r + i); }
LJ_D d3 texel0_load(const DScene &sc, size_t i, d3 *) { const float4 t = __ldg(sc.texels0_rgb + i); return mk3((double)t.x, (double)t.y, (double)t.z); }

// mipmap.h:50-69 lookup at an integer level
template <typename T>
LJ_D T mip_lookup_level(const DScene &sc, const lj_image &img, int image_id, double u, double v, int level) {
    int W = img.width[level], H = img.height[level];
    const double *base = sc.texels + img.offset[level];
    const unsigned off0 = level == 0 ? tex0_offset(sc, image_id, (T *)nullptr) : 0xffffffffu;
    u = u * W - 0.5;
    v = v * H - 0.5;
    int ufi = modulo_i((int)u, W);
    int vfi = modulo_i((int)v, H);
    int uci = modulo_i(ufi + 1, W);
    int vci = modulo_i(vfi + 1, H);
    double u_off = u - ufi;
    double v_off = v - vfi;
    T ff, fc, cf, cc;
    if (off0 != 0xffffffffu) {
        ff = texel0_load(sc, off0 + (size_t)vfi * W + ufi, (T *)nullptr);
        fc = texel0_load(sc, off0 + (size_t)vci * W + ufi, (T *)nullptr);
        cf = texel0_load(sc, off0 + (size_t)vfi * W + uci, (T *)nullptr);
        cc = texel0_load(sc, off0 + (size_t)vci * W + uci, (T *)nullptr);
    } else {
        ff = texel_load(base, (size_t)vfi * W + ufi, (T *)nullptr);
        fc = texel_load(base, (size_t)vci * W + ufi, (T *)nullptr);
        cf = texel_load(base, (size_t)vfi * W + uci, (T *)nullptr);
        cc = texel_load(base, (size_t)vci * W + uci, (T *)nullptr);
    }
    return ff * (1 - u_off) * (1 - v_off) +
           fc * (1 - u_off) * v_off +
           cf * u_off * (1 - v_off) +
           cc * u_off * v_off;
}

// mipmap.h:71-88 trilinear lookup at a real level
template <typename T>
LJ_D T mip_lookup(const DScene &sc, const lj_image &img, int image_id, double u, double v, double level) {
    int n = img.num_levels;
    if (level <= 0) {
        return mip_lookup_level<T>(sc, img, image_id, u, v, 0);
    } else if (level < (double)(n - 1)) {
        int flevel = clampi((int)floor(level), 0, n - 1);
        int clevel = clampi(flevel + 1, 0, n - 1);
        double level_off = level - flevel;
        return mip_lookup_level<T>(sc, img, image_id, u, v, flevel) * (1 - level_off) +
               mip_lookup_level<T>(sc, img, image_id, u, v, clevel) * level_off;
    } else {
        return mip_lookup_level<T>(sc, img, image_id, u, v, n - 1);
    }
}

LJ_D double tex_value(const lj_texture &t, const double *v, double *) { return v[0]; }
LJ_D d3 tex_value(const lj_texture &t, const double *v, d3 *) { return mk3(v[0], v[1], v[2]); }

// texture.h:104-154
template <typename T>
LJ_D T eval_texture_t(const DScene &sc, int tex_id, d2 uv, double footprint) {
    const lj_texture &t = sc.textures[tex_id];
    if (t.kind == LJ_TEX_CONSTANT) {
        return tex_value(t, t.value, (T *)nullptr);
    } else if (t.kind == LJ_TEX_IMAGE) {
        const lj_image &img = (sizeof(T) == sizeof(double)) ? sc.images1[t.image_id] : sc.images3[t.image_id];
        double lu = modulo_one(uv.x * t.uscale + t.uoffset);
        double lv = modulo_one(uv.y * t.vscale + t.voffset);
        int wh = img.width[0] > img.height[0] ? img.width[0] : img.height[0];
        double scaled_footprint = wh * lmax(t.uscale, t.vscale) * footprint;
        double level = log2(lmax(scaled_footprint, (double)1e-8f));
        return mip_lookup<T>(sc, img, t.image_id, lu, lv, level);
    } else {
        double lu = modulo_one(uv.x * t.uscale + t.uoffset);
        double lv = modulo_one(uv.y * t.vscale + t.voffset);
        int x = 2 * modulo_i((int)(lu * 2), 2) - 1;
        int y = 2 * modulo_i((int)(lv * 2), 2) - 1;
        if (x * y == 1) return tex_value(t, t.value, (T *)nullptr);
        return tex_value(t, t.color1, (T *)nullptr);
    }
}
LJ_D d3 eval_tex3(const DScene &sc, int tex_id, const PathVertex &v) { return eval_texture_t<d3>(sc, tex_id, v.uv, v.uv_screen_size); }
LJ_D double eval_tex1(const DScene &sc, int tex_id, const PathVertex &v) { return eval_texture_t<double>(sc, tex_id, v.uv, v.uv_screen_size); }

// ---------------- table distributions (table_dist.cpp) ----------------
// std::upper_bound(first, first+count, u) - first
LJ_D int upper_bound_d(const double *a, int count, double u) {
    int lo = 0, n = count;
    while (n > 0) {
        int half = n >> 1;
        if (!(u < __ldg(a + lo + half))) { lo += half + 1; n -= half + 1; }
        else n = half;
    }
    return lo;
}
// table_dist.cpp:27-33
LJ_D int sample_table1d(const double *cdf, int size, double u) {
    int idx = upper_bound_d(cdf, size + 1, u);
    return clampi(idx - 1, 0, size - 1);
}

// ---------------- shapes ----------------
LJ_D d3 load3(const double *p, size_t i) { return mk3(__ldg(p + 3 * i), __ldg(p + 3 * i + 1), __ldg(p + 3 * i + 2)); }
LJ_D d2 load2(const double *p, size_t i) { return mk2(__ldg(p + 2 * i), __ldg(p + 2 * i + 1)); }

struct ShadingInfo { d2 uv; Frame shading_frame; double mean_curvature, inv_uv_size; };

// shapes/triangle_mesh.inl:65-157
// the three vertices of a triangle: from its 192-byte record (DScene::tri_shade: twelve aligned 16-byte loads, issued
// together) or, without the cache, through the index triple
struct TriVerts { d3 p0, p1, p2, n0, n1, n2; d2 uv0, uv1, uv2; };
LJ_D TriVerts load_tri_verts(const DScene &sc, const lj_shape &s, int prim, bool want_normals, bool want_uvs) {
    TriVerts t;
    if (sc.tri_shade) {
        const double2 *r = reinterpret_cast<const double2 *>(sc.tri_shade + 24 * ((size_t)s.triangle_offset + prim));
        const double2 a0 = __ldg(r), a1 = __ldg(r + 1), a2 = __ldg(r + 2), a3 = __ldg(r + 3), a4 = __ldg(r + 4);
        t.p0 = mk3(a0.x, a0.y, a1.x); t.p1 = mk3(a1.y, a2.x, a2.y); t.p2 = mk3(a3.x, a3.y, a4.x);
        if (want_normals) {
            const double2 b5 = __ldg(r + 5), b6 = __ldg(r + 6), b7 = __ldg(r + 7), b8 = __ldg(r + 8);
            t.n0 = mk3(a4.y, b5.x, b5.y); t.n1 = mk3(b6.x, b6.y, b7.x); t.n2 = mk3(b7.y, b8.x, b8.y);
        }
        if (want_uvs) {
            const double2 c9 = __ldg(r + 9), c10 = __ldg(r + 10), c11 = __ldg(r + 11);
            t.uv0 = mk2(c9.x, c9.y); t.uv1 = mk2(c10.x, c10.y); t.uv2 = mk2(c11.x, c11.y);
        }
        return t;
    }
    const int *ip = sc.indices + 3 * ((size_t)s.triangle_offset + prim);
    const int i0 = __ldg(ip), i1 = __ldg(ip + 1), i2 = __ldg(ip + 2);
    const size_t vb = s.vertex_offset;
    t.p0 = load3(sc.positions, vb + i0); t.p1 = load3(sc.positions, vb + i1); t.p2 = load3(sc.positions, vb + i2);
    if (want_normals) { t.n0 = load3(sc.normals, vb + i0); t.n1 = load3(sc.normals, vb + i1); t.n2 = load3(sc.normals, vb + i2); }
    if (want_uvs) { t.uv0 = load2(sc.uvs, vb + i0); t.uv1 = load2(sc.uvs, vb + i1); t.uv2 = load2(sc.uvs, vb + i2); }
    return t;
}

LJ_D ShadingInfo shading_info_mesh(const DScene &sc, const lj_shape &s, const PathVertex &vertex) {
    const TriVerts tv = load_tri_verts(sc, s, vertex.primitive_id, s.has_normals != 0, s.has_uvs != 0);
    d2 uv0, uv1, uv2;
    if (s.has_uvs) {
        uv0 = tv.uv0; uv1 = tv.uv1; uv2 = tv.uv2;
    } else {
        uv0 = mk2(0, 0); uv1 = mk2(1, 0); uv2 = mk2(1, 1);
    }
    d2 uv = (1 - vertex.st.x - vertex.st.y) * uv0 + vertex.st.x * uv1 + vertex.st.y * uv2;
    d3 p0 = tv.p0, p1 = tv.p1, p2 = tv.p2;
    d2 duvds = uv2 - uv0;
    d2 duvdt = uv2 - uv1;
    double det = duvds.x * duvdt.y - duvdt.x * duvds.y;
    // (axis-aligned texture coordinates make most of these numerators exactly zero, which sends CUDA's f64 division into
    // its slow-path subroutine; routing them through div_z (lj_math.cuh) removed 11 % of this kernel's instructions and
    // made it 2 % SLOWER, 47.8 -> 48.7 ms per sponza step: it waits on loads, not on issue slots.  Left as plain divisions.)
    double dsdu = duvdt.y / det;
    double dtdu = -duvds.y / det;
    double dsdv = duvdt.x / det;
    double dtdv = -duvds.x / det;
    d3 dpdu, dpdv;
    if (fabs(det) > (double)1e-8f) {
        d3 dpds = p2 - p0;
        d3 dpdt = p2 - p1;
        dpdu = dpds * dsdu + dpdt * dtdu;
        dpdv = dpds * dsdv + dpdt * dtdv;
    } else {
        coordinate_system(vertex.geometry_normal, dpdu, dpdv);
    }
    d3 shading_normal = vertex.geometry_normal;
    double mean_curvature = 0;
    d3 tangent, bitangent;
    if (s.has_normals) {
        d3 n0 = tv.n0, n1 = tv.n1, n2 = tv.n2;
        shading_normal = normalize((1 - vertex.st.x - vertex.st.y) * n0 + vertex.st.x * n1 + vertex.st.y * n2);
        tangent = normalize(dpdu - shading_normal * dot(shading_normal, dpdu));
        d3 dnds = n2 - n0;
        d3 dndt = n2 - n1;
        d3 dndu = dnds * dsdu + dndt * dtdu;
        d3 dndv = dnds * dsdv + dndt * dtdv;
        bitangent = normalize(cross(shading_normal, tangent));
        mean_curvature = (dot(dndu, tangent) + dot(dndv, bitangent)) / 2.0;
    } else {
        tangent = normalize(dpdu - shading_normal * dot(shading_normal, dpdu));
        bitangent = normalize(cross(shading_normal, tangent));
    }
    ShadingInfo si;
    si.uv = uv; si.shading_frame = make_frame(tangent, bitangent, shading_normal);
    si.mean_curvature = mean_curvature;
    si.inv_uv_size = lmax(length(dpdu), length(dpdv));
    return si;
}

// shapes/sphere.inl:235-260 (note: st is fed to sin/cos as-is, as the reference does)
LJ_D ShadingInfo shading_info_sphere(const lj_shape &s, const PathVertex &vertex) {
    double r = s.radius;
    double su = sin(vertex.st.x), cu = cos(vertex.st.x), sv = sin(vertex.st.y), cv = cos(vertex.st.y);
    d3 dpdu = mk3(-r * su * sv, r * cu * sv, 0.0);
    d3 dpdv = mk3(r * cu * cv, r * su * cv, -r * sv);
    d3 tangent = normalize(dpdu - vertex.geometry_normal * dot(vertex.geometry_normal, dpdu));
    ShadingInfo si;
    si.uv = vertex.st;
    si.shading_frame = make_frame(tangent, normalize(cross(vertex.geometry_normal, tangent)), vertex.geometry_normal);
    si.mean_curvature = 1 / r;
    si.inv_uv_size = (length(dpdu) + length(dpdv)) / 2;
    return si;
}

// intersection.cpp:38-62: PathVertex from the f32 hit record and the f64 ray
LJ_D PathVertex make_vertex(const DScene &sc, const Ray &ray, const Hit &h, const RayDiff &rd) {
    PathVertex v;
    v.position = ray.org + ray.dir * (double)h.t;
    v.geometry_normal = normalize(mk3((double)h.ngx, (double)h.ngy, (double)h.ngz));
    v.shape_id = h.geom; v.primitive_id = h.prim;
    const lj_shape &s = sc.shapes[h.geom];
    v.material_id = s.material_id;
    v.interior_medium_id = s.interior_medium_id; v.exterior_medium_id = s.exterior_medium_id;
    v.st = mk2((double)h.u, (double)h.v);
    ShadingInfo si = (s.kind == LJ_SHAPE_SPHERE) ? shading_info_sphere(s, v) : shading_info_mesh(sc, s, v);
    v.shading_frame = si.shading_frame;
    v.uv = si.uv;
    v.mean_curvature = si.mean_curvature;
    v.ray_radius = rd_transfer(rd, distance(ray.org, v.position));
    v.uv_screen_size = v.ray_radius / si.inv_uv_size;
    if (dot(v.geometry_normal, v.shading_frame.n) < 0) v.geometry_normal = -v.geometry_normal;
    return v;
}

// shapes/triangle_mesh.inl:24-38 and shapes/sphere.inl:156-204
LJ_D PointAndNormal sample_point_on_shape(const DScene &sc, const lj_shape &s, d3 ref_point, d2 uv, double w) {
    PointAndNormal pn;
    if (s.kind == LJ_SHAPE_MESH) {
        int tri = sample_table1d(sc.tables + s.tri_cdf_offset, (int)s.num_triangles, w);
        const TriVerts tv = load_tri_verts(sc, s, tri, false, false);
        d3 v0 = tv.p0, v1 = tv.p1, v2 = tv.p2;
        d3 e1 = v1 - v0, e2 = v2 - v0;
        double a = sqrt(clampd(uv.x, 0.0, 1.0));
        double b1 = 1 - a;
        double b2 = a * uv.y;
        pn.position = v0 + (e1 * b1) + (e2 * b2);
        pn.normal = normalize(cross(e1, e2));
        return pn;
    }
    d3 center = mk3(s.center[0], s.center[1], s.center[2]);
    double r = s.radius;
    if (distance_squared(ref_point, center) < r * r) {
        double z = 1 - 2 * uv.x;
        double r_ = sqrt(fmax(0.0, 1 - z * z));
        double phi = 2 * LJ_PI * uv.y;
        double sp, cp; sincos(phi, &sp, &cp);
        d3 offset = mk3(r_ * cp, r_ * sp, z);
        pn.position = center + r * offset;
        pn.normal = offset;
        return pn;
    }
    d3 dir_to_center = normalize(center - ref_point);
    Frame frame = make_frame(dir_to_center);
    double sin_elevation_max_sq = r * r / distance_squared(ref_point, center);
    double cos_elevation_max = sqrt(lmax(0.0, 1 - sin_elevation_max_sq));
    double cos_elevation = (1 - uv.x) + uv.x * cos_elevation_max;
    double sin_elevation = sqrt(lmax(0.0, 1 - cos_elevation * cos_elevation));
    double azimuth = uv.y * 2 * LJ_PI;
    double dc = distance(ref_point, center);
    double ds = dc * cos_elevation - sqrt(lmax(0.0, r * r - dc * dc * sin_elevation * sin_elevation));
    double cos_alpha = (dc * dc + r * r - ds * ds) / (2 * dc * r);
    double sin_alpha = sqrt(lmax(0.0, 1 - cos_alpha * cos_alpha));
    double saz, caz; sincos(azimuth, &saz, &caz);
    d3 n_on_sphere = -to_world(frame, mk3(sin_alpha * caz, sin_alpha * saz, cos_alpha));
    pn.position = r * n_on_sphere + center;
    pn.normal = n_on_sphere;
    return pn;
}

// shapes/triangle_mesh.inl:44-46 and shapes/sphere.inl:210-230
LJ_D double pdf_point_on_shape(const lj_shape &s, const PointAndNormal &pt, d3 ref_point) {
    if (s.kind == LJ_SHAPE_MESH) return 1 / s.total_area;
    d3 center = mk3(s.center[0], s.center[1], s.center[2]);
    double r = s.radius;
    if (distance_squared(ref_point, center) < r * r) return 1 / (4 * LJ_PI * r * r);
    double sin_elevation_max_sq = r * r / distance_squared(ref_point, center);
    double cos_elevation_max = sqrt(lmax(0.0, 1 - sin_elevation_max_sq));
    double pdf_solid_angle = 1 / (2 * LJ_PI * (1 - cos_elevation_max));
    d3 dir = normalize(pt.position - ref_point);
    return pdf_solid_angle * fabs(dot(pt.normal, dir)) / distance_squared(ref_point, pt.position);
}

// ---------------- lights ----------------
// table_dist.cpp:116-139
LJ_D d2 sample_table2d(const DScene &sc, const lj_light &l, d2 rnd) {
    int w = l.dist_width, h = l.dist_height;
    const double *cdf_m = sc.tables + l.cdf_marginals_offset;
    int y_offset = clampi(upper_bound_d(cdf_m, h + 1, rnd.y) - 1, 0, h - 1);
    double m0 = __ldg(cdf_m + y_offset), m1 = __ldg(cdf_m + y_offset + 1);
    double dy = rnd.y - m0;
    if ((m1 - m0) > 0) dy /= (m1 - m0);
    const double *cdf = sc.tables + l.cdf_rows_offset + (size_t)y_offset * (w + 1);
    int x_offset = clampi(upper_bound_d(cdf, w + 1, rnd.x) - 1, 0, w - 1);
    double c0 = __ldg(cdf + x_offset), c1 = __ldg(cdf + x_offset + 1);
    double dx = rnd.x - c0;
    if (c1 - c0 > 0) dx /= (c1 - c0);
    return mk2((x_offset + dx) / w, (y_offset + dy) / h);
}
// table_dist.cpp:141-151
LJ_D double pdf_table2d(const DScene &sc, const lj_light &l, d2 xy) {
    int w = l.dist_width, h = l.dist_height;
    int x = (int)clampd(xy.x * w, 0.0, (double)(w - 1));
    int y = (int)clampd(xy.y * h, 0.0, (double)(h - 1));
    double pdf_y = __ldg(sc.tables + l.pdf_marginals_offset + y);
    double pdf_x = __ldg(sc.tables + l.pdf_rows_offset + (size_t)y * w + x);
    return pdf_y * pdf_x * w * h;
}

// lights/diffuse_area_light.inl:5-8, lights/envmap.inl:7-20
LJ_D PointAndNormal sample_point_on_light(const DScene &sc, const lj_light &l, d3 ref_point, d2 uv, double w) {
    if (l.kind == LJ_LIGHT_AREA) return sample_point_on_shape(sc, sc.shapes[l.shape_id], ref_point, uv, w);
    d2 p = sample_table2d(sc, l, uv);
    double azimuth = p.x * (2 * LJ_PI);
    double elevation = p.y * LJ_PI;
    d3 local_dir = mk3(sin(azimuth) * sin(elevation), cos(elevation), -cos(azimuth) * sin(elevation));
    d3 world_dir = xform_vector(l.to_world, local_dir);
    PointAndNormal pn; pn.position = mk3(0, 0, 0); pn.normal = -world_dir;
    return pn;
}

// lights/diffuse_area_light.inl:10-13, lights/envmap.inl:22-42
LJ_D double pdf_point_on_light(const DScene &sc, const lj_light &l, const PointAndNormal &pt, d3 ref_point) {
    if (l.kind == LJ_LIGHT_AREA) return pdf_point_on_shape(sc.shapes[l.shape_id], pt, ref_point);
    d3 world_dir = -pt.normal;
    d3 local_dir = xform_vector(l.to_local, world_dir);
    d2 uv = mk2(atan2(local_dir.x, -local_dir.z) * LJ_INVTWOPI, acos(clampd(local_dir.y, -1.0, 1.0)) * LJ_INVPI);
    if (uv.x < 0) uv.x += 1;
    double cos_elevation = local_dir.y;
    double sin_elevation = sqrt(clampd(1 - cos_elevation * cos_elevation, 0.0, 1.0));
    if (sin_elevation <= 0) return 0;
    return pdf_table2d(sc, l, uv) / (2 * LJ_PI * LJ_PI * sin_elevation);
}

// lights/diffuse_area_light.inl:15-20, lights/envmap.inl:44-73
LJ_D d3 light_emission(const DScene &sc, const lj_light &l, d3 view_dir, double view_footprint, const PointAndNormal &pt) {
    if (l.kind == LJ_LIGHT_AREA) {
        if (dot(pt.normal, view_dir) <= 0) return mk3(0, 0, 0);
        return mk3(l.intensity[0], l.intensity[1], l.intensity[2]);
    }
    d3 local_dir = xform_vector(l.to_local, -view_dir);
    d2 uv = mk2(atan2(local_dir.x, -local_dir.z) * LJ_INVTWOPI, acos(clampd(local_dir.y, -1.0, 1.0)) * LJ_INVPI);
    if (uv.x < 0) uv.x += 1;
    d3 w = local_dir;
    double dudwx = -w.z / (w.x * w.x + w.z * w.z);
    double dudwz = w.x / (w.x * w.x + w.z * w.z);
    double dvdwy = -1 / sqrt(fmax(1 - w.y * w.y, 0.0));   // std::max, envmap.inl:66
    double footprint = lmin(sqrt(dudwx * dudwx + dudwz * dudwz), dvdwy);
    return eval_texture_t<d3>(sc, l.values_texture, uv, footprint) * l.scale;
}

// intersection.cpp:87-98
LJ_D d3 vertex_emission(const DScene &sc, const PathVertex &v, d3 view_dir) {
    int light_id = sc.shapes[v.shape_id].area_light_id;
    PointAndNormal pn; pn.position = v.position; pn.normal = v.geometry_normal;
    return light_emission(sc, sc.lights[light_id], view_dir, v.uv_screen_size, pn);
}

// ---------------- camera (camera.cpp:23-47, filters/*.inl) ----------------
LJ_D d2 sample_filter(const lj_camera &cam, d2 rnd) {
    if (cam.filter_kind == LJ_FILTER_BOX) {
        double hw = cam.filter_param / 2;
        return mk2((2.0 * rnd.x - 1.0) * hw, (2.0 * rnd.y - 1.0) * hw);
    } else if (cam.filter_kind == LJ_FILTER_GAUSSIAN) {
        double r = cam.filter_param * sqrt(-2 * log(lmax(rnd.x, 1e-8)));
        return mk2(r * cos(2 * LJ_PI * rnd.y), r * sin(2 * LJ_PI * rnd.y));
    } else {
        double h = cam.filter_param / 2;
        double x = rnd.x < 0.5 ? h * (sqrt(2 * rnd.x) - 1) : h * (1 - sqrt(1 - 2 * (rnd.x - 0.5)));
        double y = rnd.y < 0.5 ? h * (sqrt(2 * rnd.y) - 1) : h * (1 - sqrt(1 - 2 * (rnd.y - 0.5)));
        return mk2(x, y);
    }
}

LJ_D Ray sample_primary(const lj_camera &cam, d2 screen_pos) {
    double px = screen_pos.x * cam.width, py = screen_pos.y * cam.height;
    double dx = px - floor(px), dy = py - floor(py);
    d2 offset = sample_filter(cam, mk2(dx, dy));
    double rx = (floor(px) + 0.5 + offset.x) / cam.width;
    double ry = (floor(py) + 0.5 + offset.y) / cam.height;
    d3 pt = xform_point(cam.sample_to_cam, mk3(rx, ry, 0.0));
    d3 dir = normalize(pt);
    Ray r;
    r.org = xform_point(cam.cam_to_world, mk3(0, 0, 0));
    r.dir = normalize(xform_vector(cam.cam_to_world, dir));
    r.tnear = 0; r.tfar = INFINITY;
    return r;
}
