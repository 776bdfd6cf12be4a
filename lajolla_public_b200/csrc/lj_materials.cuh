// BSDF evaluation / sampling / pdf, per lane, f64.
// Restates /root/reference/src/material.cpp, microfacet.h and materials/*.inl.
// Texture parameters of a material are fetched ONCE per path vertex into
// MatParams (the reference re-evaluates them inside every eval / pdf / sample
// call with identical arguments, so the values are the same).
#pragma once
#include "lj_shading.cuh"

LJ_HD d3 operator+(double s, d3 a) { return mk3(s + a.x, s + a.y, s + a.z); }

// std::pow(x, 2) in the reference: glibc returns the correctly rounded square
LJ_HD double sqr(double x) { return x * x; }

struct MatParams {
    int kind;
    double eta;
    d3 c0, c1;          // spectrum slots 0 and 1
    double s[LJ_SLOT_COUNT];   // scalar slots
};

LJ_D MatParams load_mat_params(const DScene &sc, const PathVertex &v) {
    const lj_material &m = sc.materials[v.material_id];
    MatParams p;
    p.kind = m.kind; p.eta = m.eta;
    p.c0 = mk3(0, 0, 0); p.c1 = mk3(0, 0, 0);
#pragma unroll
    for (int i = 0; i < LJ_SLOT_COUNT; i++) {
        p.s[i] = 0;
        int t = m.tex[i];
        if (t < 0) continue;
        if (sc.textures[t].channels == 3) {
            d3 c = eval_tex3(sc, t, v);
            if (i == 0) p.c0 = c; else p.c1 = c;
        } else {
            p.s[i] = eval_tex1(sc, t, v);
        }
    }
    return p;
}

struct BSDFSample { d3 dir_out; double eta, roughness; };   // material.h:127-131

// material.cpp:4-11
LJ_D d3 sample_cos_hemisphere(d2 rnd) {
    double phi = LJ_TWOPI * rnd.x;
    double tmp = sqrt(clampd(1 - rnd.y, 0.0, 1.0));
    double sp, cp; sincos(phi, &sp, &cp);      // one argument reduction for both; same values as sin() / cos()
    return mk3(cp * tmp, sp * tmp, sqrt(clampd(rnd.y, 0.0, 1.0)));
}

// ---- microfacet.h ----
LJ_D double fresnel_dielectric3(double n_dot_i, double n_dot_t, double eta) {
    double rs = (n_dot_i - eta * n_dot_t) / (n_dot_i + eta * n_dot_t);
    double rp = (eta * n_dot_i - n_dot_t) / (eta * n_dot_i + n_dot_t);
    return (rs * rs + rp * rp) / 2;
}
LJ_D double fresnel_dielectric(double n_dot_i, double eta) {
    double n_dot_t_sq = 1 - (1 - n_dot_i * n_dot_i) / (eta * eta);
    if (n_dot_t_sq < 0) return 1;
    double n_dot_t = sqrt(n_dot_t_sq);
    return fresnel_dielectric3(fabs(n_dot_i), n_dot_t, eta);
}
LJ_D double GTR2(double n_dot_h, double roughness) {
    double alpha = roughness * roughness;
    double a2 = alpha * alpha;
    double t = 1 + (a2 - 1) * n_dot_h * n_dot_h;
    return a2 / (LJ_PI * t * t);
}
LJ_D double smith_masking_gtr2(d3 v_local, double roughness) {
    double alpha = roughness * roughness;
    double a2 = alpha * alpha;
    d3 v2 = v_local * v_local;
    double Lambda = (-1 + sqrt(1 + (v2.x * a2 + v2.y * a2) / v2.z)) / 2;
    return 1 / (1 + Lambda);
}
// microfacet.h:92-114 and materials/disney_metal.inl:19-50 (ax == ay == alpha gives the isotropic one)
LJ_D d3 sample_visible_normals_aniso(d3 local_dir_in, double ax, double ay, d2 rnd) {
    bool flip = local_dir_in.z < 0;
    if (flip) local_dir_in = -local_dir_in;
    d3 hemi_dir_in = normalize(mk3(ax * local_dir_in.x, ay * local_dir_in.y, local_dir_in.z));
    double r = sqrt(rnd.x);
    double phi = 2 * LJ_PI * rnd.y;
    double sp, cp; sincos(phi, &sp, &cp);
    double t1 = r * cp;
    double t2 = r * sp;
    double s = (1 + hemi_dir_in.z) / 2;
    t2 = (1 - s) * sqrt(1 - t1 * t1) + s * t2;
    d3 disk_N = mk3(t1, t2, sqrt(lmax(0.0, 1 - t1 * t1 - t2 * t2)));
    Frame hemi_frame = make_frame(hemi_dir_in);
    d3 hemi_N = to_world(hemi_frame, disk_N);
    d3 n = normalize(mk3(ax * hemi_N.x, ay * hemi_N.y, lmax(0.0, hemi_N.z)));
    return flip ? -n : n;
}
LJ_D d3 reflect_dir(d3 dir_in, d3 h) { return normalize(-dir_in + 2 * dot(dir_in, h) * h); }

// Frame selection shared by the one-sided materials (e.g. lambertian.inl:6-9)
LJ_D Frame facing_frame(const PathVertex &v, d3 dir_in) {
    Frame f = v.shading_frame;
    if (dot(f.n, dir_in) < 0) f = neg_frame(f);
    return f;
}
// Frame selection of the two-sided dielectrics (roughdielectric.inl:6-9)
LJ_D Frame dielectric_frame(const PathVertex &v, d3 dir_in) {
    Frame f = v.shading_frame;
    if (dot(f.n, dir_in) * dot(v.geometry_normal, dir_in) < 0) f = neg_frame(f);
    return f;
}
LJ_D bool below_surface(const PathVertex &v, d3 dir_in, d3 dir_out) {
    return dot(v.geometry_normal, dir_in) < 0 || dot(v.geometry_normal, dir_out) < 0;
}

// ---------------- Lambertian (materials/lambertian.inl) ----------------
LJ_D d3 lambertian_eval(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    if (below_surface(v, in, out)) return mk3(0, 0, 0);
    Frame f = facing_frame(v, in);
    return fmax(dot(f.n, out), 0.0) * p.c0 / LJ_PI;
}
LJ_D double cosine_pdf(const PathVertex &v, d3 in, d3 out) {
    if (below_surface(v, in, out)) return 0;
    Frame f = facing_frame(v, in);
    return fmax(dot(f.n, out), 0.0) / LJ_PI;
}
LJ_D bool cosine_sample(const PathVertex &v, d3 in, d2 rnd, BSDFSample &s) {
    if (dot(v.geometry_normal, in) < 0) return false;
    Frame f = facing_frame(v, in);
    s.dir_out = to_world(f, sample_cos_hemisphere(rnd)); s.eta = 0; s.roughness = 1;
    return true;
}

// ---------------- RoughPlastic (materials/roughplastic.inl) ----------------
LJ_D d3 roughplastic_eval(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    if (below_surface(v, in, out)) return mk3(0, 0, 0);
    Frame f = facing_frame(v, in);
    d3 h = normalize(in + out);
    double n_dot_h = dot(f.n, h), n_dot_in = dot(f.n, in), n_dot_out = dot(f.n, out);
    if (n_dot_out <= 0 || n_dot_h <= 0) return mk3(0, 0, 0);
    d3 Kd = p.c0, Ks = p.c1;
    double roughness = clampd(p.s[LJ_SLOT_ROUGHNESS], 0.01, 1.0);
    double F_o = fresnel_dielectric(dot(h, out), p.eta);
    double D = GTR2(n_dot_h, roughness);
    double G = smith_masking_gtr2(to_local(f, in), roughness) * smith_masking_gtr2(to_local(f, out), roughness);
    d3 spec = Ks * (G * F_o * D) / (4 * n_dot_in * n_dot_out);
    double F_i = fresnel_dielectric(dot(h, in), p.eta);
    d3 diff = Kd * (1.0 - F_o) * (1.0 - F_i) / LJ_PI;
    return (spec + diff) * n_dot_out;
}
LJ_D double roughplastic_pdf(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    if (below_surface(v, in, out)) return 0;
    Frame f = facing_frame(v, in);
    d3 h = normalize(in + out);
    double n_dot_in = dot(f.n, in), n_dot_out = dot(f.n, out), n_dot_h = dot(f.n, h);
    if (n_dot_out <= 0 || n_dot_h <= 0) return 0;
    double lS = luminance(p.c1), lR = luminance(p.c0);
    if (lS + lR <= 0) return 0;
    double roughness = clampd(p.s[LJ_SLOT_ROUGHNESS], 0.01, 1.0);
    double spec_prob = lS / (lS + lR);
    double diff_prob = 1 - spec_prob;
    double G = smith_masking_gtr2(to_local(f, in), roughness);
    double D = GTR2(n_dot_h, roughness);
    spec_prob *= (G * D) / (4 * n_dot_in);
    diff_prob *= n_dot_out / LJ_PI;
    return spec_prob + diff_prob;
}
LJ_D bool roughplastic_sample(const MatParams &p, const PathVertex &v, d3 in, d2 rnd, double w, BSDFSample &s) {
    if (dot(v.geometry_normal, in) < 0) return false;
    Frame f = facing_frame(v, in);
    double lS = luminance(p.c1), lR = luminance(p.c0);
    if (lS + lR <= 0) return false;
    double spec_prob = lS / (lS + lR);
    if (w < spec_prob) {
        d3 local_in = to_local(f, in);
        double roughness = clampd(p.s[LJ_SLOT_ROUGHNESS], 0.01, 1.0);
        double alpha = roughness * roughness;
        d3 h = to_world(f, sample_visible_normals_aniso(local_in, alpha, alpha, rnd));
        s.dir_out = reflect_dir(in, h); s.eta = 0; s.roughness = roughness;
    } else {
        s.dir_out = to_world(f, sample_cos_hemisphere(rnd)); s.eta = 0; s.roughness = 1;
    }
    return true;
}

// ---------------- RoughDielectric (materials/roughdielectric.inl) ----------------
LJ_D d3 dielectric_half_vector(const Frame &f, d3 in, d3 out, bool reflect, double eta) {
    d3 h = reflect ? normalize(in + out) : normalize(in + out * eta);
    if (dot(h, f.n) < 0) h = -h;
    return h;
}
LJ_D d3 roughdielectric_eval(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    bool reflect = dot(v.geometry_normal, in) * dot(v.geometry_normal, out) > 0;
    Frame f = dielectric_frame(v, in);
    double eta = dot(v.geometry_normal, in) > 0 ? p.eta : 1 / p.eta;
    d3 Ks = p.c0, Kt = p.c1;
    d3 h = dielectric_half_vector(f, in, out, reflect, eta);
    double roughness = clampd(p.s[LJ_SLOT_ROUGHNESS], 0.01, 1.0);
    double h_dot_in = dot(h, in);
    double F = fresnel_dielectric(h_dot_in, eta);
    double D = GTR2(dot(f.n, h), roughness);
    double G = smith_masking_gtr2(to_local(f, in), roughness) * smith_masking_gtr2(to_local(f, out), roughness);
    if (reflect) return Ks * (F * D * G) / (4 * fabs(dot(f.n, in)));
    double eta_factor = 1 / (eta * eta);   // TransportDirection::TO_LIGHT (the only one the integrators use)
    double h_dot_out = dot(h, out);
    double sqrt_denom = h_dot_in + eta * h_dot_out;
    return Kt * (eta_factor * (1 - F) * D * G * eta * eta * fabs(h_dot_out * h_dot_in)) /
           (fabs(dot(f.n, in)) * sqrt_denom * sqrt_denom);
}
LJ_D double roughdielectric_pdf(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    bool reflect = dot(v.geometry_normal, in) * dot(v.geometry_normal, out) > 0;
    Frame f = dielectric_frame(v, in);
    double eta = dot(v.geometry_normal, in) > 0 ? p.eta : 1 / p.eta;
    d3 h = dielectric_half_vector(f, in, out, reflect, eta);
    double roughness = clampd(p.s[LJ_SLOT_ROUGHNESS], 0.01, 1.0);
    double h_dot_in = dot(h, in);
    double F = fresnel_dielectric(h_dot_in, eta);
    double D = GTR2(dot(h, f.n), roughness);
    double G_in = smith_masking_gtr2(to_local(f, in), roughness);
    if (reflect) return (F * D * G_in) / (4 * fabs(dot(f.n, in)));
    double h_dot_out = dot(h, out);
    double sqrt_denom = h_dot_in + eta * h_dot_out;
    double dh_dout = eta * eta * h_dot_out / (sqrt_denom * sqrt_denom);
    return (1 - F) * D * G_in * fabs(dh_dout * h_dot_in / dot(f.n, in));
}
// shared tail of roughdielectric.inl:143-163 / disney_glass.inl / disney_bsdf.inl glass lobe
LJ_D bool dielectric_scatter(d3 in, d3 h, double eta, double w, double roughness, BSDFSample &s) {
    double h_dot_in = dot(h, in);
    double F = fresnel_dielectric(h_dot_in, eta);
    if (w <= F) {
        s.dir_out = reflect_dir(in, h); s.eta = 0; s.roughness = roughness;
        return true;
    }
    double h_dot_out_sq = 1 - (1 - h_dot_in * h_dot_in) / (eta * eta);
    if (h_dot_out_sq <= 0) return false;
    if (h_dot_in < 0) h = -h;
    double h_dot_out = sqrt(h_dot_out_sq);
    s.dir_out = -in / eta + (fabs(h_dot_in) / eta - h_dot_out) * h;
    s.eta = eta; s.roughness = roughness;
    return true;
}
LJ_D bool roughdielectric_sample(const MatParams &p, const PathVertex &v, d3 in, d2 rnd, double w, BSDFSample &s) {
    double eta = dot(v.geometry_normal, in) > 0 ? p.eta : 1 / p.eta;
    Frame f = dielectric_frame(v, in);
    double roughness = clampd(p.s[LJ_SLOT_ROUGHNESS], 0.01, 1.0);
    double alpha = roughness * roughness;
    d3 h = to_world(f, sample_visible_normals_aniso(to_local(f, in), alpha, alpha, rnd));
    if (dot(h, f.n) < 0) h = -h;
    return dielectric_scatter(in, h, eta, w, roughness, s);
}

// ---------------- Disney lobes (materials/disney_*.inl) ----------------
// disney_metal.inl:3-17
LJ_D double smithG_GGX_aniso(double NdotW, double WdotX, double WdotY, double ax, double ay) {
    double lambda = 0.5 * (sqrt(1.0 + (sqr(WdotX * ax) + sqr(WdotY * ay)) / sqr(NdotW)) - 1.0);
    return 1.0 / (1.0 + lambda);
}
LJ_D double GTR2_aniso(double ax, double ay, const Frame &f, d3 h) {
    double ax2 = ax * ax, ay2 = ay * ay;
    double hlx2 = sqr(dot(f.x, h)), hly2 = sqr(dot(f.y, h)), hlz2 = sqr(dot(f.n, h));
    return 1.0 / (LJ_PI * ax * ay * sqr(hlx2 / ax2 + hly2 / ay2 + hlz2));
}
LJ_D void aniso_alphas(double roughness_clamped, double anisotropic, double &ax, double &ay) {
    double aspect = sqrt(1.0 - 0.9 * anisotropic);
    double a_min = 0.0001;
    ax = fmax(a_min, roughness_clamped * roughness_clamped / aspect);
    ay = fmax(a_min, roughness_clamped * roughness_clamped * aspect);
}
// disney_clearcoat.inl:3-16
LJ_D double clearcoat_fresnel(d3 h, d3 out) {
    double eta = 1.5;
    double R_0 = sqr(eta - 1.0) / sqr(eta + 1.0);
    return R_0 + (1.0 - R_0) * pow(1.0 - fabs(dot(h, out)), 5.0);
}
LJ_D double compute_Dc(double clearcoat_gloss, double hlz2) {
    double a = (1.0 - clearcoat_gloss) * 0.1 + clearcoat_gloss * 0.001;
    double a2 = a * a;
    return (a2 - 1.0) / (LJ_PI * log(a2) * (1.0 + (a2 - 1.0) * hlz2));
}

// diffuse lobe body (disney_diffuse.inl:14-36, disney_bsdf.inl:27-49); roughness is NOT clamped here
LJ_D d3 disney_diffuse_lobe(const Frame &f, d3 base_color, double roughness, double subsurface, d3 in, d3 out) {
    d3 h = normalize(in + out);
    double h_dot_out = dot(h, out), n_dot_in = dot(f.n, in), n_dot_out = dot(f.n, out);
    double FD90 = 0.5 + 2.0 * roughness * h_dot_out * h_dot_out;
    double FD_in = 1.0 + (FD90 - 1.0) * (1.0 - pow(n_dot_in, 5.0));
    double FD_out = 1.0 + (FD90 - 1.0) * (1.0 - pow(n_dot_out, 5.0));
    d3 f_d = base_color * FD_in * FD_out * fabs(n_dot_out) / LJ_PI;
    double FSS90 = roughness * h_dot_out * h_dot_out;
    double FSS_in = 1.0 + (FSS90 - 1.0) * (1.0 - pow(n_dot_in, 5.0));
    double FSS_out = 1.0 + (FSS90 - 1.0) * (1.0 - pow(n_dot_out, 5.0));
    d3 f_ss = 1.25 * base_color * (FSS_in * FSS_out * (1.0 / (fabs(n_dot_in) + fabs(n_dot_out)) - 0.5) + 0.5)
              * fabs(n_dot_out) / LJ_PI;
    return (1.0 - subsurface) * f_d + subsurface * f_ss;
}
// metal lobe given the Fresnel colour (disney_metal.inl:66-80, disney_bsdf.inl:77-104)
LJ_D d3 disney_metal_lobe(const Frame &f, d3 Fm, double roughness_raw, double anisotropic, d3 in, d3 out) {
    double roughness = clampd(roughness_raw, 0.01, 1.0);
    d3 h = normalize(in + out);
    double ax, ay; aniso_alphas(roughness, anisotropic, ax, ay);
    double Dm = GTR2_aniso(ax, ay, f, h);
    double Gin = smithG_GGX_aniso(dot(in, f.n), dot(in, f.x), dot(in, f.y), ax, ay);
    double Gout = smithG_GGX_aniso(dot(out, f.n), dot(out, f.x), dot(out, f.y), ax, ay);
    return Fm * Dm * Gin * Gout / (4.0 * fabs(dot(in, f.n)));
}
LJ_D double disney_metal_pdf_lobe(const Frame &f, double roughness_raw, double anisotropic, d3 in, d3 out) {
    double roughness = clampd(roughness_raw, 0.01, 1.0);
    d3 h = normalize(in + out);
    double ax, ay; aniso_alphas(roughness, anisotropic, ax, ay);
    double Dm = GTR2_aniso(ax, ay, f, h);
    double Gin = smithG_GGX_aniso(dot(in, f.n), dot(in, f.x), dot(in, f.y), ax, ay);
    return Dm * Gin / (4.0 * fabs(dot(in, f.n)));
}
LJ_D void disney_metal_sample_lobe(const Frame &f, double roughness_raw, double anisotropic, d3 in, d2 rnd, BSDFSample &s) {
    double roughness = clampd(roughness_raw, 0.01, 1.0);
    d3 local_in = to_local(f, in);
    double ax, ay; aniso_alphas(roughness, anisotropic, ax, ay);
    d3 h = to_world(f, sample_visible_normals_aniso(local_in, ax, ay, rnd));
    s.dir_out = reflect_dir(in, h); s.eta = 0; s.roughness = roughness;
}
// clearcoat lobe (disney_clearcoat.inl:27-44); returns false when n.h <= 0
LJ_D bool disney_clearcoat_lobe(const Frame &f, double gloss, d3 in, d3 out, double &val) {
    d3 h = normalize(in + out);
    double n_dot_h = dot(f.n, h), n_dot_in = dot(f.n, in);
    if (n_dot_h <= 0) return false;
    double F = clearcoat_fresnel(h, out);
    double D = compute_Dc(gloss, sqr(dot(f.n, h)));
    double G = smith_masking_gtr2(to_local(f, in), 0.5) * smith_masking_gtr2(to_local(f, out), 0.5);
    val = F * D * G / (4.0 * fabs(n_dot_in));
    return true;
}
LJ_D double disney_clearcoat_pdf_lobe(const Frame &f, double gloss, d3 in, d3 out) {
    d3 h = normalize(in + out);
    double n_dot_h = dot(f.n, h);
    double D = compute_Dc(gloss, sqr(dot(f.n, h)));
    return D * fabs(n_dot_h) / (4.0 * fabs(dot(h, out)));
}
LJ_D void disney_clearcoat_sample_lobe(const Frame &f, double gloss, d3 in, d2 rnd, BSDFSample &s) {
    double a = (1.0 - gloss) * 0.1 + gloss * 0.001;
    double a2 = a * a;
    double cos_h_elevation = sqrt((1.0 - pow(a2, 1.0 - rnd.x)) / (1.0 - a2));
    double h_elevation = acos(cos_h_elevation);
    double h_azimuth = 2.0 * LJ_PI * rnd.y;
    double hlx = sin(h_elevation) * cos(h_azimuth);
    double hly = sin(h_elevation) * sin(h_azimuth);
    double hlz = cos(h_elevation);
    d3 h = to_world(f, normalize(mk3(hlx, hly, hlz)));
    s.dir_out = reflect_dir(in, h); s.eta = 0; s.roughness = 1;
}
// sheen lobe (disney_sheen.inl:13-21)
LJ_D d3 disney_sheen_lobe(const Frame &f, d3 base_color, double sheen_tint, d3 in, d3 out) {
    d3 h = normalize(in + out);
    double n_dot_out = dot(f.n, out);
    d3 C_tint = base_color / luminance(base_color);
    if (luminance(base_color) <= 0) C_tint = mk3(1.0);
    d3 C_sheen = (1.0 - sheen_tint) + sheen_tint * C_tint;
    return C_sheen * pow(1.0 - fabs(dot(h, out)), 5.0) * fabs(n_dot_out);
}
// glass lobe (disney_glass.inl:3-50, disney_bsdf.inl:139-176)
LJ_D d3 disney_glass_lobe(const PathVertex &v, d3 base_color, double roughness_raw, double anisotropic, double bsdf_eta, d3 in, d3 out) {
    bool reflect = dot(v.geometry_normal, in) * dot(v.geometry_normal, out) > 0;
    Frame f = dielectric_frame(v, in);
    double eta = dot(v.geometry_normal, in) > 0 ? bsdf_eta : 1 / bsdf_eta;
    d3 h = dielectric_half_vector(f, in, out, reflect, eta);
    double roughness = clampd(roughness_raw, 0.01, 1.0);
    double ax, ay; aniso_alphas(roughness, anisotropic, ax, ay);
    double h_dot_in = dot(h, in);
    double F = fresnel_dielectric(h_dot_in, eta);
    double D = GTR2_aniso(ax, ay, f, h);
    double G = smithG_GGX_aniso(dot(in, f.n), dot(in, f.x), dot(in, f.y), ax, ay);
    if (reflect) return base_color * (F * D * G) / (4 * fabs(dot(f.n, in)));
    double h_dot_out = dot(h, out);
    return sqrt3(base_color) * (1.0 - F) * D * G * fabs(h_dot_out * h_dot_in) /
           (fabs(dot(f.n, in)) * sqr(h_dot_in + eta * h_dot_out));
}
LJ_D double disney_glass_pdf_lobe(const PathVertex &v, double roughness_raw, double anisotropic, double bsdf_eta, d3 in, d3 out) {
    bool reflect = dot(v.geometry_normal, in) * dot(v.geometry_normal, out) > 0;
    Frame f = dielectric_frame(v, in);
    double eta = dot(v.geometry_normal, in) > 0 ? bsdf_eta : 1 / bsdf_eta;
    double roughness = clampd(roughness_raw, 0.01, 1.0);
    d3 h = dielectric_half_vector(f, in, out, reflect, eta);
    double ax, ay; aniso_alphas(roughness, anisotropic, ax, ay);
    double h_dot_in = dot(h, in);
    double F = fresnel_dielectric(h_dot_in, eta);
    double D = GTR2_aniso(ax, ay, f, h);
    double G_in = smithG_GGX_aniso(dot(in, f.n), dot(in, f.x), dot(in, f.y), ax, ay);
    if (reflect) return (F * D * G_in) / (4 * fabs(dot(f.n, in)));
    double h_dot_out = dot(h, out);
    double sqrt_denom = h_dot_in + eta * h_dot_out;
    double dh_dout = eta * eta * h_dot_out / (sqrt_denom * sqrt_denom);
    return (1 - F) * D * G_in * fabs(dh_dout * h_dot_in / dot(f.n, in));
}
LJ_D bool disney_glass_sample_lobe(const PathVertex &v, double roughness_raw, double anisotropic, double bsdf_eta,
                                   d3 in, d2 rnd, double w, BSDFSample &s) {
    Frame f = dielectric_frame(v, in);
    double eta = dot(v.geometry_normal, in) > 0 ? bsdf_eta : 1 / bsdf_eta;
    double roughness = clampd(roughness_raw, 0.01, 1.0);
    double ax, ay; aniso_alphas(roughness, anisotropic, ax, ay);
    d3 h = to_world(f, sample_visible_normals_aniso(to_local(f, in), ax, ay, rnd));
    if (dot(h, f.n) < 0) h = -h;
    return dielectric_scatter(in, h, eta, w, roughness, s);
}

// disney_bsdf.inl:3-215
LJ_D d3 disney_bsdf_eval(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    d3 base_color = p.c0;
    double specular_transmission = p.s[LJ_SLOT_SPEC_TRANSMISSION], metallic = p.s[LJ_SLOT_METALLIC];
    double subsurface = p.s[LJ_SLOT_SUBSURFACE], specular = p.s[LJ_SLOT_SPECULAR_D], roughness = p.s[LJ_SLOT_ROUGHNESS];
    double specular_tint = p.s[LJ_SLOT_SPECULAR_TINT], anisotropic = p.s[LJ_SLOT_ANISOTROPIC];
    double sheen = p.s[LJ_SLOT_SHEEN], sheen_tint = p.s[LJ_SLOT_SHEEN_TINT], clearcoat = p.s[LJ_SLOT_CLEARCOAT];
    d3 f_diffuse = mk3(0.0), f_metal = mk3(0.0), f_clearcoat = mk3(0.0), f_sheen = mk3(0.0);
    if (dot(v.geometry_normal, in) >= 0 && dot(v.geometry_normal, out) >= 0) {
        Frame f = facing_frame(v, in);
        f_diffuse = disney_diffuse_lobe(f, base_color, roughness, subsurface, in, out);
        {
            d3 h = normalize(in + out);
            double h_dot_out = dot(h, out);
            d3 C_tint = base_color / luminance(base_color);
            if (luminance(base_color) <= 0) C_tint = mk3(1.0);
            double eta = 1.5;
            double R_0 = sqr(eta - 1.0) / sqr(eta + 1.0);
            d3 Ks = (1.0 - specular_tint) + specular_tint * C_tint;
            d3 C0 = specular * R_0 * (1.0 - metallic) * Ks + metallic * base_color;
            d3 Fm = C0 + (1.0 - C0) * pow(1.0 - h_dot_out, 5.0);   // no fabs here (disney_bsdf.inl:91)
            f_metal = disney_metal_lobe(f, Fm, roughness, anisotropic, in, out);
        }
        double cc;
        if (disney_clearcoat_lobe(f, p.s[LJ_SLOT_CLEARCOAT_GLOSS], in, out, cc)) f_clearcoat = mk3(cc);
        f_sheen = disney_sheen_lobe(f, base_color, sheen_tint, in, out);
    }
    d3 f_glass = disney_glass_lobe(v, base_color, roughness, anisotropic, p.eta, in, out);
    if (dot(v.geometry_normal, in) < 0) { f_diffuse = mk3(0.0); f_metal = mk3(0.0); f_sheen = mk3(0.0); f_clearcoat = mk3(0.0); }
    return (1.0 - specular_transmission) * (1.0 - metallic) * f_diffuse
         + (1.0 - metallic) * sheen * f_sheen
         + (1.0 - specular_transmission * (1.0 - metallic)) * f_metal
         + 0.25 * clearcoat * f_clearcoat
         + (1.0 - metallic) * specular_transmission * f_glass;
}

struct DisneyWeights { double diffuse, metal, glass, clearcoat; bool dead; };
// disney_bsdf.inl:228-252 / :389-428
LJ_D DisneyWeights disney_weights(const MatParams &p, const PathVertex &v, d3 in) {
    double specular_transmission = p.s[LJ_SLOT_SPEC_TRANSMISSION], metallic = p.s[LJ_SLOT_METALLIC], clearcoat = p.s[LJ_SLOT_CLEARCOAT];
    DisneyWeights w; w.dead = false;
    w.diffuse = (1.0 - metallic) * (1.0 - specular_transmission);
    w.metal = (1.0 - specular_transmission * (1.0 - metallic));
    w.glass = (1.0 - metallic) * specular_transmission;
    w.clearcoat = 0.25 * clearcoat;
    if (dot(v.geometry_normal, in) < 0) {
        w.diffuse = 0; w.metal = 0; w.clearcoat = 0;
        if (w.glass > 0) w.glass = 1; else { w.dead = true; return w; }
    }
    double total = w.diffuse + w.metal + w.glass + w.clearcoat;
    w.diffuse /= total; w.metal /= total; w.glass /= total; w.clearcoat /= total;
    return w;
}
LJ_D double disney_bsdf_pdf(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    bool reflect = dot(v.geometry_normal, in) * dot(v.geometry_normal, out) > 0;
    DisneyWeights w = disney_weights(p, v, in);
    if (w.dead) return 0;
    double roughness = p.s[LJ_SLOT_ROUGHNESS], anisotropic = p.s[LJ_SLOT_ANISOTROPIC];
    double glass_pdf = disney_glass_pdf_lobe(v, roughness, anisotropic, p.eta, in, out);
    if (!reflect) return w.glass * glass_pdf;
    Frame f = facing_frame(v, in);
    double diffuse_pdf = fmax(dot(f.n, out), 0.0) / LJ_PI;
    double metal_pdf = disney_metal_pdf_lobe(f, roughness, anisotropic, in, out);
    double clearcoat_pdf = disney_clearcoat_pdf_lobe(f, p.s[LJ_SLOT_CLEARCOAT_GLOSS], in, out);
    return w.diffuse * diffuse_pdf + w.metal * metal_pdf + w.clearcoat * clearcoat_pdf + w.glass * glass_pdf;
}
LJ_D bool disney_bsdf_sample(const MatParams &p, const PathVertex &v, d3 in, d2 rnd, double rand, BSDFSample &s) {
    DisneyWeights w = disney_weights(p, v, in);
    if (w.dead) { s.dir_out = mk3(0, 0, 0); s.eta = 0; s.roughness = 1; return true; }   // disney_bsdf.inl:417-421: not a failure
    double roughness = p.s[LJ_SLOT_ROUGHNESS], anisotropic = p.s[LJ_SLOT_ANISOTROPIC];
    if (rand < w.diffuse) {
        Frame f = facing_frame(v, in);
        s.dir_out = to_world(f, sample_cos_hemisphere(rnd)); s.eta = 0; s.roughness = 1;
        return true;
    } else if (rand < w.diffuse + w.metal) {
        disney_metal_sample_lobe(facing_frame(v, in), roughness, anisotropic, in, rnd, s);
        return true;
    } else if (rand < w.diffuse + w.metal + w.glass) {
        double rand_new = (rand - (w.diffuse + w.metal)) / w.glass;
        return disney_glass_sample_lobe(v, roughness, anisotropic, p.eta, in, rnd, rand_new, s);
    }
    disney_clearcoat_sample_lobe(facing_frame(v, in), p.s[LJ_SLOT_CLEARCOAT_GLOSS], in, rnd, s);
    return true;
}

// ---------------- dispatch (material.cpp:90-119) ----------------
LJ_D d3 bsdf_eval(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    switch (p.kind) {
        case LJ_MAT_LAMBERTIAN: return lambertian_eval(p, v, in, out);
        case LJ_MAT_ROUGHPLASTIC: return roughplastic_eval(p, v, in, out);
        case LJ_MAT_ROUGHDIELECTRIC: return roughdielectric_eval(p, v, in, out);
        case LJ_MAT_DISNEY_DIFFUSE:
            if (below_surface(v, in, out)) return mk3(0.0);
            return disney_diffuse_lobe(facing_frame(v, in), p.c0, p.s[LJ_SLOT_ROUGHNESS], p.s[LJ_SLOT_SUBSURFACE], in, out);
        case LJ_MAT_DISNEY_METAL: {
            if (below_surface(v, in, out)) return mk3(0.0);
            d3 h = normalize(in + out);
            d3 Fm = p.c0 + (1.0 - p.c0) * pow(1.0 - fabs(dot(h, out)), 5.0);
            return disney_metal_lobe(facing_frame(v, in), Fm, p.s[LJ_SLOT_ROUGHNESS], p.s[LJ_SLOT_ANISOTROPIC], in, out);
        }
        case LJ_MAT_DISNEY_GLASS: return disney_glass_lobe(v, p.c0, p.s[LJ_SLOT_ROUGHNESS], p.s[LJ_SLOT_ANISOTROPIC], p.eta, in, out);
        case LJ_MAT_DISNEY_CLEARCOAT: {
            if (below_surface(v, in, out)) return mk3(0.0);
            double cc;
            if (!disney_clearcoat_lobe(facing_frame(v, in), p.s[LJ_SLOT_CLEARCOAT_GLOSS], in, out, cc)) return mk3(0.0);
            return mk3(cc);
        }
        case LJ_MAT_DISNEY_SHEEN:
            if (below_surface(v, in, out)) return mk3(0.0);
            return disney_sheen_lobe(facing_frame(v, in), p.c0, p.s[LJ_SLOT_SHEEN_TINT], in, out);
        default: return disney_bsdf_eval(p, v, in, out);
    }
}
LJ_D double bsdf_pdf(const MatParams &p, const PathVertex &v, d3 in, d3 out) {
    switch (p.kind) {
        case LJ_MAT_LAMBERTIAN: case LJ_MAT_DISNEY_DIFFUSE: case LJ_MAT_DISNEY_SHEEN: return cosine_pdf(v, in, out);
        case LJ_MAT_ROUGHPLASTIC: return roughplastic_pdf(p, v, in, out);
        case LJ_MAT_ROUGHDIELECTRIC: return roughdielectric_pdf(p, v, in, out);
        case LJ_MAT_DISNEY_METAL:
            if (below_surface(v, in, out)) return 0;
            return disney_metal_pdf_lobe(facing_frame(v, in), p.s[LJ_SLOT_ROUGHNESS], p.s[LJ_SLOT_ANISOTROPIC], in, out);
        case LJ_MAT_DISNEY_GLASS: return disney_glass_pdf_lobe(v, p.s[LJ_SLOT_ROUGHNESS], p.s[LJ_SLOT_ANISOTROPIC], p.eta, in, out);
        case LJ_MAT_DISNEY_CLEARCOAT:
            if (below_surface(v, in, out)) return 0;
            return disney_clearcoat_pdf_lobe(facing_frame(v, in), p.s[LJ_SLOT_CLEARCOAT_GLOSS], in, out);
        default: return disney_bsdf_pdf(p, v, in, out);
    }
}
LJ_D bool bsdf_sample(const MatParams &p, const PathVertex &v, d3 in, d2 rnd, double w, BSDFSample &s) {
    switch (p.kind) {
        case LJ_MAT_LAMBERTIAN: case LJ_MAT_DISNEY_DIFFUSE: case LJ_MAT_DISNEY_SHEEN: return cosine_sample(v, in, rnd, s);
        case LJ_MAT_ROUGHPLASTIC: return roughplastic_sample(p, v, in, rnd, w, s);
        case LJ_MAT_ROUGHDIELECTRIC: return roughdielectric_sample(p, v, in, rnd, w, s);
        case LJ_MAT_DISNEY_METAL:
            if (dot(v.geometry_normal, in) < 0) return false;
            disney_metal_sample_lobe(facing_frame(v, in), p.s[LJ_SLOT_ROUGHNESS], p.s[LJ_SLOT_ANISOTROPIC], in, rnd, s);
            return true;
        case LJ_MAT_DISNEY_GLASS: return disney_glass_sample_lobe(v, p.s[LJ_SLOT_ROUGHNESS], p.s[LJ_SLOT_ANISOTROPIC], p.eta, in, rnd, w, s);
        case LJ_MAT_DISNEY_CLEARCOAT:
            if (dot(v.geometry_normal, in) < 0) return false;
            disney_clearcoat_sample_lobe(facing_frame(v, in), p.s[LJ_SLOT_CLEARCOAT_GLOSS], in, rnd, s);
            return true;
        default: return disney_bsdf_sample(p, v, in, rnd, w, s);
    }
}
