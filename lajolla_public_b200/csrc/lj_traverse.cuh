// Ray casting on the device: what Embree did for the reference
// (rtcIntersect1 / rtcOccluded1, /root/reference/src/intersection.cpp:32,:83).
//
// Rules (DESIGN.md "Ray-cast seam"), all on the f32-rounded ray:
//   * triangle: Pluecker edge test with fixed FMA placement; accepted iff tnear < t <= tfar;
//     barycentrics (u,v) weight vertices 1 and 2 (shapes/triangle_mesh.inl:81-83);
//   * sphere: the reference's own quadratic in f64 on the f32 ray (shapes/sphere.inl:15-101),
//     accepted iff tnear <= t < tfar, reported t = (float)t;
//   * closest hit = lexicographic minimum of (t, geomID, primID); boxes only cull.
#pragma once
#include "lj_scene.cuh"

struct RayF { float ox, oy, oz, tnear, dx, dy, dz, tfar; };

LJ_D RayF to_rayf(const Ray &r) {
    RayF f;
    f.ox = (float)r.org.x; f.oy = (float)r.org.y; f.oz = (float)r.org.z; f.tnear = (float)r.tnear;
    f.dx = (float)r.dir.x; f.dy = (float)r.dir.y; f.dz = (float)r.dir.z; f.tfar = (float)r.tfar;
    return f;
}

struct f3 { float x, y, z; };
LJ_D f3 f3sub(f3 a, f3 b) { f3 r; r.x = __fsub_rn(a.x, b.x); r.y = __fsub_rn(a.y, b.y); r.z = __fsub_rn(a.z, b.z); return r; }
LJ_D f3 f3add(f3 a, f3 b) { f3 r; r.x = __fadd_rn(a.x, b.x); r.y = __fadd_rn(a.y, b.y); r.z = __fadd_rn(a.z, b.z); return r; }
// a.y*b.z - a.z*b.y as fma(a.y, b.z, -(a.z*b.y)) etc.: odd in both arguments
LJ_D f3 f3cross(f3 a, f3 b) {
    f3 r;
    r.x = __fmaf_rn(a.y, b.z, -__fmul_rn(a.z, b.y));
    r.y = __fmaf_rn(a.z, b.x, -__fmul_rn(a.x, b.z));
    r.z = __fmaf_rn(a.x, b.y, -__fmul_rn(a.y, b.x));
    return r;
}
LJ_D float f3dot(f3 a, f3 b) { return __fmaf_rn(a.x, b.x, __fmaf_rn(a.y, b.y, __fmul_rn(a.z, b.z))); }

struct TriHit { float t, u, v; f3 ng; };

LJ_D bool tri_test(f3 O, f3 D, float tnear, float tfar, f3 p0, f3 p1, f3 p2, TriHit &h) {
    f3 v0 = f3sub(p0, O), v1 = f3sub(p1, O), v2 = f3sub(p2, O);
    f3 e0 = f3sub(v2, v0), e1 = f3sub(v0, v1), e2 = f3sub(v1, v2);
    float U = f3dot(f3cross(e0, f3add(v2, v0)), D);
    float V = f3dot(f3cross(e1, f3add(v0, v1)), D);
    float W = f3dot(f3cross(e2, f3add(v1, v2)), D);
    float UVW = __fadd_rn(__fadd_rn(U, V), W);
    float eps = __fmul_rn(1.1920929e-07f, fabsf(UVW));
    float mn = fminf(U, fminf(V, W));
    float mx = fmaxf(U, fmaxf(V, W));
    if (!(mn >= -eps || mx <= eps)) return false;
    if (UVW == 0.f) return false;
    f3 ng = f3cross(e0, e1);
    float den = f3dot(ng, D);
    if (den == 0.f) return false;
    float t = __fdiv_rn(f3dot(v0, ng), den);
    if (!(t > tnear && t <= tfar)) return false;
    h.t = t; h.u = __fdiv_rn(U, UVW); h.v = __fdiv_rn(V, UVW); h.ng = ng;
    return true;
}

// shapes/sphere.inl:15-38
LJ_D bool solve_quadratic(double a, double b, double c, double *t0, double *t1) {
    if (a == 0) {
        if (b == 0) return false;
        *t0 = *t1 = -c / b;
        return true;
    }
    double disc = b * b - 4 * a * c;
    if (disc < 0) return false;
    double rd = sqrt(disc);
    if (b >= 0) { *t0 = (-b - rd) / (2 * a); *t1 = 2 * c / (-b - rd); }
    else { *t0 = 2 * c / (-b + rd); *t1 = (-b + rd) / (2 * a); }
    return true;
}

// shapes/sphere.inl:40-80 : returns the accepted t (f64) or -1
LJ_D double sphere_test(const RayF &r, const SpherePrim &s, float tfar_f) {
    d3 org = mk3((double)r.ox, (double)r.oy, (double)r.oz);
    d3 dir = mk3((double)r.dx, (double)r.dy, (double)r.dz);
    double tnear = (double)r.tnear, tfar = (double)tfar_f;
    d3 v = org - mk3(s.cx, s.cy, s.cz);
    double A = dot(dir, dir);
    double B = 2 * dot(dir, v);
    double C = dot(v, v) - s.r * s.r;
    double t0, t1;
    if (!solve_quadratic(A, B, C, &t0, &t1)) return -1;
    double t = -1;
    if (t0 >= tnear && t0 < tfar) t = t0;
    if (t1 >= tnear && t1 < tfar && t < 0) t = t1;
    if (t >= tnear && t < tfar) return t;
    return -1;
}

// shapes/sphere.inl:81-99 : fill Ng, u, v for an accepted sphere hit
LJ_D void sphere_hit_details(const RayF &r, const SpherePrim &s, double t, Hit &h) {
    d3 org = mk3((double)r.ox, (double)r.oy, (double)r.oz);
    d3 dir = mk3((double)r.dx, (double)r.dy, (double)r.dz);
    d3 p = org + t * dir;
    d3 gn = p - mk3(s.cx, s.cy, s.cz);
    h.ngx = (float)gn.x; h.ngy = (float)gn.y; h.ngz = (float)gn.z;
    d3 cart = gn / s.r;
    double elevation = acos(clampd(cart.y, -1.0, 1.0));
    double azimuth = atan2(cart.z, cart.x);
    h.u = (float)(azimuth / LJ_TWOPI);
    h.v = (float)(elevation / LJ_PI);
}

struct Best {
    float t; int geom, prim; float u, v; f3 ng; int sphere; double ts;
};

LJ_D bool hit_better(float t, int g, int p, const Best &b) {
    if (b.geom < 0) return true;
    if (t != b.t) return t < b.t;
    if (g != b.geom) return g < b.geom;
    return p < b.prim;
}

// Padded slab test; returns entry distance, or +inf on a miss.
LJ_D float slab_test(float lox, float loy, float loz, float hix, float hiy, float hiz,
                     const RayF &r, float idx, float idy, float idz, float tmax) {
    float ax = (lox - r.ox) * idx, bx = (hix - r.ox) * idx;
    float ay = (loy - r.oy) * idy, by = (hiy - r.oy) * idy;
    float az = (loz - r.oz) * idz, bz = (hiz - r.oz) * idz;
    float t0 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), r.tnear));
    float t1 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tmax));
    float enter = t0 - fabsf(t0) * 4e-7f;
    float exit_ = t1 + fabsf(t1) * 4e-7f;
    return (enter <= exit_) ? enter : __int_as_float(0x7f800000);
}

// Test primitive slot `slot` (BVH leaf order) against the ray.
template <bool ANY>
LJ_D bool test_prim(const DScene &sc, int slot, const RayF &r, f3 O, f3 D, Best &best, unsigned long long *n_prims) {
    float4 a = __ldg(&sc.tri_verts[3 * slot + 0]);
    float4 b = __ldg(&sc.tri_verts[3 * slot + 1]);
    float4 c = __ldg(&sc.tri_verts[3 * slot + 2]);
    int geom = __float_as_int(a.w), prim = __float_as_int(b.w), kind = __float_as_int(c.w);
    if (kind == 0) {
        TriHit h;
        f3 p0{a.x, a.y, a.z}, p1{b.x, b.y, b.z}, p2{c.x, c.y, c.z};
        if (tri_test(O, D, r.tnear, r.tfar, p0, p1, p2, h)) {
            if (ANY) return true;
            if (hit_better(h.t, geom, prim, best)) {
                best.t = h.t; best.geom = geom; best.prim = prim; best.u = h.u; best.v = h.v; best.ng = h.ng; best.sphere = -1;
            }
        }
    } else {
        int si = __float_as_int(a.x);
        SpherePrim s = sc.spheres[si];
        double t = sphere_test(r, s, r.tfar);
        if (t >= 0) {
            if (ANY) return true;
            float tf = (float)t;
            if (hit_better(tf, geom, prim, best)) {
                best.t = tf; best.geom = geom; best.prim = prim; best.sphere = si; best.ts = t;
            }
        }
    }
    return false;
}

// ---- binary-BVH traversal (simple path; also the validation reference for the 8-wide one) ----
template <bool ANY>
LJ_D bool traverse_bvh2(const DScene &sc, const RayF &r, Hit *hit_out) {
    if (sc.num_prims == 0) return false;
    const float idx = 1.0f / r.dx, idy = 1.0f / r.dy, idz = 1.0f / r.dz;
    const f3 O{r.ox, r.oy, r.oz}, D{r.dx, r.dy, r.dz};
    Best best; best.geom = -1; best.t = r.tfar; best.sphere = -1; best.prim = -1;
    const float INF = __int_as_float(0x7f800000);
    int stack[64]; int sp = 0;
    int node = sc.bvh2_root;   // >= 0 inner, < 0 leaf
    if (sc.num_prims == 1) node = ~0;
    while (true) {
        if (node < 0) {
            if (test_prim<ANY>(sc, ~node, r, O, D, best, nullptr)) return true;
            if (sp == 0) break;
            node = stack[--sp];
            continue;
        }
        const float4 *np = reinterpret_cast<const float4 *>(&sc.bvh2[node]);
        float4 n0 = __ldg(np), n1 = __ldg(np + 1);
        int left = __float_as_int(n0.w), right = __float_as_int(n1.w);
        // children boxes
        float tl, tr;
        {
            if (left >= 0) {
                const float4 *cp = reinterpret_cast<const float4 *>(&sc.bvh2[left]);
                float4 c0 = __ldg(cp), c1 = __ldg(cp + 1);
                tl = slab_test(c0.x, c0.y, c0.z, c1.x, c1.y, c1.z, r, idx, idy, idz, best.t);
            } else {
                int s = ~left;
                float4 a = __ldg(&sc.tri_verts[3 * s]), b = __ldg(&sc.tri_verts[3 * s + 1]), c = __ldg(&sc.tri_verts[3 * s + 2]);
                if (__float_as_int(c.w) == 0)
                    tl = slab_test(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)),
                                   fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)),
                                   r, idx, idy, idz, best.t);
                else tl = r.tnear; // sphere: always test
            }
            if (right >= 0) {
                const float4 *cp = reinterpret_cast<const float4 *>(&sc.bvh2[right]);
                float4 c0 = __ldg(cp), c1 = __ldg(cp + 1);
                tr = slab_test(c0.x, c0.y, c0.z, c1.x, c1.y, c1.z, r, idx, idy, idz, best.t);
            } else {
                int s = ~right;
                float4 a = __ldg(&sc.tri_verts[3 * s]), b = __ldg(&sc.tri_verts[3 * s + 1]), c = __ldg(&sc.tri_verts[3 * s + 2]);
                if (__float_as_int(c.w) == 0)
                    tr = slab_test(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)),
                                   fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)),
                                   r, idx, idy, idz, best.t);
                else tr = r.tnear;
            }
        }
        bool hl = tl != INF, hr = tr != INF;
        if (hl && hr) {
            if (sp == 64) __trap();      // binary tree deeper than the stack (validation mode LJ_BVH=2 only): fail loudly, never drop a subtree
            if (tl <= tr) { stack[sp++] = right; node = left; }
            else { stack[sp++] = left; node = right; }
        } else if (hl) node = left;
        else if (hr) node = right;
        else { if (sp == 0) break; node = stack[--sp]; }
    }
    if (ANY) return false;
    if (best.geom < 0) return false;
    hit_out->t = best.t; hit_out->geom = best.geom; hit_out->prim = best.prim;
    if (best.sphere >= 0) {
        sphere_hit_details(r, sc.spheres[best.sphere], best.ts, *hit_out);
    } else {
        hit_out->u = best.u; hit_out->v = best.v; hit_out->ngx = best.ng.x; hit_out->ngy = best.ng.y; hit_out->ngz = best.ng.z;
    }
    return true;
}
