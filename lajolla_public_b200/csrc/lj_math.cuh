// Small f64 vector / frame / matrix toolkit for the device code.
// Semantics follow the reference's math core where they affect results
// (/root/reference/src/vector.h:193-197 `v / s` multiplies by the reciprocal,
//  vector.h:249-257 normalize(0) = 0, frame.h:11-21 branchless ONB with the
//  n.z < -1+1e-6 special case, transform.cpp:79-100 xform_point divides by w).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#define LJ_HD __host__ __device__ __forceinline__
#define LJ_D __device__ __forceinline__

typedef double Real;

#define LJ_PI 3.14159265358979323846
#define LJ_INVPI (1.0 / LJ_PI)
#define LJ_TWOPI (2.0 * LJ_PI)
#define LJ_INVTWOPI (1.0 / LJ_TWOPI)
#define LJ_FOURPI (4.0 * LJ_PI)
#define LJ_INVFOURPI (1.0 / LJ_FOURPI)

struct d2 { double x, y; };
struct d3 { double x, y, z; };

LJ_HD d3 mk3(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
LJ_HD d3 mk3(double v) { return mk3(v, v, v); }
LJ_HD d2 mk2(double x, double y) { d2 r; r.x = x; r.y = y; return r; }

LJ_HD d3 operator+(d3 a, d3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
LJ_HD d3 operator-(d3 a, d3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
LJ_HD d3 operator-(d3 a) { return mk3(-a.x, -a.y, -a.z); }
LJ_HD d3 operator*(d3 a, d3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
LJ_HD d3 operator*(d3 a, double s) { return mk3(a.x * s, a.y * s, a.z * s); }
LJ_HD d3 operator*(double s, d3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
// vector.h:193-197: division by a scalar is a multiplication by 1/s
LJ_HD d3 operator/(d3 a, double s) { double inv = 1.0 / s; return mk3(a.x * inv, a.y * inv, a.z * inv); }
// x / d for an x that is often exactly zero.  IEEE: (+-0) / d = +-0 with the sign of x xor d for every finite non-zero d,
// which is what (+-0) * d gives, so the quotient is returned without dividing.  Worth a branch because CUDA's f64
// division leaves its ~12-instruction fast path for a ~65-instruction subroutine whenever the numerator (or the
// quotient) is zero or tiny: ncu showed that subroutine as 18 % of the grid-medium tracker's instructions (sigma_t = 0 in
// empty voxels), run at 8 of 32 lanes; hetvol_colored 31.0 -> 38.5 Msamples/s.  Bit-identical by construction;
// profiles/r02_experiments.md (where it did NOT pay: the uv partials of lj_shading.cuh).
LJ_HD double div_z(double x, double d) {
    const double ad = fabs(d);
    return (x == 0.0 && ad > 0.0 && ad <= 1.7976931348623157e308) ? x * d : x / d;
}
LJ_HD d3 operator/(d3 a, d3 b) { return mk3(div_z(a.x, b.x), div_z(a.y, b.y), div_z(a.z, b.z)); }
LJ_HD d3 &operator+=(d3 &a, d3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
LJ_HD d3 operator-(double s, d3 a) { return mk3(s - a.x, s - a.y, s - a.z); }

LJ_HD d2 operator+(d2 a, d2 b) { return mk2(a.x + b.x, a.y + b.y); }
LJ_HD d2 operator-(d2 a, d2 b) { return mk2(a.x - b.x, a.y - b.y); }
LJ_HD d2 operator*(double s, d2 a) { return mk2(s * a.x, s * a.y); }

LJ_HD double dot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
LJ_HD d3 cross(d3 a, d3 b) {
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
LJ_HD double length_squared(d3 a) { return dot(a, a); }
LJ_HD double length(d3 a) { return sqrt(dot(a, a)); }
LJ_HD double distance_squared(d3 a, d3 b) { return length_squared(a - b); }
LJ_HD double distance(d3 a, d3 b) { return sqrt(distance_squared(a, b)); }
LJ_HD d3 normalize(d3 a) {
    double l = length(a);
    if (l <= 0) return mk3(0, 0, 0);
    return a / l;
}
LJ_HD double max3(d3 a) { double m = a.x > a.y ? a.x : a.y; return m > a.z ? m : a.z; }  // vector.h:266-268 on top of lajolla.h:64-66
LJ_HD double average(d3 a) { return (a.x + a.y + a.z) / 3; }
LJ_HD double luminance(d3 s) { return s.x * 0.212671 + s.y * 0.715160 + s.z * 0.072169; }   // spectrum.h:32-34
LJ_HD double comp(d3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// lajolla.h:64-72: max(a,b) = a > b ? a : b ; min(a,b) = a < b ? a : b  (NaN behaviour differs from fmax)
LJ_HD double lmax(double a, double b) { return a > b ? a : b; }
LJ_HD double lmin(double a, double b) { return a < b ? a : b; }
// std::clamp(v, lo, hi) = (v < lo) ? lo : (hi < v) ? hi : v
LJ_HD double clampd(double v, double lo, double hi) { return (v < lo) ? lo : ((hi < v) ? hi : v); }
LJ_HD int clampi(int v, int lo, int hi) { return (v < lo) ? lo : ((hi < v) ? hi : v); }
LJ_HD d3 sqrt3(d3 s) { return mk3(sqrt(lmax(s.x, 0.0)), sqrt(lmax(s.y, 0.0)), sqrt(lmax(s.z, 0.0))); }  // spectrum.h:20-24

LJ_HD double modulo_d(double a, double b) { double r = fmod(a, b); return (r < 0.0) ? r + b : r; }  // lajolla.h:59-62
// modulo_d(a, 1): fmod(a, 1) is the fractional part with the sign of a, and a - trunc(a) computes it
// exactly (same value bit for bit; fmod's software loop is ~100 instructions in f64)
LJ_HD double modulo_one(double a) {
    double r = a - trunc(a);
    if (!(fabs(a) < 4503599627370496.0)) r = fmod(a, 1.0);   // |a| >= 2^52, inf, nan: leave it to fmod
    return (r < 0.0) ? r + 1.0 : r;
}
// lajolla.h:49-52.  Texture wrap indices lie in [-b, 2b): one conditional add/sub gives a % b exactly;
// anything else takes the integer division.
LJ_HD int modulo_i(int a, int b) {
    if (a >= -b && a < 2 * b && b > 0) return a < 0 ? a + b : (a >= b ? a - b : a);
    int r = a % b; return (r < 0) ? r + b : r;
}

struct Frame { d3 x, y, n; };

// frame.h:11-21
LJ_HD void coordinate_system(d3 n, d3 &a_out, d3 &b_out) {
    if (n.z < (-1 + 1e-6)) {
        a_out = mk3(0, -1, 0); b_out = mk3(-1, 0, 0);
    } else {
        double a = 1 / (1 + n.z);
        double b = -n.x * n.y * a;
        a_out = mk3(1 - n.x * n.x * a, b, -n.x);
        b_out = mk3(b, 1 - n.y * n.y * a, -n.y);
    }
}
LJ_HD Frame make_frame(d3 n) { Frame f; f.n = n; coordinate_system(n, f.x, f.y); return f; }
LJ_HD Frame make_frame(d3 x, d3 y, d3 n) { Frame f; f.x = x; f.y = y; f.n = n; return f; }
LJ_HD Frame neg_frame(Frame f) { return make_frame(-f.x, -f.y, -f.n); }
LJ_HD d3 to_local(const Frame &f, d3 v) { return mk3(dot(v, f.x), dot(v, f.y), dot(v, f.n)); }
LJ_HD d3 to_world(const Frame &f, d3 v) { return f.x * v.x + f.y * v.y + f.n * v.z; }

// transform.cpp:79-100 (row-major 4x4)
LJ_HD d3 xform_point(const double *m, d3 p) {
    double x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    double y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    double z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    double w = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    double inv_w = 1.0 / w;
    return mk3(x * inv_w, y * inv_w, z * inv_w);
}
LJ_HD d3 xform_vector(const double *m, d3 v) {
    return mk3(m[0] * v.x + m[1] * v.y + m[2] * v.z,
               m[4] * v.x + m[5] * v.y + m[6] * v.z,
               m[8] * v.x + m[9] * v.y + m[10] * v.z);
}

// ---- PCG32 (pcg.h:15-67): pure 64-bit integer arithmetic, bit exact ----
struct Pcg32 { uint64_t state, inc; };
LJ_HD uint32_t pcg_next(Pcg32 &r) {
    uint64_t old = r.state;
    r.state = old * 6364136223846793005ULL + (r.inc | 1);
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((0u - rot) & 31));
}
LJ_HD Pcg32 pcg_init(uint64_t stream_id, uint64_t seed = 0x853c49e6748fea9bULL) {
    Pcg32 s; s.state = 0; s.inc = (stream_id << 1u) | 1u;
    pcg_next(s); s.state += seed; pcg_next(s);
    return s;
}
LJ_HD double pcg_real(Pcg32 &r) {   // pcg.h:60-67: 32 random bits on top of a 52-bit mantissa
    uint64_t u = ((uint64_t)pcg_next(r) << 20) | 0x3ff0000000000000ULL;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u) - 1.0;
#else
    double d; memcpy(&d, &u, 8); return d - 1.0;
#endif
}
