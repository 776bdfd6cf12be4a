// On-disk asset formats decoded on the device: what the reference's loaders do between the bytes
// of a file and the arrays of a TriangleMesh / GridVolume, i.e. the callers-side data formats of the hot path.
//
//   .vol         volume.cpp:6-104        header check on the host; f32 -> f64 widening, channel selection /
//                                        replication (fromRGB is the identity, spectrum.h) and max_data on the device.
//   .serialized  load_serialized.cpp:174-256   shape dictionary + zlib inflate on the host (one deflate stream is one
//                                        serial bit stream); float|double -> f64 widening, xform_point(to_world),
//                                        xform_normal(inverse(to_world)), uv / index unpacking on the device.
//   .obj         parse_obj.cpp:137-234   entirely on the device: line splitting, token classification, decimal ->
//                                        double conversion, face triangulation, (v, vt, vn) de-duplication in order of
//                                        first appearance (the std::map + push_back of get_vertex_id, :119-135),
//                                        transforms, and compute_normal (:57-92) when the file has no normals.
//
// Everything is bit-identical to the reference's loaders (tests/test_gpu_assets.py against `ref_driver mesh|vol`),
// except compute_normal's angles, which go through the device's asin() (last-ulp differences from glibc's).
// Decimal conversion is exact (correctly rounded, like strtod) for numbers with <= 19 significant digits whose
// mantissa is < 2^53 and whose decimal exponent is within +-22 (+15 more for small mantissas) -- one IEEE
// multiplication or division of two exactly represented values.  Anything else is refused with LJ_EUNSUPPORTED
// rather than approximated.  -fmad=false keeps every a*b+c as two roundings, as in the reference's x86-64 build.
#include "lj_internal.h"
#include "lj_alloc.h"
#include "lj_math.cuh"
#include <cub/cub.cuh>
#include <zlib.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

// ---------------------------------------------------------------- small helpers
struct DevBuf {                       // stream-ordered allocation freed at scope exit
    void *p = nullptr;
    ~DevBuf() { lj_free(p); }
    cudaError_t alloc(size_t bytes) { return lj_malloc(&p, bytes); }
    template <typename T> T *as() const { return (T *)p; }
};
inline unsigned blocks_for(long long n, int threads = 256) { return (unsigned)((n + threads - 1) / threads); }

enum {                                // device-side refusal reasons (first one wins, with its line number)
    ASSET_OK = 0, OBJ_E_NUMBER = 1, OBJ_E_FIELDS = 2, OBJ_E_NGON = 3, OBJ_E_INDEX = 4, OBJ_E_MIXED = 5
};
const char *obj_error_text(int code) {
    switch (code) {
        case OBJ_E_NUMBER: return "a number lies outside the exactly convertible range (<= 19 digits, mantissa < 2^53, |exp10| <= 22) or is malformed";
        case OBJ_E_FIELDS: return "a v / vt / vn / f line has too few fields";
        case OBJ_E_NGON: return "The object file contains n-gon (n>4) that we do not support.";
        case OBJ_E_INDEX: return "a face refers to a vertex / uv / normal index outside the file's pools (or a non-positive one)";
        case OBJ_E_MIXED: return "faces with and without uv (or normal) indices are mixed";
    }
    return "unknown";
}
struct ErrSlot { unsigned long long packed; };   // (code << 32 | line) of the smallest line that failed
__device__ __forceinline__ void raise(ErrSlot *e, int code, unsigned line) {
    // keep the error of the first line: store ~line in the high word so that atomicMax prefers small line numbers
    atomicMax(&e->packed, ((unsigned long long)(0xffffffffu - line) << 32) | (unsigned)code);
}

// ---------------------------------------------------------------- .vol
// volume.cpp:66-72 (target 1: first channel of every voxel) and :78-95 (target 3: grey replicated or RGB);
// max_data starts at 0 and is a component-wise running max (lajolla.h:64-66), i.e. max(0, max_i data[i]):
// for non-negative doubles the bit pattern orders like the value, so an integer atomicMax is exact.
__global__ void k_vol_decode(const float *raw, long long nvox, int channels, int target, double *data,
                             unsigned long long *max_bits) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double v[3] = {0, 0, 0};
    if (i < nvox) {
        if (target == 1) { v[0] = (double)raw[(long long)channels * i]; data[i] = v[0]; }
        else {
            if (channels == 1) { v[0] = v[1] = v[2] = (double)raw[i]; }
            else { v[0] = (double)raw[3 * i]; v[1] = (double)raw[3 * i + 1]; v[2] = (double)raw[3 * i + 2]; }
            data[3 * i] = v[0]; data[3 * i + 1] = v[1]; data[3 * i + 2] = v[2];
        }
    }
    for (int c = 0; c < target; c++) {
        double m = v[c] > 0.0 ? v[c] : 0.0;                 // NaN and negatives never replace the running max
        for (int o = 16; o > 0; o >>= 1) { double t = __shfl_xor_sync(0xffffffffu, m, o); m = t > m ? t : m; }
        if ((threadIdx.x & 31) == 0 && m > 0.0) atomicMax(&max_bits[c], (unsigned long long)__double_as_longlong(m));
    }
}

// ---------------------------------------------------------------- .serialized
// load_position<Precision> + xform_point (load_serialized.cpp:123-135, :207-214)
template <typename P>
__global__ void k_ser_positions(const P *src, long long nv, const double *to_world, double *out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nv) return;
    d3 p = xform_point(to_world, mk3((double)src[3 * i], (double)src[3 * i + 1], (double)src[3 * i + 2]));
    out[3 * i] = p.x; out[3 * i + 1] = p.y; out[3 * i + 2] = p.z;
}
// transform.cpp:95-100: n' = transpose(inv) * n (no renormalisation)
__device__ __forceinline__ d3 xform_normal_dev(const double *inv, d3 n) {
    return mk3(inv[0] * n.x + inv[4] * n.y + inv[8] * n.z,
               inv[1] * n.x + inv[5] * n.y + inv[9] * n.z,
               inv[2] * n.x + inv[6] * n.y + inv[10] * n.z);
}
template <typename P>
__global__ void k_ser_normals(const P *src, long long nv, const double *to_local, double *out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nv) return;
    d3 n = xform_normal_dev(to_local, mk3((double)src[3 * i], (double)src[3 * i + 1], (double)src[3 * i + 2]));
    out[3 * i] = n.x; out[3 * i + 1] = n.y; out[3 * i + 2] = n.z;
}
template <typename P>
__global__ void k_ser_widen(const P *src, long long n, double *out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (double)src[i];
}

// ---------------------------------------------------------------- compute_normal (parse_obj.cpp:49-92)
__device__ __forceinline__ double unit_angle_dev(d3 u, d3 v) {      // parse_obj.cpp:49-55, (pi - 2) factor kept as is
    if (dot(u, v) < 0) return (LJ_PI - 2) * asin(0.5 * length(v + u));
    return 2 * asin(0.5 * length(v - u));
}
__device__ __forceinline__ d3 ld3(const double *a, long long i) { return mk3(a[3 * i], a[3 * i + 1], a[3 * i + 2]); }

__global__ void k_corner_keys(const int32_t *indices, long long ncorner, unsigned *key, unsigned *val) {
    const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < ncorner) { key[c] = (unsigned)indices[c]; val[c] = (unsigned)c; }
}
// One thread per vertex adds the angle-weighted face normals of its corners in (face, corner) order -- the order in
// which the reference's face loop reaches `normals[index[i]] = normals[index[i]] + n * angle` -- so the f64 sum
// rounds identically; the corners of a vertex are a run of the stably sorted (vertex, face*3+corner) list.
__global__ void k_vertex_normals(const double *pos, const int32_t *indices, const unsigned *skey, const unsigned *sval,
                                 long long ncorner, long long nv, double *normals) {
    const long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nv) return;
    long long lo = 0, hi = ncorner;
    while (lo < hi) { long long mid = (lo + hi) >> 1; if (skey[mid] < (unsigned)v) lo = mid + 1; else hi = mid; }
    d3 acc = mk3(0, 0, 0);
    for (long long k = lo; k < ncorner && skey[k] == (unsigned)v; k++) {
        const unsigned c = sval[k];
        const long long t = c / 3; const int i = (int)(c % 3);
        const int32_t *ix = indices + 3 * t;
        d3 a0 = ld3(pos, ix[0]), a1 = ld3(pos, ix[1]), a2 = ld3(pos, ix[2]);
        d3 n = cross(a1 - a0, a2 - a0);
        const double l = length(n);
        if (l == 0) continue;                               // `break` at i == 0: the face contributes to no corner
        n = n / l;
        d3 v0 = ld3(pos, ix[i]), v1 = ld3(pos, ix[(i + 1) % 3]), v2 = ld3(pos, ix[(i + 2) % 3]);
        const double angle = unit_angle_dev(normalize(v1 - v0), normalize(v2 - v0));
        acc = acc + n * angle;
    }
    const double l = length(acc);
    if (l != 0) acc = acc / l; else acc = mk3(0, 0, 0);
    normals[3 * v] = acc.x; normals[3 * v + 1] = acc.y; normals[3 * v + 2] = acc.z;
}

int compute_normals_device(const double *d_pos, long long nv, const int32_t *d_idx, long long nt, double *d_normals) {
    const long long nc = 3 * nt;
    if (nv == 0) return LJ_OK;
    DevBuf key, val, skey, sval, tmp;
    LJ_CUDA(key.alloc(nc * 4)); LJ_CUDA(val.alloc(nc * 4)); LJ_CUDA(skey.alloc(nc * 4)); LJ_CUDA(sval.alloc(nc * 4));
    if (nc) k_corner_keys<<<blocks_for(nc), 256>>>(d_idx, nc, key.as<unsigned>(), val.as<unsigned>());
    size_t tb = 0;
    LJ_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, key.as<unsigned>(), skey.as<unsigned>(), val.as<unsigned>(), sval.as<unsigned>(), (int)nc));
    LJ_CUDA(tmp.alloc(tb));
    if (nc) LJ_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tb, key.as<unsigned>(), skey.as<unsigned>(), val.as<unsigned>(), sval.as<unsigned>(), (int)nc));
    k_vertex_normals<<<blocks_for(nv, 128), 128>>>(d_pos, d_idx, skey.as<unsigned>(), sval.as<unsigned>(), nc, nv, d_normals);
    LJ_CUDA(cudaGetLastError());
    return LJ_OK;
}

// ---------------------------------------------------------------- .obj on the device
__device__ __forceinline__ bool is_space(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13); }   // std::isspace, "C" locale

struct ObjCounts { int v, vt, vn, tri, corner; };
struct ObjCountsSum {
    __host__ __device__ ObjCounts operator()(const ObjCounts &a, const ObjCounts &b) const {
        ObjCounts r; r.v = a.v + b.v; r.vt = a.vt + b.vt; r.vn = a.vn + b.vn; r.tri = a.tri + b.tri; r.corner = a.corner + b.corner; return r;
    }
};

__global__ void k_obj_line_flags(const unsigned char *text, long long n, int *flag) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = (i == 0 || text[i - 1] == '\n') ? 1 : 0;
}
__global__ void k_obj_line_starts(const int *flag, const int *rank, long long n, long long *start) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flag[i]) start[rank[i]] = i;
}
// trimmed extent of line l (std::getline + trim, parse_obj.cpp:147-149)
__device__ __forceinline__ void line_extent(const unsigned char *text, long long n, const long long *start, long long nlines,
                                            long long l, long long &b, long long &e) {
    b = start[l];
    e = (l + 1 < nlines) ? start[l + 1] - 1 : ((n > 0 && text[n - 1] == '\n') ? n - 1 : n);
    while (b < e && is_space(text[b])) b++;
    while (e > b && is_space(text[e - 1])) e--;
}
enum { OBJ_NONE = 0, OBJ_V = 1, OBJ_VT = 2, OBJ_VN = 3, OBJ_F = 4 };

// pass 1: classify every line and count what it contributes (pool entries, triangles, face corners)
__global__ void k_obj_classify(const unsigned char *text, long long n, const long long *start, long long nlines,
                               unsigned char *kind, ObjCounts *counts, ErrSlot *err) {
    const long long l = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= nlines) return;
    long long b, e; line_extent(text, n, start, nlines, l, b, e);
    ObjCounts c = {0, 0, 0, 0, 0};
    int k = OBJ_NONE;
    if (b < e && text[b] != '#') {
        long long t = b; while (t < e && !is_space(text[t])) t++;
        const long long len = t - b;
        if (len == 1 && text[b] == 'v') { k = OBJ_V; c.v = 1; }
        else if (len == 2 && text[b] == 'v' && text[b + 1] == 't') { k = OBJ_VT; c.vt = 1; }
        else if (len == 2 && text[b] == 'v' && text[b + 1] == 'n') { k = OBJ_VN; c.vn = 1; }
        else if (len == 1 && text[b] == 'f') {
            k = OBJ_F;
            int fields = 0;
            while (t < e) {
                while (t < e && is_space(text[t])) t++;
                if (t < e) { fields++; while (t < e && !is_space(text[t])) t++; }
            }
            if (fields < 3) raise(err, OBJ_E_FIELDS, (unsigned)l);
            else if (fields > 4) raise(err, OBJ_E_NGON, (unsigned)l);
            c.tri = fields == 4 ? 2 : 1; c.corner = fields == 4 ? 4 : 3;
        }
    }
    kind[l] = (unsigned char)k;
    counts[l] = c;
}

// Decimal text -> double, correctly rounded in the range described at the top of the file.
// Returns the position after the number, or -1 (malformed / outside the exact range).
__device__ long long parse_real(const unsigned char *text, long long p, long long e, double &out) {
    const double p10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16,
                            1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    while (p < e && is_space(text[p])) p++;
    if (p >= e) return -1;
    bool neg = false;
    if (text[p] == '+' || text[p] == '-') { neg = text[p] == '-'; p++; }
    unsigned long long m = 0; int nd = 0, dexp = 0; bool any = false, point = false;
    for (; p < e; p++) {
        const unsigned char ch = text[p];
        if (ch == '.') { if (point) return -1; point = true; continue; }
        if (ch < '0' || ch > '9') break;
        any = true;
        const int d = ch - '0';
        if (nd == 0 && d == 0) { if (point) dexp--; continue; }      // leading zeros
        if (nd < 19) { m = m * 10 + (unsigned)d; nd++; if (point) dexp--; }
        else { if (d != 0) return -1; if (!point) dexp++; }          // only trailing zeros may follow 19 digits
    }
    if (!any) return -1;
    if (p < e && (text[p] == 'e' || text[p] == 'E')) {
        p++;
        bool eneg = false;
        if (p < e && (text[p] == '+' || text[p] == '-')) { eneg = text[p] == '-'; p++; }
        if (p >= e || text[p] < '0' || text[p] > '9') return -1;
        int ex = 0;
        for (; p < e && text[p] >= '0' && text[p] <= '9'; p++) { ex = ex * 10 + (text[p] - '0'); if (ex > 100000) return -1; }
        dexp += eneg ? -ex : ex;
    }
    if (p < e && !is_space(text[p])) return -1;                      // e.g. "1.5abc", "nan", "inf"
    double v;
    if (m == 0) v = 0.0;
    else {
        if (m >= (1ULL << 53)) return -1;
        v = (double)(long long)m;
        if (dexp > 22) {                                             // m * 10^(dexp-22) stays an exact integer < 2^53?
            const int extra = dexp - 22;
            if (extra > 15) return -1;
            v = v * p10[extra];
            if (v >= 9007199254740992.0) return -1;
            dexp = 22;
        }
        if (dexp < -22) return -1;
        v = dexp >= 0 ? v * p10[dexp] : v / p10[-dexp];
    }
    out = neg ? -v : v;
    return p;
}
// std::stoi on one "/"-separated field of a face token (split_face_str, parse_obj.cpp:31-46): empty field -> 0
__device__ __forceinline__ bool parse_face_token(const unsigned char *text, long long b, long long e, int id[3]) {
    id[0] = id[1] = id[2] = 0;
    int f = 0; long long p = b;
    while (true) {
        long long q = p; while (q < e && text[q] != '/') q++;
        if (f < 3 && q > p) {
            long long r = p; bool neg = false;
            if (text[r] == '+' || text[r] == '-') { neg = text[r] == '-'; r++; }
            if (r >= q) return false;
            long long val = 0;
            for (; r < q; r++) { if (text[r] < '0' || text[r] > '9') return false; val = val * 10 + (text[r] - '0'); if (val > 2000000000LL) return false; }
            id[f] = neg ? -(int)val : (int)val;
        }
        f++;
        if (q >= e) break;
        p = q + 1;
    }
    return true;
}

#define OBJ_KEY_BITS 21
#define OBJ_KEY_LIMIT ((1 << OBJ_KEY_BITS) - 1)

// pass 2: parse every line into its slot of the pools / corner list (slots from the exclusive scan of pass 1)
__global__ void k_obj_parse(const unsigned char *text, long long n, const long long *start, long long nlines,
                            const unsigned char *kind, const ObjCounts *offs, ObjCounts total,
                            double *pool_v, double *pool_vt, double *pool_vn,
                            unsigned long long *corner_key, int *tri_corner, ErrSlot *err) {
    const long long l = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= nlines) return;
    const int k = kind[l];
    if (k == OBJ_NONE) return;
    long long b, e; line_extent(text, n, start, nlines, l, b, e);
    long long p = b; while (p < e && !is_space(text[p])) p++;          // skip the keyword
    const ObjCounts o = offs[l];
    if (k == OBJ_V) {                     // parse_obj.cpp:155-158: Vector3{x, y, z} / w, w = 1 unless a 4th number follows
        double x, y, z, w = 1;
        if ((p = parse_real(text, p, e, x)) < 0 || (p = parse_real(text, p, e, y)) < 0 || (p = parse_real(text, p, e, z)) < 0) {
            raise(err, OBJ_E_NUMBER, (unsigned)l); return;
        }
        long long q = p; while (q < e && is_space(text[q])) q++;
        if (q < e && parse_real(text, q, e, w) < 0) { raise(err, OBJ_E_NUMBER, (unsigned)l); return; }
        d3 r = mk3(x, y, z) / w;
        pool_v[3 * o.v] = r.x; pool_v[3 * o.v + 1] = r.y; pool_v[3 * o.v + 2] = r.z;
    } else if (k == OBJ_VT) {             // :159-162: {s, 1 - t}
        double s, t;
        if ((p = parse_real(text, p, e, s)) < 0 || (p = parse_real(text, p, e, t)) < 0) { raise(err, OBJ_E_NUMBER, (unsigned)l); return; }
        pool_vt[2 * o.vt] = s; pool_vt[2 * o.vt + 1] = 1 - t;
    } else if (k == OBJ_VN) {             // :163-166: normalize({x, y, z})
        double x, y, z;
        if ((p = parse_real(text, p, e, x)) < 0 || (p = parse_real(text, p, e, y)) < 0 || (p = parse_real(text, p, e, z)) < 0) {
            raise(err, OBJ_E_NUMBER, (unsigned)l); return;
        }
        d3 r = normalize(mk3(x, y, z));
        pool_vn[3 * o.vn] = r.x; pool_vn[3 * o.vn + 1] = r.y; pool_vn[3 * o.vn + 2] = r.z;
    } else {                              // :167-229: corners in the order get_vertex_id sees them; quad -> (0,1,2) + (0,2,3)
        int c = 0;
        while (p < e && c < 4) {
            while (p < e && is_space(text[p])) p++;
            if (p >= e) break;
            long long q = p; while (q < e && !is_space(text[q])) q++;
            int id[3];
            if (!parse_face_token(text, p, q, id)) { raise(err, OBJ_E_INDEX, (unsigned)l); return; }
            const int v = id[0] - 1, vt = id[1] - 1, vn = id[2] - 1;
            if (v < 0 || v >= total.v || vt < -1 || vt >= total.vt || vn < -1 || vn >= total.vn ||
                v >= OBJ_KEY_LIMIT || vt >= OBJ_KEY_LIMIT || vn >= OBJ_KEY_LIMIT) { raise(err, OBJ_E_INDEX, (unsigned)l); return; }
            corner_key[o.corner + c] = ((unsigned long long)(v + 1) << (2 * OBJ_KEY_BITS)) | ((unsigned long long)(vt + 1) << OBJ_KEY_BITS) |
                                       (unsigned long long)(vn + 1);
            c++; p = q;
        }
        int *t = tri_corner + 3 * (long long)o.tri;
        t[0] = o.corner; t[1] = o.corner + 1; t[2] = o.corner + 2;
        if (c == 4) { t[3] = o.corner; t[4] = o.corner + 2; t[5] = o.corner + 3; }
    }
}
__global__ void k_iota(unsigned *a, long long n) { const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; if (i < n) a[i] = (unsigned)i; }
// a (v, vt, vn) triple gets its mesh vertex at its first appearance in corner order (std::map::find miss, :127-134):
// after the stable sort by key, the head of each run of equal keys is that first appearance
__global__ void k_obj_mark_first(const unsigned long long *skey, const unsigned *sseq, long long nc, int *is_first) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nc && (i == 0 || skey[i] != skey[i - 1])) is_first[sseq[i]] = 1;
}
// heads write the vertex (transformed as get_vertex_id does) and hand their id to every corner of the run
__global__ void k_obj_emit_vertices(const unsigned long long *skey, const unsigned *sseq, long long nc, const int *rank,
                                    const double *pool_v, const double *pool_vt, const double *pool_vn,
                                    const double *to_world, const double *to_local,
                                    double *pos, double *uv, double *nor, int *corner_id, int *n_with_vt, int *n_with_vn) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc || !(i == 0 || skey[i] != skey[i - 1])) return;
    const unsigned long long key = skey[i];
    const int id = rank[sseq[i]];
    const int v = (int)(key >> (2 * OBJ_KEY_BITS)) - 1, vt = (int)((key >> OBJ_KEY_BITS) & OBJ_KEY_LIMIT) - 1, vn = (int)(key & OBJ_KEY_LIMIT) - 1;
    d3 p = xform_point(to_world, ld3(pool_v, v));
    pos[3 * id] = p.x; pos[3 * id + 1] = p.y; pos[3 * id + 2] = p.z;
    if (vt >= 0) { uv[2 * id] = pool_vt[2 * vt]; uv[2 * id + 1] = pool_vt[2 * vt + 1]; atomicAdd(n_with_vt, 1); }
    if (vn >= 0) { d3 nn = xform_normal_dev(to_local, ld3(pool_vn, vn)); nor[3 * id] = nn.x; nor[3 * id + 1] = nn.y; nor[3 * id + 2] = nn.z; atomicAdd(n_with_vn, 1); }
    for (long long j = i; j < nc && skey[j] == key; j++) corner_id[sseq[j]] = id;
}
__global__ void k_obj_emit_triangles(const int *tri_corner, const int *corner_id, long long n3, int32_t *indices) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n3) indices[i] = corner_id[tri_corner[i]];
}

void mesh_release(lj_mesh_data *m) {
    free(m->positions); free(m->normals); free(m->uvs); free(m->indices);
    memset(m, 0, sizeof *m);
}
template <typename T> T *host_copy(const void *dev, size_t count, cudaError_t &e) {
    T *h = (T *)malloc((count ? count : 1) * sizeof(T));
    if (h && count && e == cudaSuccess) e = cudaMemcpy(h, dev, count * sizeof(T), cudaMemcpyDeviceToHost);
    return h;
}

struct Timer {
    cudaEvent_t a = nullptr, b = nullptr;
    void begin() { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, 0); }
    double end() { float ms = 0; cudaEventRecord(b, 0); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b); cudaEventDestroy(a); cudaEventDestroy(b); return ms; }
};

}  // namespace

// =============================================================================================== C ABI
extern "C" void lj_mesh_free(lj_mesh_data *mesh) { if (mesh) mesh_release(mesh); }
extern "C" void lj_volume_free(lj_volume_data *vol) { if (vol) { free(vol->data); memset(vol, 0, sizeof *vol); } }

extern "C" int lj_volume_load(lj_ctx *ctx, const void *file, uint64_t file_bytes, int32_t target_channels, lj_volume_data *out) {
    if (!ctx || !file || !out) { lj_set_error("lj_volume_load: null argument"); return LJ_EINVAL; }
    memset(out, 0, sizeof *out);
    if (target_channels != 1 && target_channels != 3) { lj_set_error("lj_volume_load: target_channels must be 1 or 3"); return LJ_EINVAL; }
    const unsigned char *f = (const unsigned char *)file;
    // volume.cpp:16-61: 'V' 'O' 'L' 3, int type (1 = float32), int xres yres zres, int channels, 6 float bounds = 48 bytes
    if (file_bytes < 48 || f[0] != 'V' || f[1] != 'O' || f[2] != 'L' || f[3] != 3) {
        lj_set_error("Error loading volume from a file (incorrect header)."); return LJ_EIO;
    }
    int32_t hd[5]; memcpy(hd, f + 4, 20);
    float bounds[6]; memcpy(bounds, f + 24, 24);
    if (hd[0] != 1) { lj_set_error("Unsupported volume format (only support Float32)."); return LJ_EUNSUPPORTED; }
    if (hd[4] != 1 && hd[4] != 3) { lj_set_error("Unsupported volume format (wrong number of channels)."); return LJ_EUNSUPPORTED; }
    if (hd[1] <= 0 || hd[2] <= 0 || hd[3] <= 0) { lj_set_error("lj_volume_load: empty grid"); return LJ_EIO; }
    const long long nvox = (long long)hd[1] * hd[2] * hd[3];
    const int channels = hd[4];
    if (file_bytes < 48 + (uint64_t)nvox * channels * 4) { lj_set_error("lj_volume_load: file is shorter than its header says"); return LJ_EIO; }
    LJ_CUDA(cudaSetDevice(ctx->device));
    DevBuf raw, data, mx;
    LJ_CUDA(raw.alloc(nvox * channels * 4)); LJ_CUDA(data.alloc(nvox * target_channels * 8)); LJ_CUDA(mx.alloc(3 * 8));
    LJ_CUDA(cudaMemcpyAsync(raw.p, f + 48, nvox * channels * 4, cudaMemcpyHostToDevice, 0));
    Timer tm; tm.begin();
    LJ_CUDA(cudaMemsetAsync(mx.p, 0, 24, 0));
    k_vol_decode<<<blocks_for(nvox), 256>>>(raw.as<float>(), nvox, channels, target_channels, data.as<double>(), mx.as<unsigned long long>());
    LJ_CUDA(cudaGetLastError());
    out->decode_ms = tm.end();
    cudaError_t e = cudaSuccess;
    out->data = host_copy<double>(data.p, nvox * target_channels, e);
    if (e == cudaSuccess) e = cudaMemcpy(out->max_data, mx.p, 24, cudaMemcpyDeviceToHost);   // bit patterns of doubles
    if (!out->data || e != cudaSuccess) { lj_volume_free(out); lj_set_error("lj_volume_load: %s", e == cudaSuccess ? "out of host memory" : cudaGetErrorString(e)); return LJ_ECUDA; }
    for (int i = 0; i < 3; i++) { out->resolution[i] = hd[1 + i]; out->p_min[i] = bounds[i]; out->p_max[i] = bounds[3 + i]; }
    out->channels = target_channels;
    out->num_values = (uint64_t)nvox * target_channels;
    return LJ_OK;
}

extern "C" int lj_mesh_compute_normals(lj_ctx *ctx, const double *positions, uint64_t num_vertices, const int32_t *indices,
                                       uint64_t num_triangles, double *normals_out) {
    if (!ctx || (!positions && num_vertices) || (!indices && num_triangles) || (!normals_out && num_vertices)) { lj_set_error("lj_mesh_compute_normals: null argument"); return LJ_EINVAL; }
    for (uint64_t i = 0; i < 3 * num_triangles; i++)
        if (indices[i] < 0 || (uint64_t)indices[i] >= num_vertices) { lj_set_error("lj_mesh_compute_normals: index %d out of range", indices[i]); return LJ_EINVAL; }
    LJ_CUDA(cudaSetDevice(ctx->device));
    DevBuf pos, idx, nor;
    LJ_CUDA(pos.alloc(num_vertices * 24)); LJ_CUDA(idx.alloc(num_triangles * 12)); LJ_CUDA(nor.alloc(num_vertices * 24));
    LJ_CUDA(cudaMemcpyAsync(pos.p, positions, num_vertices * 24, cudaMemcpyHostToDevice, 0));
    LJ_CUDA(cudaMemcpyAsync(idx.p, indices, num_triangles * 12, cudaMemcpyHostToDevice, 0));
    if (int rc = compute_normals_device(pos.as<double>(), (long long)num_vertices, idx.as<int32_t>(), (long long)num_triangles, nor.as<double>())) return rc;
    LJ_CUDA(cudaMemcpy(normals_out, nor.p, num_vertices * 24, cudaMemcpyDeviceToHost));
    return LJ_OK;
}

namespace {
// skip_to_idx (load_serialized.cpp:103-121) on a memory image of the file: where shape `idx`'s zlib stream begins
int serialized_stream_offset(const unsigned char *f, uint64_t n, int idx, size_t *offset, int *version) {
    if (n < 8) { lj_set_error("lj_mesh_load_serialized: file too short"); return LJ_EIO; }
    uint16_t ver; memcpy(&ver, f + 2, 2);
    *version = ver;
    if (ver != 3 && ver != 4) { lj_set_error("lj_mesh_load_serialized: unknown file format version %d", (int)ver); return LJ_EUNSUPPORTED; }
    if (idx <= 0) { *offset = 4; return LJ_OK; }
    uint32_t count; memcpy(&count, f + n - 4, 4);
    if ((uint32_t)idx >= count) { lj_set_error("lj_mesh_load_serialized: shape index %d of %u", idx, count); return LJ_EINVAL; }
    uint64_t off = 0;
    if (ver == 4) {
        const uint64_t at = n - 4 - 8ull * (count - idx);
        if (at + 8 > n) { lj_set_error("lj_mesh_load_serialized: bad dictionary"); return LJ_EIO; }
        memcpy(&off, f + at, 8);
    } else {
        const uint64_t at = n - 4ull * (count - idx + 1);
        uint32_t u; memcpy(&u, f + at, 4); off = u;
    }
    if (off + 4 > n) { lj_set_error("lj_mesh_load_serialized: bad shape offset"); return LJ_EIO; }
    *offset = off + 4;                                        // each shape repeats the 2 x u16 header
    return LJ_OK;
}
int inflate_all(const unsigned char *src, size_t n, std::vector<unsigned char> &out) {
    z_stream zs; memset(&zs, 0, sizeof zs);
    if (inflateInit2(&zs, 15) != Z_OK) { lj_set_error("Could not initialize ZLIB"); return LJ_EIO; }
    zs.next_in = (Bytef *)src; zs.avail_in = (uInt)(n > 0xffffffffu ? 0xffffffffu : n);
    out.resize(n * 4 + 65536);
    int rc;
    do {
        if (zs.total_out == out.size()) out.resize(out.size() * 2);
        zs.next_out = out.data() + zs.total_out; zs.avail_out = (uInt)(out.size() - zs.total_out);
        rc = inflate(&zs, Z_NO_FLUSH);
    } while (rc == Z_OK);
    const size_t total = zs.total_out;
    inflateEnd(&zs);
    if (rc != Z_STREAM_END) { lj_set_error("lj_mesh_load_serialized: inflate(): stream error %d", rc); return LJ_EIO; }
    out.resize(total);
    return LJ_OK;
}
}  // namespace

extern "C" int lj_mesh_load_serialized(lj_ctx *ctx, const void *file, uint64_t file_bytes, int32_t shape_index,
                                       const double *to_world, const double *to_local, lj_mesh_data *out) {
    if (!ctx || !file || !to_world || !to_local || !out) { lj_set_error("lj_mesh_load_serialized: null argument"); return LJ_EINVAL; }
    memset(out, 0, sizeof *out);
    const unsigned char *f = (const unsigned char *)file;
    size_t off = 0; int version = 0;
    if (int rc = serialized_stream_offset(f, file_bytes, shape_index, &off, &version)) return rc;
    std::vector<unsigned char> raw;
    if (int rc = inflate_all(f + off, file_bytes - off, raw)) return rc;
    // load_serialized.cpp:187-203: u32 flags, [v4: utf-8 name + NUL], u64 vertex count, u64 triangle count
    size_t p = 0;
    auto need = [&](size_t bytes) { return p + bytes <= raw.size(); };
    uint32_t flags = 0; uint64_t nv = 0, nt = 0;
    if (!need(4)) { lj_set_error("lj_mesh_load_serialized: truncated stream"); return LJ_EIO; }
    memcpy(&flags, raw.data(), 4); p = 4;
    if (version == 4) { while (p < raw.size() && raw[p] != 0) p++; p++; }
    if (!need(16)) { lj_set_error("lj_mesh_load_serialized: truncated stream"); return LJ_EIO; }
    memcpy(&nv, raw.data() + p, 8); memcpy(&nt, raw.data() + p + 8, 8); p += 16;
    const bool dbl = flags & 0x2000;                          // EDoublePrecision
    const bool has_n = flags & 0x0001, has_uv = flags & 0x0002, has_col = flags & 0x0008;
    const size_t prec = dbl ? 8 : 4;
    const size_t pos_at = p, nor_at = pos_at + nv * 3 * prec, uv_at = nor_at + (has_n ? nv * 3 * prec : 0),
                 col_at = uv_at + (has_uv ? nv * 2 * prec : 0), idx_at = col_at + (has_col ? nv * 3 * prec : 0);
    if (idx_at + nt * 12 > raw.size()) { lj_set_error("lj_mesh_load_serialized: stream shorter than its counts say"); return LJ_EIO; }

    LJ_CUDA(cudaSetDevice(ctx->device));
    DevBuf src, xf, pos, nor, uv;
    const size_t src_bytes = idx_at - pos_at;                 // vertex attributes only; indices need no conversion
    LJ_CUDA(src.alloc(src_bytes)); LJ_CUDA(xf.alloc(32 * 8));
    LJ_CUDA(pos.alloc(nv * 24)); LJ_CUDA(nor.alloc(nv * 24)); LJ_CUDA(uv.alloc(nv * 16));
    LJ_CUDA(cudaMemcpyAsync(src.p, raw.data() + pos_at, src_bytes, cudaMemcpyHostToDevice, 0));   // cudaMalloc alignment + multiples of prec: aligned
    LJ_CUDA(cudaMemcpyAsync(xf.p, to_world, 128, cudaMemcpyHostToDevice, 0));
    LJ_CUDA(cudaMemcpyAsync(xf.as<double>() + 16, to_local, 128, cudaMemcpyHostToDevice, 0));
    Timer tm; tm.begin();
    const unsigned char *s = src.as<unsigned char>();
    if (nv) {
        const unsigned g = blocks_for((long long)nv);
        if (dbl) k_ser_positions<double><<<g, 256>>>((const double *)s, (long long)nv, xf.as<double>(), pos.as<double>());
        else k_ser_positions<float><<<g, 256>>>((const float *)s, (long long)nv, xf.as<double>(), pos.as<double>());
        if (has_n) {
            if (dbl) k_ser_normals<double><<<g, 256>>>((const double *)(s + (nor_at - pos_at)), (long long)nv, xf.as<double>() + 16, nor.as<double>());
            else k_ser_normals<float><<<g, 256>>>((const float *)(s + (nor_at - pos_at)), (long long)nv, xf.as<double>() + 16, nor.as<double>());
        }
        if (has_uv) {
            if (dbl) k_ser_widen<double><<<blocks_for(2 * (long long)nv), 256>>>((const double *)(s + (uv_at - pos_at)), 2 * (long long)nv, uv.as<double>());
            else k_ser_widen<float><<<blocks_for(2 * (long long)nv), 256>>>((const float *)(s + (uv_at - pos_at)), 2 * (long long)nv, uv.as<double>());
        }
    }
    LJ_CUDA(cudaGetLastError());
    out->decode_ms = tm.end();
    cudaError_t e = cudaSuccess;
    out->positions = host_copy<double>(pos.p, nv * 3, e);
    if (has_n) out->normals = host_copy<double>(nor.p, nv * 3, e);
    if (has_uv) out->uvs = host_copy<double>(uv.p, nv * 2, e);
    out->indices = (int32_t *)malloc((nt ? nt : 1) * 12);
    if (out->indices) memcpy(out->indices, raw.data() + idx_at, nt * 12);
    if (e != cudaSuccess || !out->positions || !out->indices) { mesh_release(out); lj_set_error("lj_mesh_load_serialized: %s", e == cudaSuccess ? "out of host memory" : cudaGetErrorString(e)); return LJ_ECUDA; }
    out->num_vertices = nv; out->num_triangles = nt; out->has_normals = has_n; out->has_uvs = has_uv;
    return LJ_OK;
}

extern "C" int lj_mesh_load_obj(lj_ctx *ctx, const void *text, uint64_t text_bytes, const double *to_world, const double *to_local,
                                lj_mesh_data *out) {
    if (!ctx || (!text && text_bytes) || !to_world || !to_local || !out) { lj_set_error("lj_mesh_load_obj: null argument"); return LJ_EINVAL; }
    memset(out, 0, sizeof *out);
    if (text_bytes >= (1ull << 31)) { lj_set_error("lj_mesh_load_obj: files of 2 GiB and more are not supported"); return LJ_EUNSUPPORTED; }
    LJ_CUDA(cudaSetDevice(ctx->device));
    const long long n = (long long)text_bytes;
    if (n == 0) return LJ_OK;
    DevBuf txt, flag, rank, xf, err, tmp;
    LJ_CUDA(txt.alloc(n)); LJ_CUDA(flag.alloc(n * 4)); LJ_CUDA(rank.alloc(n * 4)); LJ_CUDA(xf.alloc(32 * 8)); LJ_CUDA(err.alloc(sizeof(ErrSlot)));
    LJ_CUDA(cudaMemcpyAsync(txt.p, text, n, cudaMemcpyHostToDevice, 0));
    LJ_CUDA(cudaMemcpyAsync(xf.p, to_world, 128, cudaMemcpyHostToDevice, 0));
    LJ_CUDA(cudaMemcpyAsync(xf.as<double>() + 16, to_local, 128, cudaMemcpyHostToDevice, 0));
    LJ_CUDA(cudaMemsetAsync(err.p, 0, sizeof(ErrSlot), 0));
    Timer tm; tm.begin();
    const unsigned char *d_text = txt.as<unsigned char>();

    // lines
    k_obj_line_flags<<<blocks_for(n), 256>>>(d_text, n, flag.as<int>());
    size_t tb = 0, tb2 = 0, tb3 = 0, tb4 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, flag.as<int>(), rank.as<int>(), (int)n);
    cub::DeviceScan::ExclusiveScan(nullptr, tb2, (ObjCounts *)nullptr, (ObjCounts *)nullptr, ObjCountsSum(), ObjCounts{0, 0, 0, 0, 0}, (int)n + 1);
    cub::DeviceRadixSort::SortPairs(nullptr, tb3, (unsigned long long *)nullptr, (unsigned long long *)nullptr, (unsigned *)nullptr, (unsigned *)nullptr, (int)n, 0, 3 * OBJ_KEY_BITS);
    cub::DeviceScan::ExclusiveSum(nullptr, tb4, (int *)nullptr, (int *)nullptr, (int)n + 1);
    tb = std::max(std::max(tb, tb2), std::max(tb3, tb4));     // upper bounds: no array here is longer than the text
    LJ_CUDA(tmp.alloc(tb));
    LJ_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, flag.as<int>(), rank.as<int>(), (int)n));
    int last_rank = 0, last_flag = 0;
    LJ_CUDA(cudaMemcpyAsync(&last_rank, rank.as<int>() + n - 1, 4, cudaMemcpyDeviceToHost, 0));
    LJ_CUDA(cudaMemcpy(&last_flag, flag.as<int>() + n - 1, 4, cudaMemcpyDeviceToHost));
    const long long nlines = (long long)last_rank + last_flag;
    DevBuf start, kind, counts, offs;
    LJ_CUDA(start.alloc(nlines * 8)); LJ_CUDA(kind.alloc(nlines)); LJ_CUDA(counts.alloc((nlines + 1) * sizeof(ObjCounts))); LJ_CUDA(offs.alloc((nlines + 1) * sizeof(ObjCounts)));
    k_obj_line_starts<<<blocks_for(n), 256>>>(flag.as<int>(), rank.as<int>(), n, start.as<long long>());

    // pass 1 + slots: one extra zero element so that the exclusive scan's last entry is the total
    LJ_CUDA(cudaMemsetAsync(counts.as<ObjCounts>() + nlines, 0, sizeof(ObjCounts), 0));
    k_obj_classify<<<blocks_for(nlines, 128), 128>>>(d_text, n, start.as<long long>(), nlines, kind.as<unsigned char>(), counts.as<ObjCounts>(), err.as<ErrSlot>());
    LJ_CUDA(cub::DeviceScan::ExclusiveScan(tmp.p, tb, counts.as<ObjCounts>(), offs.as<ObjCounts>(), ObjCountsSum(), ObjCounts{0, 0, 0, 0, 0}, (int)(nlines + 1)));
    ObjCounts total;
    LJ_CUDA(cudaMemcpy(&total, offs.as<ObjCounts>() + nlines, sizeof total, cudaMemcpyDeviceToHost));
    ErrSlot es;
    LJ_CUDA(cudaMemcpy(&es, err.p, sizeof es, cudaMemcpyDeviceToHost));
    auto refuse = [&](const ErrSlot &s) {
        const int code = (int)(s.packed & 0xffffffffu); const unsigned line = 0xffffffffu - (unsigned)(s.packed >> 32);
        lj_set_error("lj_mesh_load_obj: line %u: %s", line + 1, obj_error_text(code));
        return code == OBJ_E_NGON ? LJ_EINVAL : LJ_EUNSUPPORTED;
    };
    if (es.packed) return refuse(es);

    // pass 2
    const long long nc = total.corner, ntri = total.tri;
    DevBuf pv, pvt, pvn, ckey, tric;
    LJ_CUDA(pv.alloc((size_t)total.v * 24)); LJ_CUDA(pvt.alloc((size_t)total.vt * 16)); LJ_CUDA(pvn.alloc((size_t)total.vn * 24));
    LJ_CUDA(ckey.alloc(nc * 8)); LJ_CUDA(tric.alloc(ntri * 12));
    k_obj_parse<<<blocks_for(nlines, 128), 128>>>(d_text, n, start.as<long long>(), nlines, kind.as<unsigned char>(), offs.as<ObjCounts>(), total,
                                                  pv.as<double>(), pvt.as<double>(), pvn.as<double>(), ckey.as<unsigned long long>(), tric.as<int>(), err.as<ErrSlot>());
    LJ_CUDA(cudaMemcpy(&es, err.p, sizeof es, cudaMemcpyDeviceToHost));
    if (es.packed) return refuse(es);

    // de-duplication in order of first appearance
    DevBuf seq, skey, sseq, first, frank, cid, cnt;
    LJ_CUDA(seq.alloc(nc * 4)); LJ_CUDA(skey.alloc(nc * 8)); LJ_CUDA(sseq.alloc(nc * 4)); LJ_CUDA(first.alloc((nc + 1) * 4)); LJ_CUDA(frank.alloc((nc + 1) * 4));
    LJ_CUDA(cid.alloc(nc * 4)); LJ_CUDA(cnt.alloc(8));
    int nvtx = 0;
    if (nc) {
        k_iota<<<blocks_for(nc), 256>>>(seq.as<unsigned>(), nc);
        LJ_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tb, ckey.as<unsigned long long>(), skey.as<unsigned long long>(), seq.as<unsigned>(), sseq.as<unsigned>(), (int)nc, 0, 3 * OBJ_KEY_BITS));
        LJ_CUDA(cudaMemsetAsync(first.p, 0, (nc + 1) * 4, 0));
        k_obj_mark_first<<<blocks_for(nc), 256>>>(skey.as<unsigned long long>(), sseq.as<unsigned>(), nc, first.as<int>());
        LJ_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, first.as<int>(), frank.as<int>(), (int)(nc + 1)));
        LJ_CUDA(cudaMemcpy(&nvtx, frank.as<int>() + nc, 4, cudaMemcpyDeviceToHost));
    }
    DevBuf pos, uv, nor, idx;
    LJ_CUDA(pos.alloc((size_t)nvtx * 24)); LJ_CUDA(uv.alloc((size_t)nvtx * 16)); LJ_CUDA(nor.alloc((size_t)nvtx * 24)); LJ_CUDA(idx.alloc(ntri * 12));
    LJ_CUDA(cudaMemsetAsync(cnt.p, 0, 8, 0));
    if (nc) {
        k_obj_emit_vertices<<<blocks_for(nc), 256>>>(skey.as<unsigned long long>(), sseq.as<unsigned>(), nc, frank.as<int>(), pv.as<double>(), pvt.as<double>(), pvn.as<double>(),
                                                     xf.as<double>(), xf.as<double>() + 16, pos.as<double>(), uv.as<double>(), nor.as<double>(), cid.as<int>(),
                                                     cnt.as<int>(), cnt.as<int>() + 1);
        k_obj_emit_triangles<<<blocks_for(3 * ntri), 256>>>(tric.as<int>(), cid.as<int>(), 3 * ntri, idx.as<int32_t>());
    }
    int with[2] = {0, 0};
    LJ_CUDA(cudaMemcpy(with, cnt.p, 8, cudaMemcpyDeviceToHost));
    if ((with[0] != 0 && with[0] != nvtx) || (with[1] != 0 && with[1] != nvtx)) {
        // the reference push_backs uvs / normals only for corners that carry them (parse_obj.cpp:129-133), which
        // misaligns its arrays; such files are refused here
        lj_set_error("lj_mesh_load_obj: %s", obj_error_text(OBJ_E_MIXED)); return LJ_EUNSUPPORTED;
    }
    const bool has_uv = with[0] != 0;
    bool has_n = with[1] != 0;
    if (!has_n && nvtx) {                                    // parse_obj.cpp:230-232
        if (int rc = compute_normals_device(pos.as<double>(), nvtx, idx.as<int32_t>(), ntri, nor.as<double>())) return rc;
        has_n = true;
    }
    LJ_CUDA(cudaGetLastError());
    out->decode_ms = tm.end();
    cudaError_t e = cudaSuccess;
    out->positions = host_copy<double>(pos.p, (size_t)nvtx * 3, e);
    if (has_n) out->normals = host_copy<double>(nor.p, (size_t)nvtx * 3, e);
    if (has_uv) out->uvs = host_copy<double>(uv.p, (size_t)nvtx * 2, e);
    out->indices = host_copy<int32_t>(idx.p, (size_t)ntri * 3, e);
    if (e != cudaSuccess || !out->positions || !out->indices) { mesh_release(out); lj_set_error("lj_mesh_load_obj: %s", e == cudaSuccess ? "out of host memory" : cudaGetErrorString(e)); return LJ_ECUDA; }
    out->num_vertices = (uint64_t)nvtx; out->num_triangles = (uint64_t)ntri; out->has_normals = has_n; out->has_uvs = has_uv;
    return LJ_OK;
}
