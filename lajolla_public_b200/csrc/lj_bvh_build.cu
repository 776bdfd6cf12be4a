// GPU BVH construction: replaces Embree's rtcCommitScene
// (/root/reference/src/scene.cpp:20-27).  Pipeline, all on the device:
//   1. primitive bounds in f32 from the f64 scene arrays (the reference hands
//      Embree float vertices, shapes/triangle_mesh.inl:11-14, and float sphere
//      bounds, shapes/sphere.inl:1-10) + scene AABB reduction
//   2. 63-bit Morton codes of the centroids, radix sort (cub::DeviceRadixSort)
//   3. binary radix tree (Karras 2012), bottom-up AABB fit
//   4. collapse to an 8-wide tree and quantise into 80-byte nodes (Ylitie 2017)
#include "lj_bvh_build.h"
#include "lj_alloc.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cooperative_groups.h>
#include <cstdlib>
#include <cstring>
#include <cfloat>
#include <cstdio>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { snprintf(err, 256, "%s: %s", #x, cudaGetErrorString(e_)); return -1; } } while (0)

// LJ_BVH_TRACE=1: host wall-clock per build phase (each boundary synchronises, so only for diagnosis)
#include <chrono>
struct BvhPhaseTimer {
    bool on; std::chrono::steady_clock::time_point t0;
    BvhPhaseTimer() : on(getenv("LJ_BVH_TRACE") != nullptr) { if (on) { cudaStreamSynchronize(0); t0 = std::chrono::steady_clock::now(); } }
    void mark(const char *what) {
        if (!on) return;
        cudaStreamSynchronize(0);
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[bvh] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

namespace {

__device__ __forceinline__ void atomic_min_f(float *addr, float v) {
    // order-preserving int mapping for floats
    if (v >= 0) atomicMin((int *)addr, __float_as_int(v)); else atomicMax((unsigned *)addr, __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float *addr, float v) {
    if (v >= 0) atomicMax((int *)addr, __float_as_int(v)); else atomicMin((unsigned *)addr, __float_as_uint(v));
}

// prim i < num_tris: triangle i (global index into `indices`); else sphere i - num_tris
__global__ void k_prim_bounds(int num_tris, int num_spheres, const double *positions, const int *indices,
                              const int *tri_shape, const lj_shape *shapes, const SpherePrim *spheres,
                              float4 *prim_lo, float4 *prim_hi, float *scene_bounds /* lo[3], hi[3] */) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int n = num_tris + num_spheres;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < n) {
        if (i < num_tris) {
            const lj_shape &s = shapes[tri_shape[i]];
            for (int k = 0; k < 3; k++) {
                size_t v = (size_t)s.vertex_offset + indices[3 * (size_t)i + k];
                for (int a = 0; a < 3; a++) {
                    float f = (float)positions[3 * v + a];
                    lo[a] = fminf(lo[a], f); hi[a] = fmaxf(hi[a], f);
                }
            }
        } else {
            const SpherePrim &s = spheres[i - num_tris];
            lo[0] = (float)(s.cx - s.r); lo[1] = (float)(s.cy - s.r); lo[2] = (float)(s.cz - s.r);
            hi[0] = (float)(s.cx + s.r); hi[1] = (float)(s.cy + s.r); hi[2] = (float)(s.cz + s.r);
        }
        prim_lo[i] = make_float4(lo[0], lo[1], lo[2], 0.f);
        prim_hi[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
    // warp reduce, then one atomic per warp
    for (int a = 0; a < 3; a++) {
        float l = lo[a], h = hi[a];
        for (int o = 16; o > 0; o >>= 1) {
            l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o));
            h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o));
        }
        if ((threadIdx.x & 31) == 0 && l <= h) { atomic_min_f(&scene_bounds[a], l); atomic_max_f(&scene_bounds[3 + a], h); }
    }
}

__device__ __forceinline__ unsigned long long expand21(unsigned long long v) {
    v &= 0x1fffffULL;
    v = (v | v << 32) & 0x1f00000000ffffULL;
    v = (v | v << 16) & 0x1f0000ff0000ffULL;
    v = (v | v << 8) & 0x100f00f00f00f00fULL;
    v = (v | v << 4) & 0x10c30c30c30c30c3ULL;
    v = (v | v << 2) & 0x1249249249249249ULL;
    return v;
}

__global__ void k_morton(int n, const float4 *prim_lo, const float4 *prim_hi, const float *scene_bounds,
                         unsigned long long *keys, int *vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 lo = prim_lo[i], hi = prim_hi[i];
    float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
    unsigned long long q[3];
    for (int a = 0; a < 3; a++) {
        float ext = scene_bounds[3 + a] - scene_bounds[a];
        float t = ext > 0 ? (c[a] - scene_bounds[a]) / ext : 0.f;
        t = fminf(fmaxf(t, 0.f), 1.f);
        q[a] = (unsigned long long)fminf(t * 2097152.f, 2097151.f);
    }
    keys[i] = (expand21(q[0]) << 2) | (expand21(q[1]) << 1) | expand21(q[2]);
    vals[i] = i;
}

__device__ __forceinline__ int delta(const unsigned long long *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}

// Karras 2012: internal node i covers a key range; children are internal nodes or leaves.
__global__ void k_radix_tree(int n, const unsigned long long *keys, int *left, int *right, int *parent_inner, int *parent_leaf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int lc, rc;
    if (lo == gamma) { lc = ~gamma; parent_leaf[gamma] = i; } else { lc = gamma; parent_inner[gamma] = i; }
    if (hi == gamma + 1) { rc = ~(gamma + 1); parent_leaf[gamma + 1] = i; } else { rc = gamma + 1; parent_inner[gamma + 1] = i; }
    left[i] = lc; right[i] = rc;
    if (i == 0) parent_inner[0] = -1;
}

// bottom-up AABB fit; second arriving thread at a node continues upward
__global__ void k_fit(int n, const int *vals, const float4 *prim_lo, const float4 *prim_hi, const int *left, const int *right,
                      const int *parent_inner, const int *parent_leaf, int *flags, float4 *node_lo, float4 *node_hi) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int node = parent_leaf[i];
    while (node >= 0) {
        if (atomicAdd(&flags[node], 1) == 0) return;
        __threadfence();
        int lc = left[node], rc = right[node];
        float4 llo = lc < 0 ? prim_lo[vals[~lc]] : __ldcg(&node_lo[lc]);
        float4 lhi = lc < 0 ? prim_hi[vals[~lc]] : __ldcg(&node_hi[lc]);
        float4 rlo = rc < 0 ? prim_lo[vals[~rc]] : __ldcg(&node_lo[rc]);
        float4 rhi = rc < 0 ? prim_hi[vals[~rc]] : __ldcg(&node_hi[rc]);
        node_lo[node] = make_float4(fminf(llo.x, rlo.x), fminf(llo.y, rlo.y), fminf(llo.z, rlo.z), 0.f);
        node_hi[node] = make_float4(fmaxf(lhi.x, rhi.x), fmaxf(lhi.y, rhi.y), fmaxf(lhi.z, rhi.z), 0.f);
        __threadfence();
        node = parent_inner[node];
    }
}

__global__ void k_emit_bvh2(int n, const int *left, const int *right, const float4 *node_lo, const float4 *node_hi, Bvh2Node *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    Bvh2Node nd;
    float4 lo = node_lo[i], hi = node_hi[i];
    nd.lo[0] = lo.x; nd.lo[1] = lo.y; nd.lo[2] = lo.z; nd.left = left[i];
    nd.hi[0] = hi.x; nd.hi[1] = hi.y; nd.hi[2] = hi.z; nd.right = right[i];
    out[i] = nd;
}

// ---------------- triangle pre-splitting (early split clipping, Ernst & Greiner 2007) ----------------
// Large or diagonal triangles have bounding boxes far bigger than the triangle, which makes sibling
// boxes overlap and costs both node visits and primitive tests.  Before the Morton sort every triangle
// whose box is larger than `area_limit` is cut at the midpoint of its box's longest axis, the two halves
// get the tight box of the CLIPPED triangle, and so on (at most 2^SPLIT_MAX_DEPTH pieces).  Each piece
// is a reference to the same triangle: the leaf stores the full triangle, the ray test is the usual
// one, and a triangle found through two references gives the identical (t, geomID, primID), so the
// closest-hit result does not change.  Boxes are computed in f64 from the f32 vertices and rounded
// outward to f32, so the union of a triangle's references always covers it.
constexpr int SPLIT_MAX_DEPTH = 6;

struct SplitBox { double lo[3], hi[3]; };

// tight bounds of (triangle clipped to box b); returns false if the intersection is empty
__device__ bool clipped_bounds(const double tri[3][3], const SplitBox &b, SplitBox &out) {
    double poly[2][10][3];
    int np = 3, cur = 0;
    for (int k = 0; k < 3; k++) for (int a = 0; a < 3; a++) poly[0][k][a] = tri[k][a];
    for (int axis = 0; axis < 3 && np > 0; axis++) {
        for (int side = 0; side < 2 && np > 0; side++) {
            const double plane = side ? b.hi[axis] : b.lo[axis];
            const double sgn = side ? -1.0 : 1.0;            // inside: sgn * (x - plane) >= 0
            int nq = 0;
            for (int k = 0; k < np; k++) {
                const double *A = poly[cur][k], *B = poly[cur][(k + 1) % np];
                const double da = sgn * (A[axis] - plane), db = sgn * (B[axis] - plane);
                if (da >= 0) { for (int a = 0; a < 3; a++) poly[1 - cur][nq][a] = A[a]; nq++; }
                if ((da > 0 && db < 0) || (da < 0 && db > 0)) {
                    const double t = da / (da - db);
                    for (int a = 0; a < 3; a++) poly[1 - cur][nq][a] = (a == axis) ? plane : A[a] + t * (B[a] - A[a]);
                    nq++;
                }
            }
            np = nq; cur = 1 - cur;
        }
    }
    if (np == 0) return false;
    for (int a = 0; a < 3; a++) { out.lo[a] = 1e300; out.hi[a] = -1e300; }
    for (int k = 0; k < np; k++) for (int a = 0; a < 3; a++) {
        out.lo[a] = fmin(out.lo[a], poly[cur][k][a]); out.hi[a] = fmax(out.hi[a], poly[cur][k][a]);
    }
    // interpolated points carry rounding error: pad by a relative 1e-12 of the box scale, stay inside the cell
    for (int a = 0; a < 3; a++) {
        const double pad = 1e-12 * (fabs(out.lo[a]) + fabs(out.hi[a]) + (b.hi[a] - b.lo[a]));
        out.lo[a] = fmax(out.lo[a] - pad, b.lo[a]); out.hi[a] = fmin(out.hi[a] + pad, b.hi[a]);
    }
    return true;
}

__device__ __forceinline__ double box_area(const SplitBox &b) {
    const double dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// EMIT = false: count the references of primitive i; EMIT = true: write them at ref_offset[i]...
template <bool EMIT>
__global__ void k_split_refs(int num_tris, int num_spheres, const double *positions, const int *indices, const int *tri_shape,
                             const lj_shape *shapes, const float4 *prim_lo, const float4 *prim_hi, double area_limit,
                             int *ref_count, const int *ref_offset, float4 *ref_lo, float4 *ref_hi, int *ref_prim) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = num_tris + num_spheres;
    if (i >= n) return;
    int emitted = 0;
    const int base = EMIT ? ref_offset[i] : 0;
    if (i >= num_tris) {                      // spheres are never split
        if (EMIT) { ref_lo[base] = prim_lo[i]; ref_hi[base] = prim_hi[i]; ref_prim[base] = i; }
        else ref_count[i] = 1;
        return;
    }
    double tri[3][3];
    const lj_shape &s = shapes[tri_shape[i]];
    for (int k = 0; k < 3; k++) {
        const size_t v = (size_t)s.vertex_offset + indices[3 * (size_t)i + k];
        for (int a = 0; a < 3; a++) tri[k][a] = (double)(float)positions[3 * v + a];
    }
    SplitBox stack[SPLIT_MAX_DEPTH + 2]; int depth[SPLIT_MAX_DEPTH + 2]; int sp = 0;
    {
        const float4 l = prim_lo[i], h = prim_hi[i];
        stack[0].lo[0] = l.x; stack[0].lo[1] = l.y; stack[0].lo[2] = l.z;
        stack[0].hi[0] = h.x; stack[0].hi[1] = h.y; stack[0].hi[2] = h.z;
        depth[0] = 0; sp = 1;
    }
    while (sp > 0) {
        const SplitBox b = stack[--sp]; const int d = depth[sp];
        const double ex[3] = {b.hi[0] - b.lo[0], b.hi[1] - b.lo[1], b.hi[2] - b.lo[2]};
        const int axis = ex[0] >= ex[1] ? (ex[0] >= ex[2] ? 0 : 2) : (ex[1] >= ex[2] ? 1 : 2);
        bool leaf = d >= SPLIT_MAX_DEPTH || box_area(b) <= area_limit || !(ex[axis] > 0);
        SplitBox l, r;
        if (!leaf) {
            const double mid = 0.5 * (b.lo[axis] + b.hi[axis]);
            SplitBox cl = b, cr = b; cl.hi[axis] = mid; cr.lo[axis] = mid;
            const bool okl = clipped_bounds(tri, cl, l), okr = clipped_bounds(tri, cr, r);
            if (!okl || !okr) leaf = true;    // numerically empty half: keep the piece whole
        }
        if (leaf) {
            if (EMIT) {
                // round outward to f32
                float lo[3], hi[3];
                for (int a = 0; a < 3; a++) {
                    lo[a] = __double2float_rd(b.lo[a]); hi[a] = __double2float_ru(b.hi[a]);
                }
                ref_lo[base + emitted] = make_float4(lo[0], lo[1], lo[2], 0.f);
                ref_hi[base + emitted] = make_float4(hi[0], hi[1], hi[2], 0.f);
                ref_prim[base + emitted] = i;
            }
            emitted++;
        } else {
            stack[sp] = r; depth[sp] = d + 1; sp++;
            stack[sp] = l; depth[sp] = d + 1; sp++;
        }
    }
    if (!EMIT) ref_count[i] = emitted;
}

// primitive slots in leaf (sorted) order: 3 float4 each
__global__ void k_emit_prims(int n, int num_tris, const int *vals, const int *ref_prim, const double *positions, const int *indices,
                             const int *tri_shape, const lj_shape *shapes, const SpherePrim *spheres, float4 *tri_verts) {
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n) return;
    int i = ref_prim ? ref_prim[vals[slot]] : vals[slot];
    if (i < num_tris) {
        int sid = tri_shape[i];
        const lj_shape &s = shapes[sid];
        float p[9];
        for (int k = 0; k < 3; k++) {
            size_t v = (size_t)s.vertex_offset + indices[3 * (size_t)i + k];
            for (int a = 0; a < 3; a++) p[3 * k + a] = (float)positions[3 * v + a];
        }
        tri_verts[3 * slot + 0] = make_float4(p[0], p[1], p[2], __int_as_float(sid));
        tri_verts[3 * slot + 1] = make_float4(p[3], p[4], p[5], __int_as_float(i - (int)s.triangle_offset));
        tri_verts[3 * slot + 2] = make_float4(p[6], p[7], p[8], __int_as_float(0));
    } else {
        int si = i - num_tris;
        tri_verts[3 * slot + 0] = make_float4(__int_as_float(si), 0.f, 0.f, __int_as_float(spheres[si].geom));
        tri_verts[3 * slot + 1] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
        tri_verts[3 * slot + 2] = make_float4(0.f, 0.f, 0.f, __int_as_float(1));
    }
}

// ---------------- PLOC (parallel locally-ordered clustering, Meister & Bittner 2018) ----------------
// Clusters start as the Morton-sorted primitives.  Each round every cluster looks at its
// PLOC_R neighbours on either side (in the current order) for the partner with the smallest
// merged surface area; mutually-nearest pairs are merged into a new inner node, the array is
// compacted (order preserved), and the rounds repeat until one cluster is left.
constexpr int PLOC_R = 100;         // default search radius (sponza: 24 -> 100 = 15.0 -> 14.1 nodes, 9.8 -> 9.1 primitives per ray); LJ_PLOC_R overrides it, up to PLOC_RMAX
constexpr int PLOC_RMAX = 256;
constexpr int PLOC_B = 256;

__device__ __forceinline__ float union_area(const float4 &alo, const float4 &ahi, const float4 &blo, const float4 &bhi) {
    float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x);
    float dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y);
    float dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_ploc_init(int n, const int *vals, const float4 *prim_lo, const float4 *prim_hi,
                            int *cid, int *ccount, float4 *clo, float4 *chi) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int p = vals[i];
    cid[i] = ~i; ccount[i] = 1; clo[i] = prim_lo[p]; chi[i] = prim_hi[p];
}

// The cluster count `m` of a round lives on the device (*m_ptr): the host sizes grids from the last count it read
// (an upper bound, counts only fall) and reads the count back every few rounds instead of after each one.
__global__ void __launch_bounds__(PLOC_B) k_ploc_nn(const int *m_ptr, const float4 *clo, const float4 *chi, int *nn, const int PLOC_R) {
    __shared__ float4 slo[PLOC_B + 2 * PLOC_RMAX], shi[PLOC_B + 2 * PLOC_RMAX];
    const int m = *m_ptr;
    if (blockIdx.x * PLOC_B >= m) return;
    int base = blockIdx.x * PLOC_B - PLOC_R;
    for (int t = threadIdx.x; t < PLOC_B + 2 * PLOC_R; t += PLOC_B) {
        int g = base + t;
        if (g >= 0 && g < m) { slo[t] = clo[g]; shi[t] = chi[g]; }
    }
    __syncthreads();
    int i = blockIdx.x * PLOC_B + threadIdx.x;
    if (i >= m) return;
    int li = threadIdx.x + PLOC_R;
    float4 lo = slo[li], hi = shi[li];
    float best = FLT_MAX; int bj = -1;
    for (int o = -PLOC_R; o <= PLOC_R; o++) {
        int j = i + o;
        if (o == 0 || j < 0 || j >= m) continue;
        float a = union_area(lo, hi, slo[li + o], shi[li + o]);
        if (a < best) { best = a; bj = j; }     // ties keep the smaller j
    }
    nn[i] = bj;
}

__global__ void k_ploc_merge(const int *m_ptr, int m_upper, const int *nn, int *cid, int *ccount, float4 *clo, float4 *chi, int *valid,
                             int *left, int *right, int *count, float4 *node_lo, float4 *node_hi, int *node_counter) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int m = *m_ptr;
    if (i >= m) { if (i <= m_upper) valid[i] = 0; return; }      // the scan runs over m_upper + 1 flags
    int j = nn[i];
    int v = 1;
    if (j >= 0 && nn[j] == i) {
        if (i < j) {
            int id = atomicAdd(node_counter, 1);
            float4 lo = clo[i], hi = chi[i], lo2 = clo[j], hi2 = chi[j];
            lo = make_float4(fminf(lo.x, lo2.x), fminf(lo.y, lo2.y), fminf(lo.z, lo2.z), 0.f);
            hi = make_float4(fmaxf(hi.x, hi2.x), fmaxf(hi.y, hi2.y), fmaxf(hi.z, hi2.z), 0.f);
            left[id] = cid[i]; right[id] = cid[j];
            int c = ccount[i] + ccount[j];
            count[id] = c; node_lo[id] = lo; node_hi[id] = hi;
            cid[i] = id; ccount[i] = c; clo[i] = lo; chi[i] = hi;
        } else v = 0;
    }
    valid[i] = v;
}

__global__ void k_ploc_compact(const int *m_ptr, int m_upper, int *m_next, const int *valid, const int *pos, const int *cid, const int *ccount,
                               const float4 *clo, const float4 *chi, int *cid2, int *ccount2, float4 *clo2, float4 *chi2) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int m = *m_ptr;
    if (i == 0) *m_next = pos[m_upper];       // clusters left after this round
    if (i >= m || !valid[i]) return;
    int p = pos[i];
    cid2[p] = cid[i]; ccount2[p] = ccount[i]; clo2[p] = clo[i]; chi2[p] = chi[i];
}

// The rounds above as ONE cooperative launch (the default): the same search, merge rule and order-preserving compaction,
// with three grid barriers per round in place of five launches and a host read-back every few rounds.  A round is
//   A  nearest neighbour of every cluster (chunks of PLOC_B clusters, the chunk + 2 * radius boxes staged in shared memory)
//   B  merge decision per cluster, exclusive scan of (kept, created) inside the chunk, chunk totals to `chunk_sums`
//   C  every chunk adds the totals of the chunks before it, creates its nodes and writes the kept clusters to the other buffer
// Node ids come from the scan, so the tree is the same from run to run (the launch sequence takes them from an atomic).
// Stops once at most `stop_at` clusters are left (k_ploc_tail finishes from there) and leaves them in buffer 0.
// counter: [0] next node id, [1] clusters left, [2] rounds done, [3] 1 = a round made no progress.
struct PlocBufs { int *cid[2], *cc[2]; float4 *clo[2], *chi[2]; };
__global__ void __launch_bounds__(PLOC_B) k_ploc_rounds(PlocBufs b, int *nn, int *lpos, unsigned long long *chunk_sums, int *counter,
                                                        int *left, int *right, int *count, float4 *node_lo, float4 *node_hi,
                                                        const int radius, const int stop_at, unsigned long long *dbg /* LJ_BVH_TRACE: phase time stamps */) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    __shared__ float4 slo[PLOC_B + 2 * PLOC_RMAX], shi[PLOC_B + 2 * PLOC_RMAX];
    __shared__ int warp_sums[PLOC_B / 32];
    __shared__ unsigned long long s_pre[2];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    int m = counter[1], next_id = counter[0], cur = 0, rounds = 0;
    bool stuck = false;
    auto stamp = [&](int phase) {
        if (dbg && blockIdx.x == 0 && tid == 0 && rounds < 64) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); dbg[rounds * 4 + phase] = t; }
    };
    while (m > stop_at) {
        const int chunks = (m + PLOC_B - 1) / PLOC_B;
        const float4 *clo = b.clo[cur], *chi = b.chi[cur];
        stamp(0);
        // ---- A: nearest neighbours ----
        // T = 2^k threads per cluster once the grid has spare threads (as in k_ploc_tail): a block searches for PLOC_B / T
        // clusters, each thread takes every T-th candidate, the (area, index) minimum is the sequential scan's result
        int T = 1;
        while (T < 32 && (long long)2 * T * m <= (long long)gridDim.x * PLOC_B) T <<= 1;
        const int per_block = PLOC_B / T, chunks_a = (m + per_block - 1) / per_block;
        for (int c = blockIdx.x; c < chunks_a; c += gridDim.x) {
            const int first = c * per_block, base = first - radius;
            __syncthreads();                        // the previous chunk's searches are over
            for (int t = tid; t < per_block + 2 * radius; t += PLOC_B) {
                const int g = base + t;
                if (g >= 0 && g < m) { slo[t] = clo[g]; shi[t] = chi[g]; }
            }
            __syncthreads();
            const int i = first + tid / T, sub = tid & (T - 1);
            float best = FLT_MAX; int bj = -1;
            if (i < m) {
                const int li = tid / T + radius;
                const float4 lo = slo[li], hi = shi[li];
                const int o0 = i - radius < 0 ? -i : -radius, o1 = i + radius >= m ? m - 1 - i : radius;
                for (int o = o0 + sub; o <= o1; o += T) {
                    if (o == 0) continue;
                    const float a = union_area(lo, hi, slo[li + o], shi[li + o]);
                    if (a < best) { best = a; bj = i + o; }     // ties keep the smaller j
                }
            }
            for (int o = T >> 1; o > 0; o >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
                if (oj >= 0 && (bj < 0 || ob < best || (ob == best && oj < bj))) { best = ob; bj = oj; }
            }
            if (i < m && sub == 0) nn[i] = bj;
        }
        grid.sync();
        stamp(1);
        // ---- B: merge decisions, scan inside the chunk ----
        for (int c = blockIdx.x; c < chunks; c += gridDim.x) {
            const int i = c * PLOC_B + tid;
            int keep = 0, creates = 0;
            if (i < m) {
                const int j = nn[i];
                keep = 1;
                if (j >= 0 && nn[j] == i) { if (i < j) creates = 1; else keep = 0; }
            }
            int x = keep | (creates << 16);         // both counts stay below 2^16 inside a chunk
            const int v = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
            __syncthreads();                        // warp_sums of the previous chunk has been read
            if (lane == 31) warp_sums[w] = x;
            __syncthreads();
            int before = 0, total = 0;
#pragma unroll
            for (int k = 0; k < PLOC_B / 32; k++) { const int ws = warp_sums[k]; if (k < w) before += ws; total += ws; }
            if (i < m) lpos[i] = keep ? (int)((unsigned)(before + x - v) | (creates ? 0x80000000u : 0u)) : -1;      // creates lives in bit 31, created-before in bits 16..30
            if (tid == 0) chunk_sums[c] = (unsigned long long)(total & 0xffff) | ((unsigned long long)(total >> 16) << 32);
        }
        grid.sync();
        stamp(2);
        // ---- C: global positions, node creation, compaction into the other buffer ----
        const int *cid = b.cid[cur], *cc = b.cc[cur];
        int *cid2 = b.cid[cur ^ 1], *cc2 = b.cc[cur ^ 1];
        float4 *clo2 = b.clo[cur ^ 1], *chi2 = b.chi[cur ^ 1];
        unsigned long long tot = 0;
        for (int c = blockIdx.x; c < chunks; c += gridDim.x) {
            __syncthreads();                        // s_pre of the previous chunk has been read
            if (w == 0) {
                unsigned long long pre = 0, all = 0;
                for (int k = lane; k < chunks; k += 32) { const unsigned long long cs = chunk_sums[k]; all += cs; if (k < c) pre += cs; }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) { pre += __shfl_xor_sync(0xffffffffu, pre, o); all += __shfl_xor_sync(0xffffffffu, all, o); }
                if (lane == 0) { s_pre[0] = pre; s_pre[1] = all; }
            }
            __syncthreads();
            const unsigned long long pre = s_pre[0];
            tot = s_pre[1];
            const int i = c * PLOC_B + tid;
            if (i < m) {
                const int lp = lpos[i];
                if (lp != -1) {
                    const int p = (int)(pre & 0xffffffffu) + (lp & 0xffff);
                    int id = cid[i], cnt = cc[i];
                    float4 lo = clo[i], hi = chi[i];
                    if (lp < 0) {                   // bit 31: this cluster absorbs its partner
                        const int j = nn[i];
                        const float4 lo2 = clo[j], hi2 = chi[j];
                        const int nid = next_id + (int)(pre >> 32) + ((lp >> 16) & 0x7fff);
                        lo = make_float4(fminf(lo.x, lo2.x), fminf(lo.y, lo2.y), fminf(lo.z, lo2.z), 0.f);
                        hi = make_float4(fmaxf(hi.x, hi2.x), fmaxf(hi.y, hi2.y), fmaxf(hi.z, hi2.z), 0.f);
                        left[nid] = id; right[nid] = cid[j];
                        cnt += cc[j];
                        count[nid] = cnt; node_lo[nid] = lo; node_hi[nid] = hi;
                        id = nid;
                    }
                    cid2[p] = id; cc2[p] = cnt; clo2[p] = lo; chi2[p] = hi;
                }
            }
        }
        if (blockIdx.x >= chunks) {                 // a block without a chunk this round still needs the totals
            if (w == 0) {
                unsigned long long all = 0;
                for (int k = lane; k < chunks; k += 32) all += chunk_sums[k];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) all += __shfl_xor_sync(0xffffffffu, all, o);
                if (lane == 0) s_pre[1] = all;
            }
            __syncthreads();
            tot = s_pre[1];
            __syncthreads();
        }
        const int m_new = (int)(tot & 0xffffffffu);
        next_id += (int)(tot >> 32);
        stamp(3);
        rounds++;
        cur ^= 1;
        if (m_new >= m || m_new < 1 || rounds > 4096) { stuck = true; break; }     // uniform over the grid
        m = m_new;
        grid.sync();
    }
    if (!stuck && cur == 1) {                       // hand the remaining (<= stop_at) clusters over in buffer 0
        // every block has passed the last barrier of the loop, so buffer 1 is complete; only block 0 copies
        if (blockIdx.x == 0)
            for (int t = tid; t < m; t += PLOC_B) { b.cid[0][t] = b.cid[1][t]; b.cc[0][t] = b.cc[1][t]; b.clo[0][t] = b.clo[1][t]; b.chi[0][t] = b.chi[1][t]; }
    }
    if (blockIdx.x == 0 && tid == 0) { counter[0] = next_id; counter[1] = m; counter[2] = rounds; counter[3] = stuck ? 1 : 0; }
}

// All remaining PLOC rounds in ONE block once at most PLOC_TAIL clusters are left: the same search, merge rule
// (mutually nearest pairs, ties keep the smaller index) and order-preserving compaction as the kernels above, with the
// clusters in shared memory and block-wide scans instead of a launch sequence + a blocking read-back per round.  The
// tail used to be most of the build's ~600 launches (66 k primitives: ~9 rounds get below 1024 clusters, ~60 follow).
constexpr int PLOC_TAIL = 1024;
// the cooperative rounds hand over at this many clusters: a round costs them ~13 us whatever is left (three grid barriers),
// the one-block tail ~18 us per round at 1024 clusters (200 candidates per thread on ONE SM) but 2-5 us below 256
constexpr int PLOC_COOP_STOP = 256;
__device__ __forceinline__ int block_scan_excl_1024(int v, int *warp_sums, int *total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_sums[w] = x;
    __syncthreads();
    if (w == 0) {
        int s = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
        warp_sums[lane] = s;            // inclusive sums of the warp totals
    }
    __syncthreads();
    const int before = w > 0 ? warp_sums[w - 1] : 0;
    *total = warp_sums[31];
    __syncthreads();                    // warp_sums is reused by the next call
    return before + x - v;
}
__global__ void __launch_bounds__(PLOC_TAIL) k_ploc_tail(const int *m_ptr, int *cid, const int *ccount, const float4 *clo, const float4 *chi,
                                                         int *left, int *right, int *count, float4 *node_lo, float4 *node_hi,
                                                         int *node_counter, const int radius) {
    __shared__ float4 slo[PLOC_TAIL], shi[PLOC_TAIL];
    __shared__ int sid[PLOC_TAIL], scnt[PLOC_TAIL], snn[PLOC_TAIL], warp_sums[32];
    const int t = threadIdx.x;
    int m = *m_ptr;
    if (t < m) { slo[t] = clo[t]; shi[t] = chi[t]; sid[t] = cid[t]; scnt[t] = ccount[t]; }
    int next_id = *node_counter;
    __syncthreads();
    while (m > 1) {
        // search: T = 2^k threads per cluster once the block has spare threads (T <= 32 divides the warp size, so a
        // cluster's threads share a warp); each takes every T-th candidate, the (area, index) minimum is the sequential
        // scan's result: smallest area, ties to the smaller index
        int T = 1;
        while (T < 32 && 2 * T * m <= PLOC_TAIL) T <<= 1;
        {
            const int c = t / T, sub = t & (T - 1);
            float best = FLT_MAX; int sj = -1;
            if (c < m) {
                const float4 lo = slo[c], hi = shi[c];
                const int j0 = c - radius < 0 ? 0 : c - radius, j1 = c + radius >= m ? m - 1 : c + radius;
                for (int j = j0 + sub; j <= j1; j += T) {
                    if (j == c) continue;
                    const float a = union_area(lo, hi, slo[j], shi[j]);
                    if (a < best) { best = a; sj = j; }
                }
            }
            for (int o = T >> 1; o > 0; o >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oj = __shfl_xor_sync(0xffffffffu, sj, o);
                if (oj >= 0 && (sj < 0 || ob < best || (ob == best && oj < sj))) { best = ob; sj = oj; }
            }
            if (c < m && sub == 0) snn[c] = sj;
        }
        __syncthreads();
        const int bj = t < m ? snn[t] : -1;
        int keep = t < m ? 1 : 0, creates = 0;
        float4 lo, hi; int id_l = 0, id_r = 0, c = 0;
        if (t < m) {
            lo = slo[t]; hi = shi[t]; id_l = sid[t]; c = scnt[t];
            if (bj >= 0 && snn[bj] == t) {
                if (t < bj) {
                    creates = 1;
                    const float4 lo2 = slo[bj], hi2 = shi[bj];
                    lo = make_float4(fminf(lo.x, lo2.x), fminf(lo.y, lo2.y), fminf(lo.z, lo2.z), 0.f);
                    hi = make_float4(fmaxf(hi.x, hi2.x), fmaxf(hi.y, hi2.y), fmaxf(hi.z, hi2.z), 0.f);
                    id_r = sid[bj]; c += scnt[bj];
                } else keep = 0;
            }
        }
        int totals;
        const int packed = block_scan_excl_1024(keep | (creates << 16), warp_sums, &totals);    // both counts stay below 2^16
        // (the scan's barriers also order the reads above before the compacted writes below)
        if (keep) {
            const int p = packed & 0xffff;
            int id = id_l;
            if (creates) {
                id = next_id + (packed >> 16);
                left[id] = id_l; right[id] = id_r; count[id] = c; node_lo[id] = lo; node_hi[id] = hi;
            }
            slo[p] = lo; shi[p] = hi; sid[p] = id; scnt[p] = c;
        }
        m = totals & 0xffff;
        next_id += totals >> 16;
        __syncthreads();
    }
    if (t == 0) { cid[0] = sid[0]; *node_counter = next_id; }
}

// counts for the radix tree (LBVH path): size of the leaf range under each inner node
__global__ void k_lbvh_counts(int n, const int *left, const int *right, int *count) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int c = i; while (c >= 0) c = left[c];
    int first = ~c;
    c = i; while (c >= 0) c = right[c];
    count[i] = ~c - first + 1;
}

} // namespace

// d_info (device): [0] receives the root id, [1] is set to 1 when a round made no progress (read back by the caller
// together with the collapse counters: the build blocks on the device once, at its end).
static int ploc_build(int n, const int *vals, const float4 *prim_lo, const float4 *prim_hi,
                      int *left, int *right, int *count, float4 *node_lo, float4 *node_hi, int *d_info, char *err) {
    int *cid[2] = {nullptr, nullptr}, *cc[2] = {nullptr, nullptr}, *nn = nullptr, *valid = nullptr, *pos = nullptr, *counter = nullptr;
    float4 *clo[2] = {nullptr, nullptr}, *chi[2] = {nullptr, nullptr};
    unsigned long long *chunk_sums = nullptr;
    void *tmp = nullptr; size_t tmp_bytes = 0;
    for (int b = 0; b < 2; b++) {
        CK(lj_malloc(&cid[b], 4 * (size_t)n)); CK(lj_malloc(&cc[b], 4 * (size_t)n));
        CK(lj_malloc(&clo[b], sizeof(float4) * n)); CK(lj_malloc(&chi[b], sizeof(float4) * n));
    }
    CK(lj_malloc(&nn, 4 * (size_t)n)); CK(lj_malloc(&valid, 4 * (size_t)(n + 1))); CK(lj_malloc(&pos, 4 * (size_t)(n + 1)));
    CK(lj_malloc(&counter, 4 * 4));             // launch sequence: [0] next node id, [1], [2] cluster count before / after a round (alternating); cooperative: see k_ploc_rounds
    {
        const int init[4] = {0, n, n, 0};
        CK(cudaMemcpyAsync(counter, init, sizeof init, cudaMemcpyHostToDevice, 0));      // (pageable source: staged before the call returns)
    }
    k_ploc_init<<<(n + 255) / 256, 256>>>(n, vals, prim_lo, prim_hi, cid[0], cc[0], clo[0], chi[0]);
    int radius = PLOC_R;
    if (const char *e = getenv("LJ_PLOC_R")) { int v = atoi(e); if (v >= 1 && v <= PLOC_RMAX) radius = v; }
    const bool use_tail = !getenv("LJ_PLOC_NO_TAIL");
    const bool trace = getenv("LJ_BVH_TRACE") != nullptr;
    bool coop = use_tail;                       // LJ_PLOC_COOP=0: the launch sequence with a read-back every LJ_PLOC_BATCH rounds
    if (const char *e = getenv("LJ_PLOC_COOP")) coop = coop && atoi(e) != 0;
    if (coop) {                                 // a device (or a partition of one) without cooperative launches gets the launch sequence
        int dev = 0, ok = 0;
        CK(cudaGetDevice(&dev));
        CK(cudaDeviceGetAttribute(&ok, cudaDevAttrCooperativeLaunch, dev));
        coop = ok != 0;
    }
    int cur = 0;
    if (coop) {
        const int chunks = (n + PLOC_B - 1) / PLOC_B;
        if (n > PLOC_TAIL) {
            int per_sm = 0, dev = 0, sm_count = 0;
            CK(cudaGetDevice(&dev));
            CK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ploc_rounds, PLOC_B, 0));
            int grid = per_sm * sm_count;
            if (grid < 1) { snprintf(err, 256, "PLOC: no resident block for the cooperative rounds"); return -1; }
            // blocks beyond one per chunk still help: the search of a round spreads over all threads of the grid
            int want = sm_count * 4;
            if (const char *e = getenv("LJ_PLOC_GRID")) { int v = atoi(e); if (v >= 1) want = v; }
            if (want < 1) want = 1;
            if (grid > want) grid = want;
            CK(lj_malloc(&chunk_sums, 8 * (size_t)chunks));
            PlocBufs bufs;
            for (int b = 0; b < 2; b++) { bufs.cid[b] = cid[b]; bufs.cc[b] = cc[b]; bufs.clo[b] = clo[b]; bufs.chi[b] = chi[b]; }
            int *lpos = valid; int stop_at = PLOC_COOP_STOP; int rad = radius;
            if (const char *e = getenv("LJ_PLOC_STOP")) { int v = atoi(e); if (v >= 1 && v <= PLOC_TAIL) stop_at = v; }
            unsigned long long *dbg = nullptr;
            if (trace) { CK(lj_malloc(&dbg, 8 * 256)); CK(cudaMemsetAsync(dbg, 0, 8 * 256, 0)); }
            void *args[] = {&bufs, &nn, &lpos, &chunk_sums, &counter, &left, &right, &count, &node_lo, &node_hi, &rad, &stop_at, &dbg};
            CK(cudaLaunchCooperativeKernel((const void *)k_ploc_rounds, dim3(grid), dim3(PLOC_B), args, 0, 0));
            if (trace) {
                unsigned long long h[256];
                CK(cudaMemcpy(h, dbg, sizeof h, cudaMemcpyDeviceToHost));
                lj_free(dbg);
                double a = 0, b2 = 0, c = 0, nxt = 0; int nr = 0;
                for (int r = 0; r < 64 && h[4 * r + 3]; r++, nr++) {
                    a += (double)(h[4 * r + 1] - h[4 * r]); b2 += (double)(h[4 * r + 2] - h[4 * r + 1]); c += (double)(h[4 * r + 3] - h[4 * r + 2]);
                    if (r + 1 < 64 && h[4 * r + 4]) nxt += (double)(h[4 * r + 4] - h[4 * r + 3]);
                }
                if (nr) fprintf(stderr, "[bvh]   PLOC rounds, block 0, us per round: search + barrier %.2f, decide + barrier %.2f, compact %.2f, closing barrier %.2f (grid %d)\n",
                                a / nr * 1e-3, b2 / nr * 1e-3, c / nr * 1e-3, nxt / nr * 1e-3, grid);
            }
        }
        std::chrono::steady_clock::time_point t_a, t_b;
        int h[4] = {0, 0, 0, 0};
        if (trace) {
            CK(cudaStreamSynchronize(0)); t_a = std::chrono::steady_clock::now();
            CK(cudaMemcpy(h, counter, sizeof h, cudaMemcpyDeviceToHost));
            t_b = std::chrono::steady_clock::now();
        }
        k_ploc_tail<<<1, PLOC_TAIL>>>(counter + 1, cid[0], cc[0], clo[0], chi[0], left, right, count, node_lo, node_hi, counter, radius);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(d_info + 1, counter + 3, 4, cudaMemcpyDeviceToDevice, 0));
        if (trace) {
            CK(cudaStreamSynchronize(0));
            const double tail_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_b).count();
            fprintf(stderr, "[bvh]   PLOC: %d cooperative rounds down to %d clusters, then the one-block tail (%.3f ms)\n", n > PLOC_TAIL ? h[2] : 0, h[1], tail_ms);
        }
    } else {
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, valid, pos, n + 1));
        CK(lj_malloc(&tmp, tmp_bytes));
        int m = n, rounds = 0;              // m: the last count the host has seen
        int batch = 4;                      // rounds between two read-backs of the cluster count
        if (const char *e = getenv("LJ_PLOC_BATCH")) { int v = atoi(e); if (v >= 1 && v <= 64) batch = v; }
        while (m > 1) {
            int *m_in = counter + 1 + (rounds & 1);
            if (use_tail && m <= PLOC_TAIL) {
                k_ploc_tail<<<1, PLOC_TAIL>>>(m_in, cid[cur], cc[cur], clo[cur], chi[cur], left, right, count, node_lo, node_hi, counter, radius);
                CK(cudaGetLastError());
                break;
            }
            const int m_upper = m;
            for (int b = 0; b < batch; b++) {
                m_in = counter + 1 + (rounds & 1);
                int *m_out = counter + 1 + ((rounds + 1) & 1);
                k_ploc_nn<<<(m_upper + PLOC_B - 1) / PLOC_B, PLOC_B>>>(m_in, clo[cur], chi[cur], nn, radius);
                k_ploc_merge<<<(m_upper + 1 + 255) / 256, 256>>>(m_in, m_upper, nn, cid[cur], cc[cur], clo[cur], chi[cur], valid, left, right, count, node_lo, node_hi, counter);
                CK(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, valid, pos, m_upper + 1));
                k_ploc_compact<<<(m_upper + 255) / 256, 256>>>(m_in, m_upper, m_out, valid, pos, cid[cur], cc[cur], clo[cur], chi[cur], cid[1 - cur], cc[1 - cur], clo[1 - cur], chi[1 - cur]);
                cur = 1 - cur;
                if (++rounds > 4096) { snprintf(err, 256, "PLOC did not converge"); return -1; }
            }
            int m_new = 0;
            CK(cudaMemcpy(&m_new, counter + 1 + (rounds & 1), 4, cudaMemcpyDeviceToHost));
            if (m_new >= m || m_new < 1) { snprintf(err, 256, "PLOC made no progress (%d -> %d clusters)", m, m_new); return -1; }
            if (trace) fprintf(stderr, "[bvh]   PLOC after round %d: %d -> %d clusters\n", rounds, m, m_new);
            m = m_new;
        }
    }
    CK(cudaMemcpyAsync(d_info, cid[cur], 4, cudaMemcpyDeviceToDevice, 0));
    for (int b = 0; b < 2; b++) { lj_free(cid[b]); lj_free(cc[b]); lj_free(clo[b]); lj_free(chi[b]); }
    lj_free(nn); lj_free(valid); lj_free(pos); lj_free(counter); lj_free(tmp); lj_free(chunk_sums);
    return 0;
}

static int lj_bvh_build_refs(const BvhBuildInput &in, BvhBuildOutput &out, char *err, int n, float4 *prim_lo, float4 *prim_hi,
                             int *ref_prim, float *bounds);

int lj_bvh_build(const BvhBuildInput &in, BvhBuildOutput &out, char *err) {
    const int n = in.num_tris + in.num_spheres;
    memset(&out, 0, sizeof out);
    out.num_prims = n;
    if (n == 0) return 0;
    const int B = 256;
    const int G = (n + B - 1) / B;
    BvhPhaseTimer pt;
    float4 *prim_lo = nullptr, *prim_hi = nullptr;
    float *bounds = nullptr;
    CK(lj_malloc(&prim_lo, sizeof(float4) * n)); CK(lj_malloc(&prim_hi, sizeof(float4) * n));
    // bounds[0..5]: scene AABB; bounds[8..15] reinterpreted as ints: the results the host reads back ONCE at the end of the
    // build (root id, PLOC status, collapse counters), see lj_bvh_build_refs
    CK(lj_malloc(&bounds, sizeof(float) * 16));
    float init_b[16] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    CK(cudaMemcpyAsync(bounds, init_b, sizeof init_b, cudaMemcpyHostToDevice, 0));
    k_prim_bounds<<<G, B>>>(in.num_tris, in.num_spheres, in.positions, in.indices, in.tri_shape, in.shapes, in.spheres, prim_lo, prim_hi, bounds);
    CK(cudaGetLastError());
    pt.mark("primitive bounds");

    // ---- triangle pre-splitting: primitives -> references (see k_split_refs) ----
    const int n_prims_in = n;
    int *ref_prim = nullptr;
    {
        // area limit as a fraction of the scene AABB area; 0 disables.  OFF by default: on sponza it trades
        // 9.8 -> 7.0 primitive tests per ray for 15.0 -> 18.0 node visits and is 9 % slower (profiles/r01_experiments.md)
        const char *e = getenv("LJ_SPLIT_ALPHA");
        double alpha = e ? atof(e) : 0.0;
        if (alpha > 0) {                            // the split threshold needs the scene box on the host now
            float hb[6];
            CK(cudaMemcpy(hb, bounds, sizeof hb, cudaMemcpyDeviceToHost));
            for (int a = 0; a < 3; a++) { out.lo[a] = hb[a]; out.hi[a] = hb[3 + a]; }
        }
        const double sx = out.hi[0] - out.lo[0], sy = out.hi[1] - out.lo[1], sz = out.hi[2] - out.lo[2];
        const double scene_area = sx * sy + sy * sz + sz * sx;
        if (alpha > 0 && in.num_tris > 0 && scene_area > 0) {
            int *ref_count = nullptr, *ref_offset = nullptr; void *stmp = nullptr; size_t stmp_bytes = 0;
            CK(lj_malloc(&ref_count, 4 * (size_t)(n + 1))); CK(lj_malloc(&ref_offset, 4 * (size_t)(n + 1)));
            CK(cudaMemset(ref_count, 0, 4 * (size_t)(n + 1)));
            CK(cub::DeviceScan::ExclusiveSum(nullptr, stmp_bytes, ref_count, ref_offset, n + 1));
            CK(lj_malloc(&stmp, stmp_bytes));
            int total = n;
            for (int attempt = 0; attempt < 8; attempt++) {
                k_split_refs<false><<<G, B>>>(in.num_tris, in.num_spheres, in.positions, in.indices, in.tri_shape, in.shapes, prim_lo, prim_hi,
                                              alpha * scene_area, ref_count, nullptr, nullptr, nullptr, nullptr);
                CK(cudaGetLastError());
                CK(cub::DeviceScan::ExclusiveSum(stmp, stmp_bytes, ref_count, ref_offset, n + 1));
                CK(cudaMemcpy(&total, ref_offset + n, 4, cudaMemcpyDeviceToHost));
                if ((double)total <= 2.0 * n) break;        // reference budget: at most twice the primitives
                alpha *= 2;
            }
            if (total > n && (double)total <= 2.0 * n) {
                float4 *ref_lo = nullptr, *ref_hi = nullptr;
                CK(lj_malloc(&ref_lo, sizeof(float4) * (size_t)total)); CK(lj_malloc(&ref_hi, sizeof(float4) * (size_t)total));
                CK(lj_malloc(&ref_prim, 4 * (size_t)total));
                k_split_refs<true><<<G, B>>>(in.num_tris, in.num_spheres, in.positions, in.indices, in.tri_shape, in.shapes, prim_lo, prim_hi,
                                             alpha * scene_area, nullptr, ref_offset, ref_lo, ref_hi, ref_prim);
                CK(cudaGetLastError());
                lj_free(prim_lo); lj_free(prim_hi);
                prim_lo = ref_lo; prim_hi = ref_hi;
            }
            lj_free(ref_count); lj_free(ref_offset); lj_free(stmp);
            if (ref_prim) return lj_bvh_build_refs(in, out, err, total, prim_lo, prim_hi, ref_prim, bounds);
        }
    }
    return lj_bvh_build_refs(in, out, err, n_prims_in, prim_lo, prim_hi, nullptr, bounds);
}

// Morton sort, PLOC / LBVH, 8-wide collapse over `n` references (boxes prim_lo/prim_hi; ref_prim maps a
// reference to its primitive, nullptr = identity).  Takes ownership of prim_lo, prim_hi, ref_prim, bounds.
static int lj_bvh_build_refs(const BvhBuildInput &in, BvhBuildOutput &out, char *err, int n, float4 *prim_lo, float4 *prim_hi,
                             int *ref_prim, float *bounds) {
    const int B = 256;
    const int G = (n + B - 1) / B;
    out.num_prims = n;
    BvhPhaseTimer pt;
    int *d_info = (int *)(bounds + 8);      // [0] root id, [1] PLOC status, [2..7] collapse counters
    float4 *node_lo = nullptr, *node_hi = nullptr;
    unsigned long long *keys = nullptr, *keys2 = nullptr;
    int *vals = nullptr, *vals2 = nullptr, *left = nullptr, *right = nullptr, *par_i = nullptr, *par_l = nullptr, *flags = nullptr, *count = nullptr;
    void *tmp = nullptr;
    auto free_temps = [&]() {
        lj_free(ref_prim);
        lj_free(prim_lo); lj_free(prim_hi); lj_free(bounds); lj_free(keys); lj_free(keys2); lj_free(vals); lj_free(vals2);
        lj_free(left); lj_free(right); lj_free(par_i); lj_free(par_l); lj_free(flags); lj_free(node_lo); lj_free(node_hi); lj_free(tmp); lj_free(count);
    };
    CK(lj_malloc(&keys, 8 * (size_t)n)); CK(lj_malloc(&keys2, 8 * (size_t)n));
    CK(lj_malloc(&vals, 4 * (size_t)n)); CK(lj_malloc(&vals2, 4 * (size_t)n));
    k_morton<<<G, B>>>(n, prim_lo, prim_hi, bounds, keys, vals);
    CK(cudaGetLastError());
    // LJ_SORT_BITS: sort on the top bits of the code only (fewer radix passes).  Measured and not adopted: 32 bits take
    // 0.2 ms off sponza's sort and cost 5 % of its traversal time (extend 74.4 -> 78.4 ms), profiles/r02_experiments.md.
    const char *bsel = getenv("LJ_BUILDER");       // "lbvh" keeps the plain radix tree; default PLOC
    const bool lbvh = bsel && bsel[0] == 'l';
    int sort_bits = 63;
    if (const char *e = getenv("LJ_SORT_BITS")) { int v = atoi(e); if (v >= 8 && v <= 63 && !lbvh) sort_bits = v; }
    size_t tmp_bytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, vals, vals2, n, 63 - sort_bits, 63));
    CK(lj_malloc(&tmp, tmp_bytes));
    pt.mark("morton codes");
    CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, vals, vals2, n, 63 - sort_bits, 63));
    pt.mark("radix sort");

    CK(lj_malloc(&out.tri_verts, sizeof(float4) * 3 * (size_t)n));
    k_emit_prims<<<G, B>>>(n, in.num_tris, vals2, ref_prim, in.positions, in.indices, in.tri_shape, in.shapes, in.spheres, out.tri_verts);
    CK(cudaGetLastError());
    pt.mark("emit primitives");

    if (n >= 2) {
        CK(lj_malloc(&left, 4 * (size_t)n)); CK(lj_malloc(&right, 4 * (size_t)n));
        CK(lj_malloc(&par_i, 4 * (size_t)n)); CK(lj_malloc(&par_l, 4 * (size_t)n));
        CK(lj_malloc(&flags, 4 * (size_t)n)); CK(cudaMemset(flags, 0, 4 * (size_t)n));
        CK(lj_malloc(&node_lo, sizeof(float4) * n)); CK(lj_malloc(&node_hi, sizeof(float4) * n));
        CK(lj_malloc(&count, 4 * (size_t)n));
        if (lbvh) {
            k_radix_tree<<<G, B>>>(n, keys2, left, right, par_i, par_l);
            CK(cudaGetLastError());
            k_fit<<<G, B>>>(n, vals2, prim_lo, prim_hi, left, right, par_i, par_l, flags, node_lo, node_hi);
            CK(cudaGetLastError());
            k_lbvh_counts<<<G, B>>>(n, left, right, count);
            CK(cudaMemsetAsync(d_info, 0, 2 * sizeof(int), 0));         // root id 0
        } else {
            if (ploc_build(n, vals2, prim_lo, prim_hi, left, right, count, node_lo, node_hi, d_info, err) != 0) { free_temps(); return -1; }
        }
        pt.mark("PLOC");
        if (pt.on) {
            // id-independent fingerprint of the binary tree (the builders number their nodes differently): a node hashes
            // its children's hashes in order, a leaf its sorted slot
            std::vector<int> hl(n), hr(n); int hroot = 0;
            CK(cudaMemcpy(hl.data(), left, 4 * (size_t)n, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(hr.data(), right, 4 * (size_t)n, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(&hroot, d_info, 4, cudaMemcpyDeviceToHost));
            std::vector<unsigned long long> hv(n, 0ull); std::vector<char> done(n, 0); std::vector<int> stk;
            auto leaf_h = [](int c) { unsigned long long x = (unsigned long long)(unsigned)~c * 0x9E3779B97F4A7C15ull; return x ^ (x >> 29); };
            stk.push_back(hroot);
            while (!stk.empty()) {
                const int x = stk.back();
                const int l = hl[x], r = hr[x];
                const bool lr = l < 0 || done[l], rr = r < 0 || done[r];
                if (lr && rr) {
                    const unsigned long long a = l < 0 ? leaf_h(l) : hv[l], b = r < 0 ? leaf_h(r) : hv[r];
                    unsigned long long h = (a * 0xBF58476D1CE4E5B9ull) ^ ((b + 0x94D049BB133111EBull) * 0xD6E8FEB86659FD93ull);
                    hv[x] = h ^ (h >> 31); done[x] = 1; stk.pop_back();
                } else { if (!lr) stk.push_back(l); if (!rr) stk.push_back(r); }
            }
            fprintf(stderr, "[bvh]   binary tree fingerprint %016llx\n", hv[hroot]);
            pt.mark("(fingerprint)");
        }
        CK(lj_malloc(&out.bvh2, sizeof(Bvh2Node) * (size_t)(n - 1)));
        k_emit_bvh2<<<G, B>>>(n, left, right, node_lo, node_hi, out.bvh2);
        CK(cudaGetLastError());
        out.num_bvh2 = n - 1;
    }
    pt.mark("emit binary tree");
    // the 8-wide collapse consumes the binary tree (lj_bvh8_build.cu)
    int rc = lj_bvh8_collapse(n, d_info, out.tri_verts, left, right, count, node_lo, node_hi, vals2, prim_lo, prim_hi, out, d_info + 2, err);
    pt.mark("8-wide collapse");
    // the one blocking read-back of the build: scene box, root id, PLOC status, collapse counters
    float h_res[16];
    if (rc == 0) {
        cudaError_t e_ = cudaMemcpy(h_res, bounds, sizeof h_res, cudaMemcpyDeviceToHost);
        if (e_ != cudaSuccess) { snprintf(err, 256, "BVH build: %s", cudaGetErrorString(e_)); rc = -1; }
    }
    free_temps();
    if (rc != 0) return rc;
    int h_info[8];
    memcpy(h_info, h_res + 8, sizeof h_info);
    for (int a = 0; a < 3; a++) { out.lo[a] = h_res[a]; out.hi[a] = h_res[3 + a]; }
    out.bvh2_root = h_info[0];
    if (h_info[1] != 0) { snprintf(err, 256, "PLOC made no progress (a round merged nothing)"); return -1; }
    out.num_bvh8 = h_info[2];
    out.num_slots8 = h_info[3];
    out.levels8 = h_info[7];
    if (pt.on) fprintf(stderr, "[bvh]   8-wide tree: %d nodes, %d levels\n", out.num_bvh8, out.levels8);
    if (out.levels8 > 128) { snprintf(err, 256, "bvh8 collapse did not terminate"); return -1; }
    if (out.num_slots8 != n) { snprintf(err, 256, "bvh8 collapse lost primitives: %d of %d", out.num_slots8, n); return -1; }
    return 0;
}

void lj_bvh_free(BvhBuildOutput &out) {
    lj_free(out.tri_verts); lj_free(out.bvh2); lj_free(out.bvh8); lj_free(out.tri_verts8);
    memset(&out, 0, sizeof out);
}
