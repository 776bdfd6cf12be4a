// TEST INFRASTRUCTURE (oracle) -- not part of the product path.
//
// A CPU stand-in for the 22 Embree 3.13.2 entry points the reference links
// against; the real libembree3.so.3 is a missing blob in
// /root/reference (.MISSING_LARGE_BLOBS).  Linked with the UNMODIFIED
// reference sources it yields the parity oracle, the golden-image generator
// and the timed CPU baseline ("reference code + BVH shim, not Embree").
//
// Call sites this file serves (all paths under /root/reference/src):
//   main.cpp:30,48            rtcNewDevice / rtcReleaseDevice
//   scene.cpp:20-34,64,70     scene create / quality / flags / commit / bounds / retain / release
//   shapes/triangle_mesh.inl:1-22   FLOAT3 vertex buffer (16-byte stride), UINT3 index buffer
//   shapes/sphere.inl:143-154       user geometry: bounds / intersect / occluded callbacks
//   intersection.cpp:32,83    rtcIntersect1 / rtcOccluded1
//
// Ray/primitive rules (documented in DESIGN.md "Ray-cast seam"; the CUDA
// traversal restates the same arithmetic independently):
//   * triangles: Pluecker edge test in f32 with explicit FMA placement
//     (tri_test below), accepted iff tnear < t <= tfar0;
//   * user primitives (spheres): the reference callback is invoked with the
//     ORIGINAL ray segment, so acceptance does not depend on visit order;
//   * closest hit = lexicographic minimum of (t, geomID, primID) over all
//     accepted primitives, so the answer is independent of the BVH shape;
//   * boxes only cull, with a padded slab test.
//
// Parity status: the single reference test crossing this seam is
// src/tests/intersection.cpp (one triangle, one ray) -- it passes against
// this shim (tests/test_oracle.py).  Beyond that the Embree seam is
// "parity unpinned" (real Embree arithmetic is not available here).

#include <embree3/rtcore.h>

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <vector>

namespace {

struct V3 { float x, y, z; };
static inline V3 sub3(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 add3(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline float fma1(float a, float b, float c) { return __builtin_fmaf(a, b, c); }
// cross / dot with fixed FMA placement (odd in each argument, so shared
// edges of neighbouring triangles get exactly negated edge functions).
static inline V3 cross_f(V3 a, V3 b) {
    return {fma1(a.y, b.z, -(a.z * b.y)),
            fma1(a.z, b.x, -(a.x * b.z)),
            fma1(a.x, b.y, -(a.y * b.x))};
}
static inline float dot_f(V3 a, V3 b) { return fma1(a.x, b.x, fma1(a.y, b.y, a.z * b.z)); }

struct TriHit { float t, u, v; V3 ng; };

// Pluecker edge test.  p = (1-u-v) p0 + u p1 + v p2 (triangle_mesh.inl:81-83).
static inline bool tri_test(V3 O, V3 D, float tnear, float tfar,
                            V3 p0, V3 p1, V3 p2, TriHit &h) {
    V3 v0 = sub3(p0, O), v1 = sub3(p1, O), v2 = sub3(p2, O);
    V3 e0 = sub3(v2, v0), e1 = sub3(v0, v1), e2 = sub3(v1, v2);
    float U = dot_f(cross_f(e0, add3(v2, v0)), D);
    float V = dot_f(cross_f(e1, add3(v0, v1)), D);
    float W = dot_f(cross_f(e2, add3(v1, v2)), D);
    float UVW = (U + V) + W;
    float eps = FLT_EPSILON * std::fabs(UVW);
    float mn = std::fmin(U, std::fmin(V, W));
    float mx = std::fmax(U, std::fmax(V, W));
    if (!(mn >= -eps || mx <= eps)) return false;
    if (UVW == 0.f) return false;
    V3 ng = cross_f(e0, e1); // == cross(p1-p0, p2-p0)
    float den = dot_f(ng, D);
    if (den == 0.f) return false;
    float t = dot_f(v0, ng) / den;
    if (!(t > tnear && t <= tfar)) return false;
    h.t = t; h.u = U / UVW; h.v = V / UVW; h.ng = ng;
    return true;
}

struct Prim {            // one BVH primitive
    float lo[3], hi[3];
    uint32_t geom, prim;
    int is_user;
};

struct Geom {
    std::atomic<int> refs{1};
    RTCGeometryType type;
    unsigned geomID = 0;
    // triangles
    std::vector<char> vbuf, ibuf;
    size_t vstride = 0, vcount = 0, istride = 0, icount = 0;
    // user
    unsigned user_count = 0;
    void *user_ptr = nullptr;
    RTCBoundsFunction bounds_fn = nullptr;
    void *bounds_user = nullptr;
    RTCIntersectFunctionN isect_fn = nullptr;
    RTCOccludedFunctionN occl_fn = nullptr;
};

struct Node {            // BVH2, 32 bytes
    float lo[3]; uint32_t left_or_first;   // inner: index of left child (right = left+1)
    float hi[3]; uint32_t count;           // 0 => inner
};

struct LeafTri { V3 p0, p1, p2; uint32_t geom, prim; };

struct SceneImpl {
    std::atomic<int> refs{1};
    std::vector<Geom *> geoms;
    std::vector<Node> nodes;
    std::vector<Prim> prims;      // in leaf order
    std::vector<LeafTri> tris;    // parallel to prims (valid when !is_user)
    float lo[3], hi[3];
    bool committed = false;
};

// ---- ray counters (per thread, summed on demand) ----
struct alignas(64) Counter { std::atomic<uint64_t> isect{0}, occl{0}; };
std::mutex g_counter_mu;
std::vector<Counter *> g_counters;
Counter *my_counter() {
    thread_local Counter *c = nullptr;
    if (!c) {
        c = new Counter();
        std::lock_guard<std::mutex> g(g_counter_mu);
        g_counters.push_back(c);
    }
    return c;
}

// ---- binned SAH builder ----
struct BuildCtx {
    std::vector<Prim> &prims;
    std::vector<Node> &nodes;
};

static inline float half_area(const float *lo, const float *hi) {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

void build_rec(BuildCtx &c, uint32_t node_idx, uint32_t first, uint32_t count) {
    Node &n0 = c.nodes[node_idx];
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = first; i < first + count; i++) {
        const Prim &p = c.prims[i];
        for (int a = 0; a < 3; a++) {
            lo[a] = std::min(lo[a], p.lo[a]); hi[a] = std::max(hi[a], p.hi[a]);
            float ctr = 0.5f * (p.lo[a] + p.hi[a]);
            clo[a] = std::min(clo[a], ctr); chi[a] = std::max(chi[a], ctr);
        }
    }
    for (int a = 0; a < 3; a++) { n0.lo[a] = lo[a]; n0.hi[a] = hi[a]; }
    auto make_leaf = [&]() { c.nodes[node_idx].left_or_first = first; c.nodes[node_idx].count = count; };
    if (count <= 2) { make_leaf(); return; }

    constexpr int NB = 16;
    float best_cost = FLT_MAX; int best_axis = -1, best_split = -1;
    for (int a = 0; a < 3; a++) {
        float ext = chi[a] - clo[a];
        if (!(ext > 0)) continue;
        struct Bin { float lo[3], hi[3]; int n; } bins[NB];
        for (auto &b : bins) { for (int k = 0; k < 3; k++) { b.lo[k] = FLT_MAX; b.hi[k] = -FLT_MAX; } b.n = 0; }
        float scale = NB / ext;
        for (uint32_t i = first; i < first + count; i++) {
            const Prim &p = c.prims[i];
            int b = std::min(NB - 1, std::max(0, int((0.5f * (p.lo[a] + p.hi[a]) - clo[a]) * scale)));
            bins[b].n++;
            for (int k = 0; k < 3; k++) { bins[b].lo[k] = std::min(bins[b].lo[k], p.lo[k]); bins[b].hi[k] = std::max(bins[b].hi[k], p.hi[k]); }
        }
        float la[NB], ra[NB]; int ln[NB], rn[NB];
        float l0[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, h0[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX}; int cnt = 0;
        for (int b = 0; b < NB - 1; b++) {
            cnt += bins[b].n;
            for (int k = 0; k < 3; k++) { l0[k] = std::min(l0[k], bins[b].lo[k]); h0[k] = std::max(h0[k], bins[b].hi[k]); }
            ln[b] = cnt; la[b] = cnt ? half_area(l0, h0) : 0.f;
        }
        float l1[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, h1[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX}; cnt = 0;
        for (int b = NB - 1; b > 0; b--) {
            cnt += bins[b].n;
            for (int k = 0; k < 3; k++) { l1[k] = std::min(l1[k], bins[b].lo[k]); h1[k] = std::max(h1[k], bins[b].hi[k]); }
            rn[b - 1] = cnt; ra[b - 1] = cnt ? half_area(l1, h1) : 0.f;
        }
        for (int b = 0; b < NB - 1; b++) {
            if (ln[b] == 0 || rn[b] == 0) continue;
            float cost = la[b] * ln[b] + ra[b] * rn[b];
            if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = b; }
        }
    }
    uint32_t mid;
    if (best_axis < 0) {
        if (count <= 8) { make_leaf(); return; }
        mid = first + count / 2; // all centroids coincide: median split
    } else {
        float leaf_cost = half_area(lo, hi) * count;
        if (count <= 4 && leaf_cost <= best_cost + 0.5f * half_area(lo, hi)) { make_leaf(); return; }
        int a = best_axis; float ext = chi[a] - clo[a]; float scale = NB / ext;
        auto it = std::partition(c.prims.begin() + first, c.prims.begin() + first + count, [&](const Prim &p) {
            int b = std::min(NB - 1, std::max(0, int((0.5f * (p.lo[a] + p.hi[a]) - clo[a]) * scale)));
            return b <= best_split;
        });
        mid = uint32_t(it - c.prims.begin());
        if (mid == first || mid == first + count) mid = first + count / 2;
    }
    uint32_t left = (uint32_t)c.nodes.size();
    c.nodes.emplace_back(); c.nodes.emplace_back();
    c.nodes[node_idx].left_or_first = left; c.nodes[node_idx].count = 0;
    build_rec(c, left, first, mid - first);
    build_rec(c, left + 1, mid, first + count - mid);
}

// Padded slab test: returns entry distance or +inf on miss.
static inline float slab(const float *lo, const float *hi, const float *o, const float *id,
                         float tnear, float tfar) {
    float t0 = tnear, t1 = tfar;
    for (int a = 0; a < 3; a++) {
        float ta = (lo[a] - o[a]) * id[a], tb = (hi[a] - o[a]) * id[a];
        float tn = std::fmin(ta, tb), tx = std::fmax(ta, tb); // fmin/fmax drop NaN (0*inf)
        t0 = std::fmax(t0, tn); t1 = std::fmin(t1, tx);
    }
    // pad both ends by a few ulps and a tiny absolute term
    float enter = t0 - std::fabs(t0) * 4e-7f;
    float exit_ = t1 + std::fabs(t1) * 4e-7f;
    return (enter <= exit_) ? enter : std::numeric_limits<float>::infinity();
}

struct Best {
    float t; uint32_t geom, prim; float u, v; V3 ng; bool valid;
};
static inline bool better(float t, uint32_t g, uint32_t p, const Best &b) {
    if (!b.valid) return true;
    if (t != b.t) return t < b.t;
    if (g != b.geom) return g < b.geom;
    return p < b.prim;
}

template <bool ANY>
bool traverse(SceneImpl *s, RTCIntersectContext *ctx, RTCRay &ray, RTCHit *hit_o) {
    if (s->nodes.empty()) return false;
    const float o[3] = {ray.org_x, ray.org_y, ray.org_z};
    const float d[3] = {ray.dir_x, ray.dir_y, ray.dir_z};
    const float id[3] = {1.0f / d[0], 1.0f / d[1], 1.0f / d[2]};
    const V3 O{o[0], o[1], o[2]}, D{d[0], d[1], d[2]};
    const float tnear = ray.tnear, tfar0 = ray.tfar;
    Best best; best.valid = false; best.t = tfar0;
    uint32_t stack[128]; int sp = 0;
    if (slab(s->nodes[0].lo, s->nodes[0].hi, o, id, tnear, tfar0) == std::numeric_limits<float>::infinity()) return false;
    stack[sp++] = 0;
    while (sp) {
        const Node &n = s->nodes[stack[--sp]];
        if (n.count == 0) {
            const Node &l = s->nodes[n.left_or_first], &r = s->nodes[n.left_or_first + 1];
            float tl = slab(l.lo, l.hi, o, id, tnear, best.t);
            float tr = slab(r.lo, r.hi, o, id, tnear, best.t);
            bool hl = tl != std::numeric_limits<float>::infinity(), hr = tr != std::numeric_limits<float>::infinity();
            if (hl && hr) {
                if (tl <= tr) { stack[sp++] = n.left_or_first + 1; stack[sp++] = n.left_or_first; }
                else { stack[sp++] = n.left_or_first; stack[sp++] = n.left_or_first + 1; }
            } else if (hl) stack[sp++] = n.left_or_first;
            else if (hr) stack[sp++] = n.left_or_first + 1;
            continue;
        }
        for (uint32_t i = n.left_or_first; i < n.left_or_first + n.count; i++) {
            const Prim &p = s->prims[i];
            if (!p.is_user) {
                const LeafTri &tr = s->tris[i];
                TriHit h;
                if (tri_test(O, D, tnear, tfar0, tr.p0, tr.p1, tr.p2, h)) {
                    if (ANY) return true;
                    if (better(h.t, tr.geom, tr.prim, best)) {
                        best.valid = true; best.t = h.t; best.geom = tr.geom; best.prim = tr.prim;
                        best.u = h.u; best.v = h.v; best.ng = h.ng;
                    }
                }
            } else {
                Geom *g = s->geoms[p.geom];
                int valid = -1;
                if (ANY) {
                    RTCRay r2 = ray; r2.tfar = tfar0;
                    RTCOccludedFunctionNArguments a;
                    a.valid = &valid; a.geometryUserPtr = g->user_ptr; a.primID = p.prim; a.context = ctx;
                    a.ray = (RTCRayN *)&r2; a.N = 1; a.geomID = p.geom;
                    if (g->occl_fn) g->occl_fn(&a);
                    if (r2.tfar < 0) return true;
                } else {
                    RTCRayHit rh; rh.ray = ray; rh.ray.tfar = tfar0;
                    rh.hit.geomID = RTC_INVALID_GEOMETRY_ID; rh.hit.primID = RTC_INVALID_GEOMETRY_ID;
                    rh.hit.instID[0] = RTC_INVALID_GEOMETRY_ID; rh.hit.u = rh.hit.v = 0; rh.hit.Ng_x = rh.hit.Ng_y = rh.hit.Ng_z = 0;
                    RTCIntersectFunctionNArguments a;
                    a.valid = &valid; a.geometryUserPtr = g->user_ptr; a.primID = p.prim; a.context = ctx;
                    a.rayhit = (RTCRayHitN *)&rh; a.N = 1; a.geomID = p.geom;
                    if (g->isect_fn) g->isect_fn(&a);
                    if (rh.hit.geomID != RTC_INVALID_GEOMETRY_ID) {
                        float t = rh.ray.tfar;
                        if (better(t, p.geom, p.prim, best)) {
                            best.valid = true; best.t = t; best.geom = p.geom; best.prim = p.prim;
                            best.u = rh.hit.u; best.v = rh.hit.v; best.ng = {rh.hit.Ng_x, rh.hit.Ng_y, rh.hit.Ng_z};
                        }
                    }
                }
            }
        }
    }
    if (ANY) return false;
    if (!best.valid) return false;
    ray.tfar = best.t;
    hit_o->Ng_x = best.ng.x; hit_o->Ng_y = best.ng.y; hit_o->Ng_z = best.ng.z;
    hit_o->u = best.u; hit_o->v = best.v; hit_o->primID = best.prim; hit_o->geomID = best.geom;
    hit_o->instID[0] = RTC_INVALID_GEOMETRY_ID;
    return true;
}

void release_geom(Geom *g) { if (--g->refs == 0) delete g; }

} // namespace

// Exposed for the oracle driver (ray counting for Mrays/s).
extern "C" void lj_shim_ray_counts(uint64_t *isect, uint64_t *occl) {
    std::lock_guard<std::mutex> g(g_counter_mu);
    uint64_t a = 0, b = 0;
    for (Counter *c : g_counters) { a += c->isect.load(std::memory_order_relaxed); b += c->occl.load(std::memory_order_relaxed); }
    *isect = a; *occl = b;
}
extern "C" void lj_shim_reset_ray_counts() {
    std::lock_guard<std::mutex> g(g_counter_mu);
    for (Counter *c : g_counters) { c->isect = 0; c->occl = 0; }
}
// Brute-force closest hit over every primitive (validates the BVH; tests only).
extern "C" int lj_shim_intersect_bruteforce(RTCScene scene, RTCRayHit *rh) {
    SceneImpl *s = (SceneImpl *)scene;
    // temporarily view all prims as one leaf
    std::vector<Node> saved; saved.swap(s->nodes);
    Node root; for (int a = 0; a < 3; a++) { root.lo[a] = -FLT_MAX; root.hi[a] = FLT_MAX; }
    root.left_or_first = 0; root.count = (uint32_t)s->prims.size();
    s->nodes.push_back(root);
    RTCIntersectContext ctx; rtcInitIntersectContext(&ctx);
    bool h = traverse<false>(s, &ctx, rh->ray, &rh->hit);
    s->nodes.swap(saved);
    return h ? 1 : 0;
}

RTC_API RTCDevice rtcNewDevice(const char *) { return (RTCDevice) new int(1); }
RTC_API void rtcReleaseDevice(RTCDevice d) { delete (int *)d; }

RTC_API RTCScene rtcNewScene(RTCDevice) { return (RTCScene) new SceneImpl(); }
RTC_API void rtcRetainScene(RTCScene s) { ((SceneImpl *)s)->refs++; }
RTC_API void rtcReleaseScene(RTCScene s_) {
    SceneImpl *s = (SceneImpl *)s_;
    if (!s) return;
    if (--s->refs == 0) { for (Geom *g : s->geoms) release_geom(g); delete s; }
}
RTC_API void rtcSetSceneBuildQuality(RTCScene, enum RTCBuildQuality) {}
RTC_API void rtcSetSceneFlags(RTCScene, enum RTCSceneFlags) {}

RTC_API RTCGeometry rtcNewGeometry(RTCDevice, enum RTCGeometryType type) {
    Geom *g = new Geom(); g->type = type; return (RTCGeometry)g;
}
RTC_API unsigned int rtcAttachGeometry(RTCScene s_, RTCGeometry g_) {
    SceneImpl *s = (SceneImpl *)s_; Geom *g = (Geom *)g_;
    g->refs++; g->geomID = (unsigned)s->geoms.size(); s->geoms.push_back(g);
    return g->geomID;
}
RTC_API void rtcCommitGeometry(RTCGeometry) {}
RTC_API void rtcReleaseGeometry(RTCGeometry g) { release_geom((Geom *)g); }
RTC_API void rtcSetGeometryVertexAttributeCount(RTCGeometry, unsigned int) {}
RTC_API void *rtcSetNewGeometryBuffer(RTCGeometry g_, enum RTCBufferType type, unsigned int,
                                      enum RTCFormat, size_t byteStride, size_t itemCount) {
    Geom *g = (Geom *)g_;
    if (type == RTC_BUFFER_TYPE_VERTEX) {
        g->vstride = byteStride; g->vcount = itemCount; g->vbuf.assign(byteStride * itemCount + 16, 0);
        return g->vbuf.data();
    } else if (type == RTC_BUFFER_TYPE_INDEX) {
        g->istride = byteStride; g->icount = itemCount; g->ibuf.assign(byteStride * itemCount + 16, 0);
        return g->ibuf.data();
    }
    return nullptr;
}
RTC_API void rtcSetGeometryUserPrimitiveCount(RTCGeometry g, unsigned int n) { ((Geom *)g)->user_count = n; }
RTC_API void rtcSetGeometryUserData(RTCGeometry g, void *p) { ((Geom *)g)->user_ptr = p; }
RTC_API void rtcSetGeometryBoundsFunction(RTCGeometry g, RTCBoundsFunction f, void *u) { ((Geom *)g)->bounds_fn = f; ((Geom *)g)->bounds_user = u; }
RTC_API void rtcSetGeometryIntersectFunction(RTCGeometry g, RTCIntersectFunctionN f) { ((Geom *)g)->isect_fn = f; }
RTC_API void rtcSetGeometryOccludedFunction(RTCGeometry g, RTCOccludedFunctionN f) { ((Geom *)g)->occl_fn = f; }

RTC_API void rtcCommitScene(RTCScene s_) {
    SceneImpl *s = (SceneImpl *)s_;
    s->prims.clear(); s->tris.clear(); s->nodes.clear();
    std::vector<LeafTri> tri_src; // indexed by (geom, prim) lookup after partitioning
    for (Geom *g : s->geoms) {
        if (g->type == RTC_GEOMETRY_TYPE_TRIANGLE) {
            for (size_t i = 0; i < g->icount; i++) {
                const uint32_t *idx = (const uint32_t *)(g->ibuf.data() + i * g->istride);
                Prim p; p.geom = g->geomID; p.prim = (uint32_t)i; p.is_user = 0;
                for (int a = 0; a < 3; a++) { p.lo[a] = FLT_MAX; p.hi[a] = -FLT_MAX; }
                for (int k = 0; k < 3; k++) {
                    const float *v = (const float *)(g->vbuf.data() + (size_t)idx[k] * g->vstride);
                    for (int a = 0; a < 3; a++) { p.lo[a] = std::min(p.lo[a], v[a]); p.hi[a] = std::max(p.hi[a], v[a]); }
                }
                s->prims.push_back(p);
            }
        } else if (g->type == RTC_GEOMETRY_TYPE_USER) {
            for (unsigned i = 0; i < g->user_count; i++) {
                RTCBounds b; RTCBoundsFunctionArguments a;
                a.geometryUserPtr = g->user_ptr; a.primID = i; a.timeStep = 0; a.bounds_o = &b;
                g->bounds_fn(&a);
                Prim p; p.geom = g->geomID; p.prim = i; p.is_user = 1;
                p.lo[0] = b.lower_x; p.lo[1] = b.lower_y; p.lo[2] = b.lower_z;
                p.hi[0] = b.upper_x; p.hi[1] = b.upper_y; p.hi[2] = b.upper_z;
                s->prims.push_back(p);
            }
        }
    }
    for (int a = 0; a < 3; a++) { s->lo[a] = FLT_MAX; s->hi[a] = -FLT_MAX; }
    for (const Prim &p : s->prims)
        for (int a = 0; a < 3; a++) { s->lo[a] = std::min(s->lo[a], p.lo[a]); s->hi[a] = std::max(s->hi[a], p.hi[a]); }
    if (!s->prims.empty()) {
        s->nodes.reserve(s->prims.size() * 2);
        s->nodes.emplace_back();
        BuildCtx c{s->prims, s->nodes};
        build_rec(c, 0, 0, (uint32_t)s->prims.size());
    }
    s->tris.resize(s->prims.size());
    for (size_t i = 0; i < s->prims.size(); i++) {
        const Prim &p = s->prims[i];
        if (p.is_user) continue;
        Geom *g = s->geoms[p.geom];
        const uint32_t *idx = (const uint32_t *)(g->ibuf.data() + (size_t)p.prim * g->istride);
        const float *a = (const float *)(g->vbuf.data() + (size_t)idx[0] * g->vstride);
        const float *b = (const float *)(g->vbuf.data() + (size_t)idx[1] * g->vstride);
        const float *c = (const float *)(g->vbuf.data() + (size_t)idx[2] * g->vstride);
        s->tris[i] = LeafTri{{a[0], a[1], a[2]}, {b[0], b[1], b[2]}, {c[0], c[1], c[2]}, p.geom, p.prim};
    }
    s->committed = true;
}

RTC_API void rtcGetSceneBounds(RTCScene s_, struct RTCBounds *b) {
    SceneImpl *s = (SceneImpl *)s_;
    b->lower_x = s->lo[0]; b->lower_y = s->lo[1]; b->lower_z = s->lo[2]; b->align0 = 0;
    b->upper_x = s->hi[0]; b->upper_y = s->hi[1]; b->upper_z = s->hi[2]; b->align1 = 0;
}

// What the calling thread's last rtcIntersect1 returned (the seam dump of ref_driver reads the raw f32
// tfar / u / v / Ng that the reference's PathVertex no longer carries).
static thread_local RTCRayHit g_last_rayhit;
extern "C" void lj_shim_last_rayhit(RTCRayHit *out) { *out = g_last_rayhit; }

RTC_API void rtcIntersect1(RTCScene s, struct RTCIntersectContext *ctx, struct RTCRayHit *rh) {
    my_counter()->isect.fetch_add(1, std::memory_order_relaxed);
    traverse<false>((SceneImpl *)s, ctx, rh->ray, &rh->hit);
    g_last_rayhit = *rh;
}
RTC_API void rtcOccluded1(RTCScene s, struct RTCIntersectContext *ctx, struct RTCRay *ray) {
    my_counter()->occl.fetch_add(1, std::memory_order_relaxed);
    RTCHit dummy;
    if (traverse<true>((SceneImpl *)s, ctx, *ray, &dummy)) ray->tfar = -std::numeric_limits<float>::infinity();
}
