// Host-side stand-in for the Embree entry points the reference's Scene
// constructor and loaders call (src/scene.cpp:20-34, src/shapes/*.inl,
// src/main.cpp:30,48) in the GPU drop-in build.  It only records what the
// reference hands over and answers rtcGetSceneBounds (f32 AABB of all
// primitives, which defines Scene::bounds and hence every epsilon,
// src/scene.h:99-105).  It builds no acceleration structure and cannot trace:
// ray casting lives on the GPU (csrc/lj_bvh_build.cu, csrc/lj_traverse.cuh).
// rtcIntersect1 / rtcOccluded1 abort loudly -- there is no CPU fallback.
#include <embree3/rtcore.h>

#include <algorithm>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace {
struct CapGeom {
    int refs = 1;
    RTCGeometryType type;
    std::vector<char> vbuf, ibuf;
    size_t vstride = 0, istride = 0, icount = 0;
    unsigned user_count = 0;
    void *user_ptr = nullptr;
    RTCBoundsFunction bounds_fn = nullptr;
};
struct CapScene {
    int refs = 1;
    std::vector<CapGeom *> geoms;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
};
void drop(CapGeom *g) { if (--g->refs == 0) delete g; }
[[noreturn]] void no_cpu_path(const char *what) {
    fprintf(stderr, "lajolla_b200: %s called on the host -- the GPU build has no CPU ray casting path\n", what);
    abort();
}
} // namespace

RTC_API RTCDevice rtcNewDevice(const char *) { return (RTCDevice) new int(0); }
RTC_API void rtcReleaseDevice(RTCDevice d) { delete (int *)d; }
RTC_API RTCScene rtcNewScene(RTCDevice) { return (RTCScene) new CapScene(); }
RTC_API void rtcRetainScene(RTCScene s) { ((CapScene *)s)->refs++; }
RTC_API void rtcReleaseScene(RTCScene s_) {
    CapScene *s = (CapScene *)s_;
    if (s && --s->refs == 0) { for (CapGeom *g : s->geoms) drop(g); delete s; }
}
RTC_API void rtcSetSceneBuildQuality(RTCScene, enum RTCBuildQuality) {}
RTC_API void rtcSetSceneFlags(RTCScene, enum RTCSceneFlags) {}
RTC_API RTCGeometry rtcNewGeometry(RTCDevice, enum RTCGeometryType type) {
    CapGeom *g = new CapGeom(); g->type = type; return (RTCGeometry)g;
}
RTC_API unsigned int rtcAttachGeometry(RTCScene s_, RTCGeometry g_) {
    CapScene *s = (CapScene *)s_; CapGeom *g = (CapGeom *)g_;
    g->refs++; s->geoms.push_back(g);
    return (unsigned)s->geoms.size() - 1;
}
RTC_API void rtcCommitGeometry(RTCGeometry) {}
RTC_API void rtcReleaseGeometry(RTCGeometry g) { drop((CapGeom *)g); }
RTC_API void rtcSetGeometryVertexAttributeCount(RTCGeometry, unsigned int) {}
RTC_API void *rtcSetNewGeometryBuffer(RTCGeometry g_, enum RTCBufferType type, unsigned int, enum RTCFormat,
                                      size_t byteStride, size_t itemCount) {
    CapGeom *g = (CapGeom *)g_;
    if (type == RTC_BUFFER_TYPE_VERTEX) { g->vstride = byteStride; g->vbuf.assign(byteStride * itemCount + 16, 0); return g->vbuf.data(); }
    if (type == RTC_BUFFER_TYPE_INDEX) { g->istride = byteStride; g->icount = itemCount; g->ibuf.assign(byteStride * itemCount + 16, 0); return g->ibuf.data(); }
    return nullptr;
}
RTC_API void rtcSetGeometryUserPrimitiveCount(RTCGeometry g, unsigned int n) { ((CapGeom *)g)->user_count = n; }
RTC_API void rtcSetGeometryUserData(RTCGeometry g, void *p) { ((CapGeom *)g)->user_ptr = p; }
RTC_API void rtcSetGeometryBoundsFunction(RTCGeometry g, RTCBoundsFunction f, void *) { ((CapGeom *)g)->bounds_fn = f; }
RTC_API void rtcSetGeometryIntersectFunction(RTCGeometry, RTCIntersectFunctionN) {}
RTC_API void rtcSetGeometryOccludedFunction(RTCGeometry, RTCOccludedFunctionN) {}

RTC_API void rtcCommitScene(RTCScene s_) {
    CapScene *s = (CapScene *)s_;
    for (int a = 0; a < 3; a++) { s->lo[a] = FLT_MAX; s->hi[a] = -FLT_MAX; }
    for (CapGeom *g : s->geoms) {
        if (g->type == RTC_GEOMETRY_TYPE_TRIANGLE) {
            for (size_t i = 0; i < g->icount; i++) {
                const unsigned *idx = (const unsigned *)(g->ibuf.data() + i * g->istride);
                for (int k = 0; k < 3; k++) {
                    const float *v = (const float *)(g->vbuf.data() + (size_t)idx[k] * g->vstride);
                    for (int a = 0; a < 3; a++) { s->lo[a] = std::min(s->lo[a], v[a]); s->hi[a] = std::max(s->hi[a], v[a]); }
                }
            }
        } else if (g->type == RTC_GEOMETRY_TYPE_USER && g->bounds_fn) {
            for (unsigned i = 0; i < g->user_count; i++) {
                RTCBounds b; RTCBoundsFunctionArguments a;
                a.geometryUserPtr = g->user_ptr; a.primID = i; a.timeStep = 0; a.bounds_o = &b;
                g->bounds_fn(&a);
                const float l[3] = {b.lower_x, b.lower_y, b.lower_z}, u[3] = {b.upper_x, b.upper_y, b.upper_z};
                for (int k = 0; k < 3; k++) { s->lo[k] = std::min(s->lo[k], l[k]); s->hi[k] = std::max(s->hi[k], u[k]); }
            }
        }
    }
}
RTC_API void rtcGetSceneBounds(RTCScene s_, struct RTCBounds *b) {
    CapScene *s = (CapScene *)s_;
    b->lower_x = s->lo[0]; b->lower_y = s->lo[1]; b->lower_z = s->lo[2]; b->align0 = 0;
    b->upper_x = s->hi[0]; b->upper_y = s->hi[1]; b->upper_z = s->hi[2]; b->align1 = 0;
}
RTC_API void rtcIntersect1(RTCScene, struct RTCIntersectContext *, struct RTCRayHit *) { no_cpu_path("rtcIntersect1"); }
RTC_API void rtcOccluded1(RTCScene, struct RTCIntersectContext *, struct RTCRay *) { no_cpu_path("rtcOccluded1"); }
