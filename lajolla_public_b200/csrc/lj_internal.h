// Internal host-side structures behind the opaque C-ABI handles.
#pragma once
#include "lj_scene.cuh"
#include "lj_bvh_build.h"
#include <string>
#include <vector>

void lj_set_error(const char *fmt, ...);

struct LjDeviceGroup;
struct lj_ctx {
    int device;
    int sm_count;
    // wavefront path-slot arena: grown on demand, shared by every scene of the context (renders are
    // synchronous and the ABI is single-threaded per context), freed by lj_shutdown
    void *wf_pool; size_t wf_pool_bytes;
    // read-back ring of the wavefront engines (pinned host buffer + events), created once per context: creating
    // them per render cost more than a small render itself
    int *h_ring;                       // LJ_RING_DEPTH x LJ_RING_INTS ints, pinned
    cudaEvent_t ring_done[8];          // no-timing events, one per ring slot
    cudaEvent_t ring_ev[8][5];         // timing events per ring slot (stage times, lj_render_params.flags bit1)
    // second stream of the surface wavefront engine (shadow + finish of an iteration beside generate + extend of the next,
    // lj_wavefront.cu) and the events that join the two per ring slot
    cudaStream_t stream_b;
    cudaEvent_t ring_a[8], ring_b[8];
    bool ring_ready;
    // multi-device contexts (lj_init_devices, lj_multi.cu): the per-device contexts, worker threads and NCCL
    // communicators; null for a one-device context.  `this` is then only the handle and mirrors device 0.
    LjDeviceGroup *group;
};
int lj_ctx_arena(lj_ctx *ctx, size_t bytes, void **out);
enum { LJ_RING_DEPTH = 8, LJ_RING_INTS = 64 };
int lj_ctx_ring(lj_ctx *ctx);          // creates the ring on first use

struct lj_scene {
    lj_ctx *ctx;
    DScene d;                       // device pointers inside
    std::vector<void *> allocs;     // every cudaMalloc owned by the scene
    BvhBuildOutput bvh;
    lj_stats build_stats;
    bool kind_present[LJ_MAT_KIND_COUNT];   // material kinds the scene uses (one shade launch per kind)
    bool grid_media;                        // any heterogeneous medium (the tracking kernels run)
    bool media_grey;                        // every medium has equal channels everywhere (one-channel tracking, lj_volpath.cuh)
    std::vector<lj_scene *> replicas;       // multi-device contexts: one scene per device (this object is only the handle)
};

// one-device implementations behind the C ABI (lj_api.cu); the multi-device layer (lj_multi.cu) calls them per device
int lj_init_single(int device_id, lj_ctx **out);
void lj_shutdown_single(lj_ctx *ctx);
int lj_scene_create_single(lj_ctx *ctx, const lj_scene_desc *desc, uint32_t flags, lj_scene **out);
void lj_scene_destroy_single(lj_scene *s);
int lj_render_device_single(lj_scene *s, const lj_render_params *params, double *dev_sum, lj_stats *stats);
// multi-device layer
void lj_group_shutdown(lj_ctx *ctx);
int lj_group_scene_create(lj_ctx *ctx, const lj_scene_desc *desc, uint32_t flags, lj_scene **out);
void lj_group_scene_destroy(lj_scene *s);
int lj_group_render_device(lj_scene *s, const lj_render_params *params, double *dev_sum, lj_stats *stats);
const char *lj_error_text();

struct RenderLaunch {
    lj_render_params p;
    int spp, s_begin, s_end;
    int width, height, ntx, nty;
};

// engines (each fills dev_sum with per-pixel radiance sums over the rendered samples)
int lj_render_tiles_mode_a(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st);
int lj_render_megakernel(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st);
int lj_render_wavefront(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st);
int lj_render_aux(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st);
int lj_render_volpath(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st);
int lj_render_volwave(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st);
// scene.cpp:36-52 on the device (lj_scene_build.cu)
int lj_scene_build_tables(lj_scene *s, const lj_scene_desc *desc, bool poison, float *elapsed_ms);
// per-triangle shading records + f32 level-0 texels (DScene::tri_shade, texels0_*), lj_scene_build.cu
int lj_scene_build_shading_cache(lj_scene *s, const lj_scene_desc *desc, const int *d_tri_shape);

#define LJ_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { lj_set_error("%s failed: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); return LJ_ECUDA; } } while (0)
