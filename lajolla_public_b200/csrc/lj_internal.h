// Internal host-side structures behind the opaque C-ABI handles.
#pragma once
#include "lj_scene.cuh"
#include "lj_bvh_build.h"
#include <string>
#include <vector>

void lj_set_error(const char *fmt, ...);

struct lj_ctx {
    int device;
    int sm_count;
    // wavefront path-slot arena: grown on demand, shared by every scene of the context (renders are
    // synchronous and the ABI is single-threaded per context), freed by lj_shutdown
    void *wf_pool; size_t wf_pool_bytes;
};
int lj_ctx_arena(lj_ctx *ctx, size_t bytes, void **out);

struct lj_scene {
    lj_ctx *ctx;
    DScene d;                       // device pointers inside
    std::vector<void *> allocs;     // every cudaMalloc owned by the scene
    BvhBuildOutput bvh;
    lj_stats build_stats;
};

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

#define LJ_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { lj_set_error("%s failed: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); return LJ_ECUDA; } } while (0)
