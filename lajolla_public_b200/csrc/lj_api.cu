// C ABI of include/lajolla_gpu.h: context, scene upload + BVH build, render
// dispatch, batch ray casting, scene blobs.
#include "lj_internal.h"
#include "lj_alloc.h"
#include "lj_traverse8.cuh"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

static thread_local char g_err[1024] = "";

void lj_set_error(const char *fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}
extern "C" const char *lj_last_error(void) { return g_err; }
const char *lj_error_text() { return g_err; }

extern "C" int lj_init(int device_id, lj_ctx **out) { return lj_init_single(device_id, out); }
extern "C" void lj_shutdown(lj_ctx *ctx) {
    if (!ctx) return;
    if (ctx->group) lj_group_shutdown(ctx); else lj_shutdown_single(ctx);
}

int lj_init_single(int device_id, lj_ctx **out) {
    if (!out) { lj_set_error("lj_init: out is null"); return LJ_EINVAL; }
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        lj_set_error("lj_init: no CUDA device (%s); this library has no CPU path", e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return LJ_ENODEVICE;
    }
    if (device_id < 0 || device_id >= n) { lj_set_error("lj_init: device %d out of range [0,%d)", device_id, n); return LJ_EINVAL; }
    LJ_CUDA(cudaSetDevice(device_id));
    lj_ctx *c = new lj_ctx();
    c->device = device_id; c->wf_pool = nullptr; c->wf_pool_bytes = 0; c->group = nullptr;
    c->h_ring = nullptr; c->ring_ready = false;
    {   // keep freed blocks in the device's default memory pool (see lj_alloc.h)
        cudaMemPool_t pool;
        LJ_CUDA(cudaDeviceGetDefaultMemPool(&pool, device_id));
        unsigned long long keep = ~0ULL;
        LJ_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    LJ_CUDA(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device_id));
    *out = c;
    return LJ_OK;
}
void lj_shutdown_single(lj_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->wf_pool) lj_free(ctx->wf_pool);
    cudaDeviceSynchronize();
    if (ctx->ring_ready) {
        for (int i = 0; i < LJ_RING_DEPTH; i++) {
            cudaEventDestroy(ctx->ring_done[i]); for (auto &e : ctx->ring_ev[i]) cudaEventDestroy(e);
            cudaEventDestroy(ctx->ring_a[i]); cudaEventDestroy(ctx->ring_b[i]);
        }
        cudaStreamDestroy(ctx->stream_b);
        cudaFreeHost(ctx->h_ring);
    }
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, ctx->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);   // hand the cached blocks back
    delete ctx;
}

int lj_ctx_ring(lj_ctx *ctx) {
    if (ctx->ring_ready) return LJ_OK;
    LJ_CUDA(cudaMallocHost(&ctx->h_ring, sizeof(int) * LJ_RING_INTS * LJ_RING_DEPTH));
    for (int i = 0; i < LJ_RING_DEPTH; i++) {
        LJ_CUDA(cudaEventCreateWithFlags(&ctx->ring_done[i], cudaEventDisableTiming));
        for (auto &e : ctx->ring_ev[i]) LJ_CUDA(cudaEventCreate(&e));
        LJ_CUDA(cudaEventCreateWithFlags(&ctx->ring_a[i], cudaEventDisableTiming));
        LJ_CUDA(cudaEventCreateWithFlags(&ctx->ring_b[i], cudaEventDisableTiming));
    }
    LJ_CUDA(cudaStreamCreateWithFlags(&ctx->stream_b, cudaStreamNonBlocking));
    ctx->ring_ready = true;
    return LJ_OK;
}

int lj_ctx_arena(lj_ctx *ctx, size_t bytes, void **out) {
    if (bytes > ctx->wf_pool_bytes) {
        if (ctx->wf_pool) lj_free(ctx->wf_pool);
        ctx->wf_pool = nullptr; ctx->wf_pool_bytes = 0;
        LJ_CUDA(lj_malloc(&ctx->wf_pool, bytes));
        ctx->wf_pool_bytes = bytes;
    }
    *out = ctx->wf_pool;
    return LJ_OK;
}

// All scene arrays go into ONE device allocation (sub-allocated at 256-byte boundaries) filled with
// asynchronous copies: a scene is ~20 arrays, and a cudaMalloc + blocking copy per array cost more than
// the transfer itself (sponza: 73 MB).
// Arrays marked `late` (the texels: 66 of sponza's 73 MB) are not needed by the BVH build; they are copied on the context's
// second stream so that the build kernels run while they are still in flight (from pinned host memory; a pageable source
// is staged by the driver and simply takes its usual time).  `late_done` is recorded behind them.
struct Uploader {
    struct Item { const void *host; size_t bytes, offset; const void **dev; bool late; };
    std::vector<Item> items;
    size_t total = 0;
    template <typename T>
    void add(const T *host, size_t count, const T **dev, bool late = false) {
        Item it; it.host = host; it.bytes = sizeof(T) * count; it.offset = total; it.dev = (const void **)dev; it.late = late;
        total += (sizeof(T) * (count ? count : 1) + 255) & ~(size_t)255;
        items.push_back(it);
    }
    int commit(lj_scene *s, cudaStream_t late_stream, cudaEvent_t late_done) {
        void *base = nullptr;
        LJ_CUDA(lj_malloc(&base, total ? total : 256));
        s->allocs.push_back(base);
        LJ_CUDA(cudaEventRecord(late_done));                       // the allocation is ordered on this thread's stream
        LJ_CUDA(cudaStreamWaitEvent(late_stream, late_done, 0));
        for (const Item &it : items) {
            *it.dev = (const char *)base + it.offset;
            if (it.bytes && !it.late) LJ_CUDA(cudaMemcpyAsync((char *)base + it.offset, it.host, it.bytes, cudaMemcpyHostToDevice));
        }
        for (const Item &it : items)
            if (it.bytes && it.late) LJ_CUDA(cudaMemcpyAsync((char *)base + it.offset, it.host, it.bytes, cudaMemcpyHostToDevice, late_stream));
        LJ_CUDA(cudaEventRecord(late_done, late_stream));
        return LJ_OK;
    }
};

// Every id, kind and offset of a description is range-checked here, once, on the host: the kernels index with them
// unchecked, and descriptions also come from .ljs files (lj_scene_desc_load).
static int validate_desc(const lj_scene_desc *d) {
#define BAD(...) do { lj_set_error(__VA_ARGS__); return LJ_EINVAL; } while (0)
    typedef unsigned long long ull;
    if (d->camera.width <= 0 || d->camera.height <= 0) BAD("scene: bad resolution %dx%d", d->camera.width, d->camera.height);
    if ((long long)d->camera.width * d->camera.height > (1LL << 28)) BAD("scene: film of %dx%d pixels is too large", d->camera.width, d->camera.height);
    if (d->camera.filter_kind < LJ_FILTER_BOX || d->camera.filter_kind > LJ_FILTER_GAUSSIAN) BAD("scene: unknown pixel filter %d", d->camera.filter_kind);
    if (d->camera.medium_id < -1 || d->camera.medium_id >= (long long)d->num_media) BAD("scene: camera medium id out of range");
    if (d->options.integrator < LJ_INTEGRATOR_DEPTH || d->options.integrator > LJ_INTEGRATOR_VOLPATH) BAD("scene: unknown integrator %d", d->options.integrator);
    if (d->envmap_light_id < -1 || d->envmap_light_id >= (long long)d->num_lights) BAD("scene: envmap light id out of range");
    if (d->num_shapes > 0x7fffffffu || d->num_lights > 0x7fffffffu || d->num_materials > 0x7fffffffu || d->num_media > 0x7fffffffu ||
        d->num_textures > 0x7fffffffu || d->num_images1 > 0x7fffffffu || d->num_images3 > 0x7fffffffu || d->num_indices > 0x7fffffffu / 3 ||
        d->num_positions > 0x7fffffffu) BAD("scene: record count out of range");
    auto in_tables = [&](uint64_t off, uint64_t n) { return off <= d->num_tables && n <= d->num_tables - off; };
    for (uint64_t i = 0; i < d->num_images1 + d->num_images3; i++) {
        const bool one = i < d->num_images1;
        const lj_image &im = one ? d->images1[i] : d->images3[i - d->num_images1];
        if (im.channels != (one ? 1 : 3) || im.num_levels < 1 || im.num_levels > 8) BAD("scene: image %llu has bad channels / level count", (ull)i);
        for (int l = 0; l < im.num_levels; l++) {
            if (im.width[l] <= 0 || im.height[l] <= 0) BAD("scene: image %llu level %d is empty", (ull)i, l);
            const uint64_t n = (uint64_t)im.width[l] * (uint64_t)im.height[l] * (uint64_t)im.channels;
            if (im.offset[l] > d->num_texels || n > d->num_texels - im.offset[l]) BAD("scene: image %llu level %d lies outside the texel array", (ull)i, l);
        }
    }
    for (uint64_t i = 0; i < d->num_textures; i++) {
        const lj_texture &t = d->textures[i];
        if (t.kind < LJ_TEX_CONSTANT || t.kind > LJ_TEX_CHECKERBOARD || (t.channels != 1 && t.channels != 3)) BAD("scene: texture %llu has an unknown kind / channel count", (ull)i);
        if (t.kind == LJ_TEX_IMAGE && (t.image_id < 0 || (uint64_t)t.image_id >= (t.channels == 1 ? d->num_images1 : d->num_images3)))
            BAD("scene: texture %llu image id out of range", (ull)i);
    }
    for (uint64_t i = 0; i < d->num_materials; i++) {
        const lj_material &m = d->materials[i];
        if (m.kind < 0 || m.kind >= LJ_MAT_KIND_COUNT) BAD("scene: material %llu has unknown kind %d", (ull)i, m.kind);
        for (int k = 0; k < LJ_SLOT_COUNT; k++)
            if (m.tex[k] < -1 || m.tex[k] >= (long long)d->num_textures) BAD("scene: material %llu texture slot %d out of range", (ull)i, k);
    }
    for (uint64_t i = 0; i < d->num_media; i++) {
        const lj_medium &m = d->media[i];
        if (m.kind < LJ_MEDIUM_HOMOGENEOUS || m.kind > LJ_MEDIUM_HETEROGENEOUS || m.phase_kind < LJ_PHASE_ISOTROPIC || m.phase_kind > LJ_PHASE_HG)
            BAD("scene: medium %llu has an unknown kind / phase function", (ull)i);
        const lj_volume *vols[2] = {&m.albedo, &m.density};
        for (const lj_volume *v : vols) {
            if (m.kind != LJ_MEDIUM_HETEROGENEOUS) continue;
            if (v->kind < LJ_VOLUME_CONSTANT || v->kind > LJ_VOLUME_GRID) BAD("scene: medium %llu has an unknown volume kind", (ull)i);
            if (v->kind != LJ_VOLUME_GRID) continue;
            if (v->resolution[0] <= 0 || v->resolution[1] <= 0 || v->resolution[2] <= 0) BAD("scene: medium %llu has an empty grid", (ull)i);
            const uint64_t n = 3ull * (uint64_t)v->resolution[0] * (uint64_t)v->resolution[1] * (uint64_t)v->resolution[2];
            if (v->resolution[0] > 4096 || v->resolution[1] > 4096 || v->resolution[2] > 4096 || v->data_offset > d->num_grid_data || n > d->num_grid_data - v->data_offset)
                BAD("scene: medium %llu grid lies outside the grid array", (ull)i);
        }
    }
    for (uint64_t i = 0; i < d->num_shapes; i++) {
        const lj_shape &s = d->shapes[i];
        if (s.kind != LJ_SHAPE_SPHERE && s.kind != LJ_SHAPE_MESH) BAD("scene: shape %llu has unknown kind %d", (ull)i, s.kind);
        if (s.material_id < -1 || s.material_id >= (long long)d->num_materials) BAD("scene: shape %llu material id out of range", (ull)i);
        if (s.area_light_id < -1 || s.area_light_id >= (long long)d->num_lights) BAD("scene: shape %llu area light id out of range", (ull)i);
        if (s.interior_medium_id < -1 || s.interior_medium_id >= (long long)d->num_media || s.exterior_medium_id < -1 || s.exterior_medium_id >= (long long)d->num_media)
            BAD("scene: shape %llu medium id out of range", (ull)i);
        if (s.material_id < 0 && d->options.integrator == LJ_INTEGRATOR_PATH) BAD("scene: shape %llu has no material (index-matched surfaces need the volpath integrator)", (ull)i);
        if (s.kind == LJ_SHAPE_MESH) {
            if ((uint64_t)s.vertex_offset + s.num_vertices > d->num_positions || (uint64_t)s.triangle_offset + s.num_triangles > d->num_indices)
                BAD("scene: shape %llu mesh ranges out of bounds", (ull)i);
            for (uint64_t t = 0; t < (uint64_t)s.num_triangles * 3; t++) {
                int idx = d->indices[3 * (size_t)s.triangle_offset + t];
                if (idx < 0 || (uint32_t)idx >= s.num_vertices) BAD("scene: shape %llu has an out-of-range vertex index", (ull)i);
            }
            if (s.area_light_id >= 0 && (!in_tables(s.tri_cdf_offset, (uint64_t)s.num_triangles + 1) || !in_tables(s.tri_pmf_offset, s.num_triangles)))
                BAD("scene: shape %llu sampling tables lie outside the table array", (ull)i);
        }
    }
    for (uint64_t i = 0; i < d->num_lights; i++) {
        const lj_light &l = d->lights[i];
        if (l.kind == LJ_LIGHT_AREA) {
            if (l.shape_id < 0 || (uint64_t)l.shape_id >= d->num_shapes) BAD("scene: light %llu shape id out of range", (ull)i);
        } else if (l.kind == LJ_LIGHT_ENVMAP) {
            if (l.values_texture < 0 || (uint64_t)l.values_texture >= d->num_textures) BAD("scene: envmap %llu texture id out of range", (ull)i);
            if (l.dist_width <= 0 || l.dist_height <= 0) BAD("scene: envmap %llu has an empty sampling table", (ull)i);
            const uint64_t w = (uint64_t)l.dist_width, h = (uint64_t)l.dist_height;
            if (!in_tables(l.cdf_rows_offset, h * (w + 1)) || !in_tables(l.pdf_rows_offset, h * w) ||
                !in_tables(l.cdf_marginals_offset, h + 1) || !in_tables(l.pdf_marginals_offset, h))
                BAD("scene: envmap %llu sampling tables lie outside the table array", (ull)i);
        } else BAD("scene: light %llu has unknown kind %d", (ull)i, l.kind);
    }
    if (d->num_lights && (!in_tables(d->light_pmf_offset, d->num_lights) || !in_tables(d->light_cdf_offset, d->num_lights + 1)))
        BAD("scene: light distribution lies outside the table array");
    if ((d->options.integrator == LJ_INTEGRATOR_PATH || d->options.integrator == LJ_INTEGRATOR_VOLPATH) && d->num_lights == 0)
        BAD("scene: no lights (the reference asserts in sample(TableDist1D), table_dist.cpp:29)");
#undef BAD
    return LJ_OK;
}

extern "C" int lj_scene_create(lj_ctx *ctx, const lj_scene_desc *desc, lj_scene **out) {
    uint32_t flags = 0;
    if (const char *e = getenv("LJ_BUILD_TABLES")) if (atoi(e) != 0) flags |= LJ_SCENE_BUILD_TABLES;
    return lj_scene_create_ex(ctx, desc, flags, out);
}

extern "C" int lj_scene_create_ex(lj_ctx *ctx, const lj_scene_desc *desc, uint32_t flags, lj_scene **out) {
    if (!ctx || !desc || !out) { lj_set_error("lj_scene_create: null argument"); return LJ_EINVAL; }
    int rc = validate_desc(desc);
    if (rc) return rc;
    return ctx->group ? lj_group_scene_create(ctx, desc, flags, out) : lj_scene_create_single(ctx, desc, flags, out);
}

// `desc` has been validated
int lj_scene_create_single(lj_ctx *ctx, const lj_scene_desc *desc, uint32_t flags, lj_scene **out) {
    int rc;
    LJ_CUDA(cudaSetDevice(ctx->device));
    lj_scene *s = new lj_scene();
    s->ctx = ctx;
    memset(&s->build_stats, 0, sizeof s->build_stats);
    memset(&s->bvh, 0, sizeof s->bvh);
    memset(s->kind_present, 0, sizeof s->kind_present); s->grid_media = false;
    for (uint64_t i = 0; i < desc->num_materials; i++) s->kind_present[desc->materials[i].kind] = true;   // kinds validated above
    for (uint64_t i = 0; i < desc->num_media; i++) s->grid_media = s->grid_media || desc->media[i].kind == LJ_MEDIUM_HETEROGENEOUS;
    DScene &d = s->d;
    memset(&d, 0, sizeof d);
    cudaEvent_t e0, e1, e2;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
    cudaEventRecord(e0);
    d.camera = desc->camera; d.options = desc->options;
    d.bounds_radius = desc->bounds_radius;
    double e = desc->bounds_radius * 1e-5;            // scene.h:99-105
    d.eps = e < 0.01 ? e : 0.01;
    d.envmap_light_id = desc->envmap_light_id;
    d.num_lights = (int)desc->num_lights; d.num_shapes = (int)desc->num_shapes; d.num_materials = (int)desc->num_materials; d.num_media = (int)desc->num_media;
    // volume grids: f32 (one or three channels) where that is exact, see DScene::grid_f32
    std::vector<lj_medium> dev_media(desc->media, desc->media + desc->num_media);
    std::vector<float> grid_f32;
    bool need_f64_grids = false;
    if (!getenv("LJ_GRID_F64")) {
        for (lj_medium &m : dev_media) {
            if (m.kind != LJ_MEDIUM_HETEROGENEOUS) continue;
            // density and albedo on the SAME cells (resolution and box equal), at least one of them coloured: one record of
            // 8 floats per voxel (density rgb, 0, albedo rgb, 0), so that a collision fetches 8 corners x 32 contiguous
            // bytes with one cell computation instead of 2 x 8 x 3 scattered floats (storage marker 8, _pad2 = which half)
            if (m.density.kind == LJ_VOLUME_GRID && m.albedo.kind == LJ_VOLUME_GRID && !getenv("LJ_NO_GRID_PAIR") &&
                memcmp(m.density.resolution, m.albedo.resolution, sizeof m.density.resolution) == 0 &&
                memcmp(m.density.p_min, m.albedo.p_min, sizeof m.density.p_min) == 0 && memcmp(m.density.p_max, m.albedo.p_max, sizeof m.density.p_max) == 0) {
                const size_t nvox = (size_t)m.density.resolution[0] * m.density.resolution[1] * m.density.resolution[2];
                const double *gd = desc->grid_data + m.density.data_offset, *ga = desc->grid_data + m.albedo.data_offset;
                bool exact = true, grey = true;
                for (size_t i = 0; i < nvox * 3 && exact; i++) exact = (double)(float)gd[i] == gd[i] && (double)(float)ga[i] == ga[i];
                for (size_t i = 0; i < nvox && grey; i++) grey = gd[3 * i] == gd[3 * i + 1] && gd[3 * i] == gd[3 * i + 2] && ga[3 * i] == ga[3 * i + 1] && ga[3 * i] == ga[3 * i + 2];
                if (exact && !grey) {
                    while (grid_f32.size() % 8) grid_f32.push_back(0.f);      // 32-byte aligned records
                    m.density.data_offset = m.albedo.data_offset = grid_f32.size();
                    m.density._pad = m.albedo._pad = 8; m.density._pad2 = 0; m.albedo._pad2 = 1;
                    grid_f32.reserve(grid_f32.size() + nvox * 8);
                    for (size_t i = 0; i < nvox; i++) {
                        for (int c = 0; c < 3; c++) grid_f32.push_back((float)gd[3 * i + c]);
                        grid_f32.push_back(0.f);
                        for (int c = 0; c < 3; c++) grid_f32.push_back((float)ga[3 * i + c]);
                        grid_f32.push_back(0.f);
                    }
                    continue;
                }
            }
            lj_volume *vols[2] = {&m.albedo, &m.density};
            for (lj_volume *v : vols) {
                if (v->kind != LJ_VOLUME_GRID) continue;
                const size_t nvox = (size_t)v->resolution[0] * v->resolution[1] * v->resolution[2];
                const double *g = desc->grid_data + v->data_offset;
                bool exact = true, grey = true;
                for (size_t i = 0; i < nvox * 3 && exact; i++) exact = (double)(float)g[i] == g[i];
                for (size_t i = 0; i < nvox && grey; i++) grey = g[3 * i] == g[3 * i + 1] && g[3 * i] == g[3 * i + 2];
                if (!exact) { need_f64_grids = true; continue; }
                while (grid_f32.size() % 4) grid_f32.push_back(0.f);          // 16-byte aligned grids
                v->data_offset = grid_f32.size();
                v->_pad = grey ? 1 : 3;
                if (grey) for (size_t i = 0; i < nvox; i++) grid_f32.push_back((float)g[3 * i]);
                else for (size_t i = 0; i < nvox * 3; i++) grid_f32.push_back((float)g[i]);
            }
        }
    } else need_f64_grids = true;
    {   // grey scene: equal channels in every medium (host mirror of medium_is_grey, after the grids got their storage marker)
        auto vol_grey = [](const lj_volume &v) { return v.kind == LJ_VOLUME_CONSTANT ? (v.value[0] == v.value[1] && v.value[0] == v.value[2]) : v._pad == 1; };
        s->media_grey = !dev_media.empty();
        for (const lj_medium &m : dev_media) {
            bool g;
            if (m.kind == LJ_MEDIUM_HOMOGENEOUS)
                g = m.sigma_a[0] == m.sigma_a[1] && m.sigma_a[0] == m.sigma_a[2] && m.sigma_s[0] == m.sigma_s[1] && m.sigma_s[0] == m.sigma_s[2];
            else
                g = vol_grey(m.density) && vol_grey(m.albedo) &&
                    (m.density.kind == LJ_VOLUME_CONSTANT || (m.density.max_data[0] == m.density.max_data[1] && m.density.max_data[0] == m.density.max_data[2]));
            s->media_grey = s->media_grey && g;
        }
    }
    Uploader up;
#define UP(field, cnt) up.add(desc->field, (size_t)(cnt), &d.field);
    UP(shapes, desc->num_shapes) UP(lights, desc->num_lights) UP(materials, desc->num_materials)
    up.add(dev_media.data(), dev_media.size(), &d.media);
    up.add(grid_f32.data(), grid_f32.size(), &d.grid_f32);
    UP(textures, desc->num_textures) UP(images1, desc->num_images1) UP(images3, desc->num_images3)
    UP(positions, desc->num_positions * 3) UP(normals, desc->num_positions * 3) UP(uvs, desc->num_positions * 2)
    UP(indices, desc->num_indices * 3) up.add(desc->texels, (size_t)desc->num_texels, &d.texels, true); UP(tables, desc->num_tables) UP(grid_data, need_f64_grids ? desc->num_grid_data : 0)
#undef UP

    // primitive lists for the BVH builder
    std::vector<int> tri_shape(desc->num_indices, 0);
    std::vector<SpherePrim> spheres;
    for (uint64_t i = 0; i < desc->num_shapes; i++) {
        const lj_shape &sh = desc->shapes[i];
        if (sh.kind == LJ_SHAPE_MESH) {
            for (uint32_t t = 0; t < sh.num_triangles; t++) tri_shape[sh.triangle_offset + t] = (int)i;
        } else {
            SpherePrim sp; sp.cx = sh.center[0]; sp.cy = sh.center[1]; sp.cz = sh.center[2]; sp.r = sh.radius; sp.geom = (int)i; sp._pad = 0;
            spheres.push_back(sp);
        }
    }
    const int *d_tri_shape; const SpherePrim *d_spheres;
    up.add(tri_shape.data(), tri_shape.size(), &d_tri_shape);
    up.add(spheres.data(), spheres.size(), &d_spheres);
    if ((rc = lj_ctx_ring(ctx)) != LJ_OK) { cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); lj_scene_destroy_single(s); return rc; }
    cudaEvent_t late_done = ctx->ring_a[0];         // (no render is in flight during a scene create)
    if ((rc = up.commit(s, ctx->stream_b, late_done)) != LJ_OK) { cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); lj_scene_destroy_single(s); return rc; }
    d.light_pmf = d.tables + desc->light_pmf_offset;
    d.light_cdf = d.tables + desc->light_cdf_offset;
    cudaEventRecord(e1);
    BvhBuildInput bi;
    bi.num_tris = (int)desc->num_indices; bi.num_spheres = (int)spheres.size();
    bi.positions = d.positions; bi.indices = d.indices; bi.tri_shape = d_tri_shape; bi.shapes = d.shapes; bi.spheres = d_spheres;
    char berr[256] = "";
    if (lj_bvh_build(bi, s->bvh, berr) != 0) {
        lj_set_error("BVH build failed: %s", berr);
        cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); lj_scene_destroy_single(s); return LJ_ECUDA;
    }
    // Trav8 pushes at most two entries per level it descends (the rest of the node group and a postponed primitive
    // group, lj_traverse8.cuh), and drops work when its stack is full: a tree deeper than half the stack is refused
    // here instead of losing hits silently.  (PLOC trees of the shipped scenes: 6-9 levels; limit 16.)
    {
        int max_levels = LJ_BVH8_STACK / 2;
        if (const char *e = getenv("LJ_BVH8_MAX_LEVELS")) max_levels = atoi(e);      // test hook
        if (s->bvh.bvh8 && s->bvh.levels8 > max_levels) {
            lj_set_error("BVH is %d levels deep; the traversal stack (%d entries) covers %d levels", s->bvh.levels8, LJ_BVH8_STACK, max_levels);
            cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); lj_scene_destroy_single(s); return LJ_EUNSUPPORTED;
        }
    }
    cudaEventRecord(e2);
    // the texels have been arriving on the second stream meanwhile; everything below reads them
    cudaStreamWaitEvent(0, late_done, 0);
    if (flags & LJ_SCENE_BUILD_TABLES) {
        float tms = 0;
        if ((rc = lj_scene_build_tables(s, desc, (flags & LJ_SCENE_POISON_DERIVED) != 0, &tms)) != LJ_OK) {
            cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); lj_scene_destroy_single(s); return rc;
        }
        s->build_stats.tables_build_ms = tms;
    }

    if ((rc = lj_scene_build_shading_cache(s, desc, d_tri_shape)) != LJ_OK) {
        cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); lj_scene_destroy_single(s); return rc;
    }

    cudaEvent_t e3; cudaEventCreate(&e3);
    cudaEventRecord(e3);
    cudaEventSynchronize(e3);

    d.num_tris = bi.num_tris; d.num_spheres = bi.num_spheres; d.num_prims = s->bvh.num_prims;
    d.tri_verts = s->bvh.tri_verts; d.spheres = d_spheres;
    d.bvh2 = s->bvh.bvh2; d.bvh2_root = s->bvh.bvh2_root;
    d.bvh8 = s->bvh.bvh8;
    d.m15_bits = 0x47000000u;
    // LJ_BVH=2 keeps the binary tree for traversal (validation of the 8-wide path against it)
    if (const char *bv = getenv("LJ_BVH")) if (atoi(bv) == 2) d.bvh8 = nullptr;
    if (d.bvh8) d.tri_verts = s->bvh.tri_verts8;
    float ms = 0;
    // upload = the arrays the build needs (e0..e1) + whatever of the texel copy, the table build and the shading caches was
    // still to do when the BVH build finished (e2..e3)
    float ms_tail = 0;
    cudaEventElapsedTime(&ms, e0, e1); cudaEventElapsedTime(&ms_tail, e2, e3);
    s->build_stats.upload_ms = ms + ms_tail - s->build_stats.tables_build_ms;
    cudaEventElapsedTime(&ms, e1, e2); s->build_stats.bvh_build_ms = ms;
    cudaEventDestroy(e3);
    s->build_stats.bvh_nodes = s->bvh.bvh8 ? s->bvh.num_bvh8 : s->bvh.num_bvh2;
    s->build_stats.bvh_prims = s->bvh.num_prims;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    *out = s;
    return LJ_OK;
}

extern "C" void lj_scene_destroy(lj_scene *s) {
    if (!s) return;
    if (!s->replicas.empty()) lj_group_scene_destroy(s); else lj_scene_destroy_single(s);
}
void lj_scene_destroy_single(lj_scene *s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    if (s->ctx->ring_ready) cudaStreamSynchronize(s->ctx->stream_b);      // an upload of this scene may still be in flight there (failed create)
    for (void *p : s->allocs) lj_free(p);
    lj_bvh_free(s->bvh);
    delete s;
}

extern "C" int lj_scene_read_derived(lj_scene *s, double *tables, uint64_t num_tables, double *texels, uint64_t num_texels,
                                     lj_shape *shapes, uint64_t num_shapes, lj_light *lights, uint64_t num_lights) {
    if (!s) { lj_set_error("lj_scene_read_derived: null scene"); return LJ_EINVAL; }
    if (!s->replicas.empty()) s = s->replicas[0];
    LJ_CUDA(cudaSetDevice(s->ctx->device));
    if (tables && num_tables) LJ_CUDA(cudaMemcpy(tables, s->d.tables, sizeof(double) * num_tables, cudaMemcpyDeviceToHost));
    if (texels && num_texels) LJ_CUDA(cudaMemcpy(texels, s->d.texels, sizeof(double) * num_texels, cudaMemcpyDeviceToHost));
    if (shapes && num_shapes) LJ_CUDA(cudaMemcpy(shapes, s->d.shapes, sizeof(lj_shape) * num_shapes, cudaMemcpyDeviceToHost));
    if (lights && num_lights) LJ_CUDA(cudaMemcpy(lights, s->d.lights, sizeof(lj_light) * num_lights, cudaMemcpyDeviceToHost));
    return LJ_OK;
}

extern "C" int lj_scene_info(lj_scene *s, int32_t *width, int32_t *height, int32_t *spp, lj_stats *build_stats) {
    if (!s) { lj_set_error("lj_scene_info: null scene"); return LJ_EINVAL; }
    const lj_stats handle_stats = s->build_stats;
    const bool multi = !s->replicas.empty();
    if (multi) s = s->replicas[0];
    if (width) *width = s->d.camera.width;
    if (height) *height = s->d.camera.height;
    if (spp) *spp = s->d.options.samples_per_pixel;
    if (build_stats) *build_stats = multi ? handle_stats : s->build_stats;
    return LJ_OK;
}

// ---------------- render dispatch (render.cpp:155-170) ----------------
__global__ void k_finalize(int npix, const double *sum, double inv_spp_num, float *rgb) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix * 3) return;
    // render.cpp:94: img = radiance / Real(spp)  (vector / scalar = multiply by reciprocal)
    rgb[i] = (float)(sum[i] * inv_spp_num);
}

static int make_launch(lj_scene *s, const lj_render_params *params, RenderLaunch &rl) {
    lj_render_params p;
    if (params) p = *params; else { memset(&p, 0, sizeof p); p.max_depth_override = -100; p.rng_mode = LJ_RNG_SAMPLE_STREAMS; }
    rl.p = p;
    rl.spp = p.spp > 0 ? p.spp : s->d.options.samples_per_pixel;
    rl.s_begin = 0; rl.s_end = rl.spp;
    if (p.sample_end > p.sample_begin) { rl.s_begin = p.sample_begin; rl.s_end = p.sample_end; }
    if (rl.spp <= 0) { lj_set_error("render: spp must be positive"); return LJ_EINVAL; }
    if (rl.s_begin < 0 || rl.s_end > rl.spp) { lj_set_error("render: sample slice [%d,%d) outside [0,%d)", rl.s_begin, rl.s_end, rl.spp); return LJ_EINVAL; }
    const int integ = s->d.options.integrator;
    const bool sampling = integ == LJ_INTEGRATOR_PATH || integ == LJ_INTEGRATOR_VOLPATH;
    // the tile-serial engine consumes ONE stream per tile over all samples (render.cpp:82-96): a slice of it does not exist
    if ((rl.s_begin != 0 || rl.s_end != rl.spp) && (p.rng_mode == LJ_RNG_TILE_STREAMS || !sampling)) {
        lj_set_error("render: a sample slice needs per-sample streams (LJ_RNG_SAMPLE_STREAMS) and a sampling integrator"); return LJ_EINVAL;
    }
    if (p.tile_world < 0 || (p.tile_world > 0 && (p.tile_rank < 0 || p.tile_rank >= p.tile_world))) { lj_set_error("render: bad tile partition %d/%d", p.tile_rank, p.tile_world); return LJ_EINVAL; }
    rl.width = s->d.camera.width; rl.height = s->d.camera.height;
    rl.ntx = (rl.width + 15) / 16; rl.nty = (rl.height + 15) / 16;
    return LJ_OK;
}

extern "C" int lj_render_device(lj_scene *s, const lj_render_params *params, double *dev_sum, lj_stats *stats) {
    if (!s || !dev_sum) { lj_set_error("lj_render_device: null argument"); return LJ_EINVAL; }
    return s->replicas.empty() ? lj_render_device_single(s, params, dev_sum, stats) : lj_group_render_device(s, params, dev_sum, stats);
}

int lj_render_device_single(lj_scene *s, const lj_render_params *params, double *dev_sum, lj_stats *stats) {
    LJ_CUDA(cudaSetDevice(s->ctx->device));
    RenderLaunch rl;
    int rc = make_launch(s, params, rl);
    if (rc) return rc;
    lj_stats st; memset(&st, 0, sizeof st);
    st.upload_ms = s->build_stats.upload_ms; st.bvh_build_ms = s->build_stats.bvh_build_ms;
    st.bvh_nodes = s->build_stats.bvh_nodes; st.bvh_prims = s->build_stats.bvh_prims;
    DScene saved = s->d;
    if (rl.p.max_depth_override != -100) s->d.options.max_depth = rl.p.max_depth_override;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    int integ = s->d.options.integrator;
    if (integ == LJ_INTEGRATOR_PATH) {
        if (rl.p.rng_mode == LJ_RNG_TILE_STREAMS) rc = lj_render_tiles_mode_a(s, rl, dev_sum, &st);
        else if (rl.p.engine == LJ_ENGINE_MEGAKERNEL) rc = lj_render_megakernel(s, rl, dev_sum, &st);
        else rc = lj_render_wavefront(s, rl, dev_sum, &st);
    } else if (integ == LJ_INTEGRATOR_VOLPATH) {
        rc = lj_render_volpath(s, rl, dev_sum, &st);
    } else {
        rc = lj_render_aux(s, rl, dev_sum, &st);
    }
    cudaEventRecord(e1);
    cudaError_t ce = cudaEventSynchronize(e1);
    s->d = saved;
    if (rc == LJ_OK && ce != cudaSuccess) { lj_set_error("render kernels failed: %s", cudaGetErrorString(ce)); rc = LJ_ECUDA; }
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    st.render_ms = ms;
    st.n_devices = 1; st.partition = LJ_PARTITION_AUTO; st.reduce_ms = 0; st.device_ms_min = st.device_ms_max = ms;
    if (stats) *stats = st;
    return rc;
}

extern "C" int lj_render(lj_scene *s, const lj_render_params *params, float *rgb_out, lj_stats *stats) {
    if (!s || !rgb_out) { lj_set_error("lj_render: null argument"); return LJ_EINVAL; }
    lj_scene *s0 = s->replicas.empty() ? s : s->replicas[0];
    LJ_CUDA(cudaSetDevice(s0->ctx->device));
    const int npix = s0->d.camera.width * s0->d.camera.height;
    RenderLaunch rl;
    int rc = make_launch(s0, params, rl);
    if (rc) return rc;
    double *sum = nullptr; float *rgb = nullptr;
    LJ_CUDA(lj_malloc(&sum, sizeof(double) * 3 * (size_t)npix));
    if (lj_malloc(&rgb, sizeof(float) * 3 * (size_t)npix) != cudaSuccess) { lj_free(sum); lj_set_error("lj_render: out of device memory"); return LJ_ECUDA; }
    // the buffers are allocated in this thread's stream; a multi-device render writes them from its worker threads
    if (!s->replicas.empty()) cudaStreamSynchronize(0);
    rc = lj_render_device(s, params, sum, stats);
    if (rc == LJ_OK) {
        cudaSetDevice(s0->ctx->device);
        const int integ = s0->d.options.integrator;
        const double denom = (integ == LJ_INTEGRATOR_PATH || integ == LJ_INTEGRATOR_VOLPATH) ? (double)(rl.s_end - rl.s_begin) : 1.0;
        k_finalize<<<(npix * 3 + 255) / 256, 256>>>(npix, sum, 1.0 / denom, rgb);
        cudaError_t e = cudaMemcpy(rgb_out, rgb, sizeof(float) * 3 * (size_t)npix, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { lj_set_error("lj_render: D2H failed: %s", cudaGetErrorString(e)); rc = LJ_ECUDA; }
    }
    lj_free(sum); lj_free(rgb);
    return rc;
}

// ---------------- batch ray casting (rtcIntersect1 / rtcOccluded1) ----------------
__global__ void k_intersect_batch(DScene sc, const lj_ray *rays, long long n, lj_hit *hits, int any_hit) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    lj_ray r = rays[i];
    RayF rf; rf.ox = r.org[0]; rf.oy = r.org[1]; rf.oz = r.org[2]; rf.tnear = r.tnear;
    rf.dx = r.dir[0]; rf.dy = r.dir[1]; rf.dz = r.dir[2]; rf.tfar = r.tfar;
    lj_hit out; out.t = r.tfar; out.u = out.v = 0; out.geom_id = -1; out.prim_id = -1; out.ng[0] = out.ng[1] = out.ng[2] = 0;
    if (any_hit) {
        if (trace_any(sc, rf)) { out.t = -INFINITY; out.geom_id = 0; }
    } else {
        Hit h;
        if (trace_closest(sc, rf, &h)) {
            out.t = h.t; out.u = h.u; out.v = h.v; out.geom_id = h.geom; out.prim_id = h.prim;
            out.ng[0] = h.ngx; out.ng[1] = h.ngy; out.ng[2] = h.ngz;
        }
    }
    hits[i] = out;
}

extern "C" int lj_intersect_batch(lj_scene *s, const lj_ray *rays, int64_t n, lj_hit *hits, int any_hit) {
    if (!s || (n > 0 && (!rays || !hits))) { lj_set_error("lj_intersect_batch: null argument"); return LJ_EINVAL; }
    if (n <= 0) return LJ_OK;
    if (!s->replicas.empty()) s = s->replicas[0];
    LJ_CUDA(cudaSetDevice(s->ctx->device));
    lj_ray *dr = nullptr; lj_hit *dh = nullptr;
    LJ_CUDA(lj_malloc(&dr, sizeof(lj_ray) * n));
    cudaError_t e = lj_malloc(&dh, sizeof(lj_hit) * n);
    if (e == cudaSuccess) e = cudaMemcpy(dr, rays, sizeof(lj_ray) * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_intersect_batch<<<(unsigned)((n + 127) / 128), 128>>>(s->d, dr, n, dh, any_hit);
        e = cudaMemcpy(hits, dh, sizeof(lj_hit) * n, cudaMemcpyDeviceToHost);
    }
    lj_free(dr); lj_free(dh);
    if (e != cudaSuccess) { lj_set_error("lj_intersect_batch: %s", cudaGetErrorString(e)); return LJ_ECUDA; }
    return LJ_OK;
}

// ---------------- pinned host scene arrays ----------------
extern "C" int lj_scene_desc_pin(lj_scene_desc *d, int pin) {
    if (!d) { lj_set_error("lj_scene_desc_pin: null desc"); return LJ_EINVAL; }
    struct A { const void *p; size_t bytes; } arr[] = {
        {d->positions, d->num_positions * 3 * sizeof(double)}, {d->normals, d->num_positions * 3 * sizeof(double)},
        {d->uvs, d->num_positions * 2 * sizeof(double)}, {d->indices, d->num_indices * 3 * sizeof(int32_t)},
        {d->texels, d->num_texels * sizeof(double)}, {d->tables, d->num_tables * sizeof(double)},
        {d->grid_data, d->num_grid_data * sizeof(double)}};
    int rc = LJ_OK;
    for (auto &a : arr) {
        if (!a.p || a.bytes < 4096) continue;
        cudaError_t e = pin ? cudaHostRegister((void *)a.p, a.bytes, cudaHostRegisterDefault) : cudaHostUnregister((void *)a.p);
        if (e != cudaSuccess) { cudaGetLastError(); lj_set_error("lj_scene_desc_pin: %s", cudaGetErrorString(e)); rc = LJ_ECUDA; }
    }
    return rc;
}
