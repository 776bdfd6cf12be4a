// Whole-path engines: the tile-serial parity kernel (mode A), the megakernel
// (mode B, one lane per pixel-sample slot) and the primary-ray AOV renderers.
// The wavefront engine lives in lj_wavefront.cu.
#include "lj_internal.h"
#include "lj_alloc.h"
#include "lj_path.cuh"
#include "lj_volpath.cuh"

typedef d3 (*SampleFnDummy)(const DScene &, int, int, Pcg32 &, RayCounters &);

template <int INTEGRATOR>
LJ_D d3 radiance_sample(const DScene &sc, int x, int y, Pcg32 &rng, RayCounters &rc, bool &keep) {
    keep = true;
    if (INTEGRATOR == LJ_INTEGRATOR_PATH) return path_tracing_sample(sc, x, y, rng, rc);
    d3 L = vol_path_tracing_dispatch(sc, x, y, rng, rc);
    // render.cpp:140 with vector.h:289-292: a sample is dropped only if ALL channels are non-finite
    keep = isfinite(L.x) || isfinite(L.y) || isfinite(L.z);
    return L;
}

LJ_D void flush_counters(const RayCounters &rc, unsigned long long *counters) {
    if (rc.primary) atomicAdd(&counters[0], rc.primary);
    if (rc.extension) atomicAdd(&counters[1], rc.extension);
    if (rc.shadow) atomicAdd(&counters[2], rc.shadow);
}

// Mode A: one lane per 16x16 tile, reproducing render.cpp:80-96 exactly: the
// tile's pcg32 stream is consumed pixel-major / sample-minor.
template <int INTEGRATOR>
__global__ void k_tiles_mode_a(DScene sc, int ntx, int nty, int spp, int tile_rank, int tile_world,
                               double *sum, unsigned long long *counters) {
    int tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= ntx * nty) return;
    int tx = tile % ntx, ty = tile / ntx;
    if (tile_world > 0 && (tile_dealt_index(tx, ty, ntx) % tile_world) != tile_rank) return;
    Pcg32 rng = pcg_init((uint64_t)(ty * ntx + tx));
    int w = sc.camera.width, h = sc.camera.height;
    int x0 = tx * 16, x1 = min(x0 + 16, w), y0 = ty * 16, y1 = min(y0 + 16, h);
    RayCounters rc; rc.primary = rc.extension = rc.shadow = 0;
    for (int y = y0; y < y1; y++) for (int x = x0; x < x1; x++) {
        d3 radiance = mk3(0.0);
        for (int s = 0; s < spp; s++) {
            bool keep;
            d3 L = radiance_sample<INTEGRATOR>(sc, x, y, rng, rc, keep);
            if (keep) radiance += L;
        }
        size_t o = 3 * ((size_t)y * w + x);
        sum[o] = radiance.x; sum[o + 1] = radiance.y; sum[o + 2] = radiance.z;
    }
    flush_counters(rc, counters);
}

// Mode B megakernel: lane (pixel, k) renders samples s_begin+k, s_begin+k+K, ...
template <int INTEGRATOR>
__global__ void k_mega_mode_b(DScene sc, int ntx, int s_begin, int s_end, int K, int tile_rank, int tile_world,
                              double *partial, unsigned long long *counters) {
    long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int w = sc.camera.width, h = sc.camera.height;
    long long npix = (long long)w * h;
    if (gid >= npix * K) return;
    int k = (int)(gid / npix);
    int pix = (int)(gid % npix);
    int x = pix % w, y = pix / w;
    d3 radiance = mk3(0.0);
    RayCounters rc; rc.primary = rc.extension = rc.shadow = 0;
    bool mine = true;
    if (tile_world > 0) mine = (tile_dealt_index(x / 16, y / 16, ntx) % tile_world) == tile_rank;
    if (mine) {
        for (int s = s_begin + k; s < s_end; s += K) {
            Pcg32 rng = pcg_init((uint64_t)s * (uint64_t)npix + (uint64_t)pix);
            bool keep;
            d3 L = radiance_sample<INTEGRATOR>(sc, x, y, rng, rc, keep);
            if (keep) radiance += L;
        }
    }
    size_t o = 3 * ((size_t)k * npix + pix);
    partial[o] = radiance.x; partial[o + 1] = radiance.y; partial[o + 2] = radiance.z;
    flush_counters(rc, counters);
}

__global__ void k_sum_partials(long long n3, int K, const double *partial, double *sum) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    double a = 0;
    for (int k = 0; k < K; k++) a += partial[(size_t)k * n3 + i];
    sum[i] = a;
}

static int finish_counters(unsigned long long *d_counters, lj_stats *st) {
    unsigned long long c[3];
    LJ_CUDA(cudaMemcpy(c, d_counters, sizeof c, cudaMemcpyDeviceToHost));
    st->rays_primary += c[0]; st->rays_extension += c[1]; st->rays_shadow += c[2];
    lj_free(d_counters);
    return LJ_OK;
}

template <int INTEGRATOR>
static int run_tiles(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    size_t npix = (size_t)rl.width * rl.height;
    unsigned long long *d_counters;
    LJ_CUDA(lj_malloc(&d_counters, 3 * sizeof(unsigned long long)));
    LJ_CUDA(cudaMemset(d_counters, 0, 3 * sizeof(unsigned long long)));
    LJ_CUDA(cudaMemsetAsync(dev_sum, 0, sizeof(double) * 3 * npix));
    int ntiles = rl.ntx * rl.nty;
    // one warp-lane per tile, small blocks so the tiles spread over all SMs
    k_tiles_mode_a<INTEGRATOR><<<(ntiles + 31) / 32, 32>>>(s->d, rl.ntx, rl.nty, rl.spp, rl.p.tile_rank, rl.p.tile_world, dev_sum, d_counters);
    LJ_CUDA(cudaGetLastError());
    st->kernel_launches += 1;
    st->samples = npix * (size_t)rl.spp;
    return finish_counters(d_counters, st);
}

template <int INTEGRATOR>
static int run_mega(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    size_t npix = (size_t)rl.width * rl.height;
    int nsamp = rl.s_end - rl.s_begin;
    int K = nsamp < 8 ? nsamp : 8;
    unsigned long long *d_counters; double *partial;
    LJ_CUDA(lj_malloc(&d_counters, 3 * sizeof(unsigned long long)));
    LJ_CUDA(cudaMemset(d_counters, 0, 3 * sizeof(unsigned long long)));
    LJ_CUDA(lj_malloc(&partial, sizeof(double) * 3 * npix * K));
    long long total = (long long)npix * K;
    k_mega_mode_b<INTEGRATOR><<<(unsigned)((total + 127) / 128), 128>>>(s->d, rl.ntx, rl.s_begin, rl.s_end, K, rl.p.tile_rank, rl.p.tile_world, partial, d_counters);
    LJ_CUDA(cudaGetLastError());
    k_sum_partials<<<(unsigned)((npix * 3 + 255) / 256), 256>>>((long long)npix * 3, K, partial, dev_sum);
    LJ_CUDA(cudaGetLastError());
    st->kernel_launches += 2;
    st->samples = npix * (size_t)nsamp;
    int rc = finish_counters(d_counters, st);
    lj_free(partial);
    return rc;
}

int lj_render_tiles_mode_a(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    return run_tiles<LJ_INTEGRATOR_PATH>(s, rl, dev_sum, st);
}
int lj_render_megakernel(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    return run_mega<LJ_INTEGRATOR_PATH>(s, rl, dev_sum, st);
}
int lj_render_volpath(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    if (rl.p.rng_mode == LJ_RNG_TILE_STREAMS) return run_tiles<LJ_INTEGRATOR_VOLPATH>(s, rl, dev_sum, st);
    // versions 1 and 2 are one or two rays per sample: nothing to gain from stages
    const int v = s->d.options.vol_path_version;
    if (rl.p.engine == LJ_ENGINE_WAVEFRONT && v != 1 && v != 2 && s->d.bvh8) return lj_render_volwave(s, rl, dev_sum, st);
    return run_mega<LJ_INTEGRATOR_VOLPATH>(s, rl, dev_sum, st);
}

// ---------------- primary-ray AOV integrators (render.cpp:12-69) ----------------
__global__ void k_aux(DScene sc, int ntx, int tile_rank, int tile_world, double *sum) {
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    int w = sc.camera.width, h = sc.camera.height;
    if (pix >= w * h) return;
    int x = pix % w, y = pix / w;
    if (tile_world > 0 && (tile_dealt_index(x / 16, y / 16, ntx) % tile_world) != tile_rank) {     // another rank's tile
        sum[3 * (size_t)pix] = 0; sum[3 * (size_t)pix + 1] = 0; sum[3 * (size_t)pix + 2] = 0;
        return;
    }
    Ray ray = sample_primary(sc.camera, mk2((x + 0.5) / w, (y + 0.5) / h));
    RayDiff rd = init_ray_differential(w, h);
    d3 color = mk3(0.0);
    PathVertex v;
    if (scene_intersect(sc, ray, rd, v)) {
        int integ = sc.options.integrator;
        if (integ == LJ_INTEGRATOR_DEPTH) { double dist = distance(v.position, ray.org); color = mk3(dist); }
        else if (integ == LJ_INTEGRATOR_SHADING_NORMAL) color = v.shading_frame.n;
        else if (integ == LJ_INTEGRATOR_MEAN_CURVATURE) color = mk3(v.mean_curvature);
        else if (integ == LJ_INTEGRATOR_RAY_DIFFERENTIAL) color = mk3(rd.radius, rd.spread, 0.0);
        else if (integ == LJ_INTEGRATOR_MIPMAP_LEVEL) {
            // get_texture(material): slot 0 of every material except clearcoat (constant zero)
            const lj_material &m = sc.materials[v.material_id];
            int t = m.kind == LJ_MAT_DISNEY_CLEARCOAT ? -1 : m.tex[0];
            if (t >= 0 && sc.textures[t].kind == LJ_TEX_IMAGE) {
                const lj_texture &tex = sc.textures[t];
                const lj_image &img = sc.images3[tex.image_id];
                int wh = img.width[0] > img.height[0] ? img.width[0] : img.height[0];
                double scaled = wh * lmax(tex.uscale, tex.vscale) * v.uv_screen_size;
                color = mk3(log2(lmax(scaled, (double)1e-8f)));
            }
        }
    }
    sum[3 * (size_t)pix] = color.x; sum[3 * (size_t)pix + 1] = color.y; sum[3 * (size_t)pix + 2] = color.z;
}

int lj_render_aux(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    int npix = rl.width * rl.height;
    k_aux<<<(npix + 127) / 128, 128>>>(s->d, rl.ntx, rl.p.tile_rank, rl.p.tile_world, dev_sum);
    LJ_CUDA(cudaGetLastError());
    st->kernel_launches += 1; st->samples = npix; st->rays_primary = npix;
    return LJ_OK;
}
