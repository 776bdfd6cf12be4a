// Scene build step on the device: everything Scene::Scene does besides the Embree build
// (scene.cpp:36-52), from the primary data already uploaded by lj_scene_create:
//   * mip chains           make_mipmap, mipmap.h:25-48         (2x2 box filter, up to 8 levels, f64)
//   * per-mesh area tables init_sampling_dist, shapes/triangle_mesh.inl:48-63 + make_table_dist_1d, table_dist.cpp:3-25
//   * envmap distribution  init_sampling_dist, lights/envmap.inl:75-98 + make_table_dist_2d, table_dist.cpp:40-114
//   * light distribution   light_power (diffuse_area_light.inl:1-3, envmap.inl:1-5) + make_table_dist_1d, scene.cpp:46-51
// The reference accumulates its CDFs with one running f64 sum; f64 addition does not reassociate, so the scans
// keep that order: a tile of the input is staged in shared memory by the whole block, one thread adds it up,
// the block writes the tile back.  Everything that is a pure map (areas, box filter, normalisation) is parallel.
// Results are bit-identical to the host-built tables except where the reference calls libm's sin (envmap rows).
#include "lj_internal.h"
#include "lj_alloc.h"
#include "lj_shading.cuh"
#include <vector>

namespace {

constexpr int TILE = 2048;

__global__ void k_fill(double *p, size_t n, double v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

// ---- make_mipmap (mipmap.h:25-48): level l from level l-1 of every image of one class, blockIdx.y = image ----
template <int C>
__global__ void k_mip_level(const lj_image *images, double *texels, int level) {
    const lj_image im = images[blockIdx.y];
    if (level >= im.num_levels) return;
    const int pw = im.width[level - 1], ph = im.height[level - 1];
    const int nw = im.width[level], nh = im.height[level];
    const double *prev = texels + im.offset[level - 1];
    double *next = texels + im.offset[level];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)nw * nh; i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % nw), y = (int)(i / nw);
        // Once a dimension has shrunk to 1 the reference's 2y+1 (2x+1) indexes past the image (mipmap.h:38-41 reads
        // beyond the buffer: undefined values end up in those levels); the edge texel is replicated here instead.
        const int x1 = min(2 * x + 1, pw - 1), y1 = min(2 * y + 1, ph - 1);
        const long long i00 = (long long)(2 * y) * pw + 2 * x, i10 = (long long)(2 * y) * pw + x1;
        const long long i01 = (long long)y1 * pw + 2 * x, i11 = (long long)y1 * pw + x1;
#pragma unroll
        for (int c = 0; c < C; c++) {
            const double s = ((prev[i00 * C + c] + prev[i10 * C + c]) + prev[i01 * C + c]) + prev[i11 * C + c];
            next[i * C + c] = s * (1.0 / 4.0);          // Vector3 / Real multiplies by the reciprocal (vector.h:193-197); exact for 4
        }
    }
}

// running sum of in[0..n) in the reference's order: out[0] = 0 (written by the caller), out[i + 1] = out[i] + in[i].
// Block-cooperative; returns the total to every thread.
__device__ double block_serial_scan(const double *in, double *out, int n, double *tile /* TILE + 1 doubles of shared memory */) {
    __shared__ double carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += TILE) {
        const int m = min(TILE, n - base);
        for (int i = threadIdx.x; i < m; i += blockDim.x) tile[i] = in[base + i];
        __syncthreads();
        if (threadIdx.x == 0) {
            double acc = carry;
            for (int i = 0; i < m; i++) { acc = acc + tile[i]; tile[i] = acc; }
            carry = acc;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < m; i += blockDim.x) out[base + i + 1] = tile[i];
        __syncthreads();
    }
    return carry;
}

// make_table_dist_1d's normalisation (table_dist.cpp:12-24): note that cdf[n] keeps the un-normalised total
__device__ void normalise_table_1d(double *pmf, double *cdf, int n, double total) {
    if (total > 0) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) { pmf[i] = pmf[i] / total; cdf[i] = cdf[i] / total; }
    } else {
        for (int i = threadIdx.x; i < n; i += blockDim.x) { pmf[i] = 1.0 / (double)n; cdf[i] = (double)i / (double)n; }
        if (threadIdx.x == 0) cdf[n] = 1;
    }
}

// ---- init_sampling_dist(TriangleMesh): one block per shape ----
__global__ void __launch_bounds__(256) k_mesh_tables(lj_shape *shapes, const double *positions, const int *indices, double *tables) {
    __shared__ double tile[TILE];
    lj_shape *sh = &shapes[blockIdx.x];
    if (sh->kind != LJ_SHAPE_MESH) return;
    const int n = (int)sh->num_triangles;
    double *pmf = tables + sh->tri_pmf_offset, *cdf = tables + sh->tri_cdf_offset;
    const double *pos = positions + 3 * (size_t)sh->vertex_offset;
    const int *idx = indices + 3 * (size_t)sh->triangle_offset;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        const d3 v0 = mk3(pos[3 * idx[3 * t]], pos[3 * idx[3 * t] + 1], pos[3 * idx[3 * t] + 2]);
        const d3 v1 = mk3(pos[3 * idx[3 * t + 1]], pos[3 * idx[3 * t + 1] + 1], pos[3 * idx[3 * t + 1] + 2]);
        const d3 v2 = mk3(pos[3 * idx[3 * t + 2]], pos[3 * idx[3 * t + 2] + 1], pos[3 * idx[3 * t + 2] + 2]);
        pmf[t] = length(cross(v1 - v0, v2 - v0)) / 2;
    }
    if (threadIdx.x == 0) cdf[0] = 0;
    __syncthreads();
    const double total = block_serial_scan(pmf, cdf, n, tile);   // == total_area: same additions in the same order
    if (threadIdx.x == 0) sh->total_area = total;
    __syncthreads();
    normalise_table_1d(pmf, cdf, n, total);
}

// ---- init_sampling_dist(Envmap), envmap.inl:75-98: f = luminance(texel lookup at the pixel centre) * sin(elevation) ----
__global__ void k_env_f(DScene sc, lj_image img, int w, int h, double *f) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)w * h) return;
    const int x = (int)(i % w), y = (int)(i / w);
    const double v = (y + 0.5) / (double)h;
    const double sin_elevation = sin(LJ_PI * v);
    const double u = (x + 0.5) / (double)w;
    f[i] = luminance(mip_lookup_level<d3>(sc, img, -1, u, v, 0)) * sin_elevation;   // (the f32 level-0 cache is built later)
}

// rows of make_table_dist_2d (table_dist.cpp:44-73): one block per row; pdf_rows holds f on entry
__global__ void __launch_bounds__(256) k_env_rows(int w, double *cdf_rows, double *pdf_rows) {
    __shared__ double tile[TILE];
    const int y = blockIdx.x;
    double *cdf = cdf_rows + (size_t)y * (w + 1), *pdf = pdf_rows + (size_t)y * w;
    if (threadIdx.x == 0) cdf[0] = 0;
    __syncthreads();
    const double integral = block_serial_scan(pdf, cdf, w, tile);
    if (integral > 0) {
        for (int x = threadIdx.x; x < w; x += blockDim.x) { cdf[x] = cdf[x] / integral; pdf[x] = pdf[x] / integral; }
        // cdf[w] stays the row integral until the marginal has been built
    } else {
        for (int x = threadIdx.x; x < w; x += blockDim.x) { pdf[x] = 1.0 / (double)w; cdf[x] = (double)x / (double)w; }
        if (threadIdx.x == 0) cdf[w] = 1;
    }
}

// marginal of make_table_dist_2d (table_dist.cpp:74-107): one block
__global__ void __launch_bounds__(256) k_env_marginal(lj_light *light, int w, int h, double *cdf_rows, double *cdf_m, double *pdf_m) {
    __shared__ double tile[TILE];
    // the row weights, gathered into pdf_m first
    for (int y = threadIdx.x; y < h; y += blockDim.x) pdf_m[y] = cdf_rows[(size_t)y * (w + 1) + w];
    if (threadIdx.x == 0) cdf_m[0] = 0;
    __syncthreads();
    const double total = block_serial_scan(pdf_m, cdf_m, h, tile);
    if (total > 0) {
        for (int y = threadIdx.x; y < h; y += blockDim.x) { cdf_m[y] = cdf_m[y] / total; pdf_m[y] = pdf_m[y] / total; }
    } else {
        for (int y = threadIdx.x; y < h; y += blockDim.x) { pdf_m[y] = 1.0 / (double)h; cdf_m[y] = (double)y / (double)h; }
    }
    __syncthreads();
    if (threadIdx.x == 0) { cdf_m[h] = 1; light->total_values = total; }
    for (int y = threadIdx.x; y < h; y += blockDim.x) cdf_rows[(size_t)y * (w + 1) + w] = 1;
}

// ---- scene.cpp:46-51: light_power of every light, then make_table_dist_1d ----
__global__ void k_light_dist(const lj_light *lights, const lj_shape *shapes, int n, double bounds_radius, double *pmf, double *cdf) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    cdf[0] = 0;
    for (int i = 0; i < n; i++) {
        const lj_light &l = lights[i];
        double power;
        if (l.kind == LJ_LIGHT_AREA) {
            const lj_shape &sh = shapes[l.shape_id];
            const double area = sh.kind == LJ_SHAPE_MESH ? sh.total_area : 4 * LJ_PI * sh.radius * sh.radius;
            power = luminance(mk3(l.intensity[0], l.intensity[1], l.intensity[2])) * area * LJ_PI;
        } else {
            power = LJ_PI * bounds_radius * bounds_radius * l.total_values / (double)(l.dist_width * l.dist_height);
        }
        pmf[i] = power;
        cdf[i + 1] = cdf[i] + power;
    }
    const double total = cdf[n];
    if (total > 0) {
        for (int i = 0; i < n; i++) { pmf[i] = pmf[i] / total; cdf[i] = cdf[i] / total; }
    } else {
        for (int i = 0; i < n; i++) { pmf[i] = 1.0 / (double)n; cdf[i] = (double)i / (double)n; }
        cdf[n] = 1;
    }
}

__global__ void k_poison_records(lj_shape *shapes, int ns, lj_light *lights, int nl, double v) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ns && shapes[i].kind == LJ_SHAPE_MESH) shapes[i].total_area = v;
    if (i < nl && lights[i].kind == LJ_LIGHT_ENVMAP) lights[i].total_values = v;
}

}  // namespace

// Rebuilds every derived table of the scene on the device.  `desc` supplies the host copies of the record
// arrays (ranges and offsets only; the derived values in it are not read).  With `poison` the derived regions
// are first overwritten with NaN, so that a later comparison proves the kernels produced every value.
int lj_scene_build_tables(lj_scene *s, const lj_scene_desc *desc, bool poison, float *elapsed_ms) {
    DScene &d = s->d;
    double *tables = const_cast<double *>(d.tables), *texels = const_cast<double *>(d.texels);
    lj_shape *shapes = const_cast<lj_shape *>(d.shapes);
    lj_light *lights = const_cast<lj_light *>(d.lights);
    const int ns = (int)desc->num_shapes, nl = (int)desc->num_lights;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    const double nan = __builtin_nan("");
    auto fill = [&](double *p, size_t n) { if (n) k_fill<<<(unsigned)std::min<size_t>((n + 255) / 256, 1024), 256>>>(p, n, nan); };

    if (poison) {
        for (int c = 0; c < 2; c++) {
            const lj_image *imgs = c ? desc->images3 : desc->images1;
            const uint64_t ni = c ? desc->num_images3 : desc->num_images1;
            for (uint64_t i = 0; i < ni; i++)
                for (int l = 1; l < imgs[i].num_levels && l < 8; l++)
                    fill(texels + imgs[i].offset[l], (size_t)imgs[i].width[l] * imgs[i].height[l] * imgs[i].channels);
        }
        for (int i = 0; i < ns; i++) if (desc->shapes[i].kind == LJ_SHAPE_MESH) {
            fill(tables + desc->shapes[i].tri_pmf_offset, desc->shapes[i].num_triangles);
            fill(tables + desc->shapes[i].tri_cdf_offset, (size_t)desc->shapes[i].num_triangles + 1);
        }
        for (int i = 0; i < nl; i++) if (desc->lights[i].kind == LJ_LIGHT_ENVMAP && desc->lights[i].dist_width > 0) {
            const lj_light &l = desc->lights[i];
            const size_t w = l.dist_width, h = l.dist_height;
            fill(tables + l.cdf_rows_offset, h * (w + 1)); fill(tables + l.pdf_rows_offset, h * w);
            fill(tables + l.cdf_marginals_offset, h + 1); fill(tables + l.pdf_marginals_offset, h);
        }
        fill(tables + desc->light_pmf_offset, nl); fill(tables + desc->light_cdf_offset, (size_t)nl + 1);
        const int nrec = ns > nl ? ns : nl;
        if (nrec) k_poison_records<<<(nrec + 127) / 128, 128>>>(shapes, ns, lights, nl, nan);
    }

    // mip chains: level by level, all images of a class in one launch
    for (int c = 0; c < 2; c++) {
        const lj_image *imgs = c ? desc->images3 : desc->images1;
        const uint64_t ni = c ? desc->num_images3 : desc->num_images1;
        int max_levels = 0;
        for (uint64_t i = 0; i < ni; i++) max_levels = imgs[i].num_levels > max_levels ? imgs[i].num_levels : max_levels;
        for (int l = 1; l < max_levels && l < 8; l++) {
            const dim3 grid(64, (unsigned)ni);
            if (c) k_mip_level<3><<<grid, 256>>>(d.images3, texels, l);
            else k_mip_level<1><<<grid, 256>>>(d.images1, texels, l);
        }
    }
    // mesh area tables
    if (ns) k_mesh_tables<<<ns, 256>>>(shapes, d.positions, d.indices, tables);
    // envmap distributions (need the level-0 texels only)
    for (int i = 0; i < nl; i++) {
        const lj_light &l = desc->lights[i];
        if (l.kind != LJ_LIGHT_ENVMAP || l.dist_width <= 0 || l.values_texture < 0) continue;
        const lj_texture &t = desc->textures[l.values_texture];
        if (t.kind != LJ_TEX_IMAGE) continue;
        const lj_image &img = desc->images3[t.image_id];
        const int w = l.dist_width, h = l.dist_height;
        k_env_f<<<(unsigned)(((long long)w * h + 255) / 256), 256>>>(d, img, w, h, tables + l.pdf_rows_offset);
        k_env_rows<<<h, 256>>>(w, tables + l.cdf_rows_offset, tables + l.pdf_rows_offset);
        k_env_marginal<<<1, 256>>>(lights + i, w, h, tables + l.cdf_rows_offset, tables + l.cdf_marginals_offset, tables + l.pdf_marginals_offset);
    }
    // light distribution
    if (nl) k_light_dist<<<1, 32>>>(lights, shapes, nl, d.bounds_radius, tables + desc->light_pmf_offset, tables + desc->light_cdf_offset);
    LJ_CUDA(cudaGetLastError());
    cudaEventRecord(e1);
    LJ_CUDA(cudaEventSynchronize(e1));
    if (elapsed_ms) cudaEventElapsedTime(elapsed_ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return LJ_OK;
}

// ---------------- shading caches (DScene::tri_shade, texels0_*) ----------------
namespace {

// one 192-byte record per triangle: p0 p1 p2 | n0 n1 n2 | uv0 uv1 uv2 (f64; zeros where the mesh has no normals / uvs)
__global__ void k_tri_shade(int num_tris, const int *tri_shape, const lj_shape *shapes, const int *indices,
                            const double *positions, const double *normals, const double *uvs, double *out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_tris) return;
    const lj_shape &s = shapes[tri_shape[i]];
    double *o = out + 24 * (size_t)i;
    for (int k = 0; k < 3; k++) {
        const size_t v = (size_t)s.vertex_offset + indices[3 * (size_t)i + k];
        for (int a = 0; a < 3; a++) {
            o[3 * k + a] = positions[3 * v + a];
            o[9 + 3 * k + a] = s.has_normals ? normals[3 * v + a] : 0.0;
        }
        for (int a = 0; a < 2; a++) o[18 + 2 * k + a] = s.has_uvs ? uvs[2 * v + a] : 0.0;
    }
}

struct Tex0Image { unsigned long long src; unsigned start, count; };    // level-0 texels in `texels` (doubles), first texel / texel count in the f32 array

// image of the t-th level-0 texel (images are few: linear search over the prefix table)
__device__ __forceinline__ int tex0_image_of(const Tex0Image *imgs, int n, unsigned t) {
    int i = 0;
    while (i + 1 < n && t >= imgs[i + 1].start) i++;
    return i;
}
template <int C>
__global__ void k_tex0_check(const Tex0Image *imgs, int n, unsigned total, const double *texels, int *inexact) {
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
        const int i = tex0_image_of(imgs, n, t);
        const double *p = texels + imgs[i].src + (size_t)C * (t - imgs[i].start);
        bool ok = true;
        for (int c = 0; c < C; c++) ok = ok && (double)(float)p[c] == p[c];
        if (!ok) inexact[i] = 1;
    }
}
template <int C>
__global__ void k_tex0_convert(const Tex0Image *imgs, int n, unsigned total, const double *texels, float4 *rgb, float *r) {
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
        const int i = tex0_image_of(imgs, n, t);
        const double *p = texels + imgs[i].src + (size_t)C * (t - imgs[i].start);
        if (C == 3) rgb[t] = make_float4((float)p[0], (float)p[1], (float)p[2], 0.f);
        else r[t] = (float)p[0];
    }
}
__global__ void k_tex0_offsets(const Tex0Image *imgs, int n, const int *inexact, unsigned *off) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) off[i] = inexact[i] ? 0xffffffffu : imgs[i].start;
}

}  // namespace

// Builds DScene::tri_shade and the f32 level-0 texel arrays; everything stays on the device (the exactness of an image
// is decided by a kernel and lands in tex0_off*, the host never waits).  LJ_NO_SHADE_CACHE=1 skips both (A/B runs).
int lj_scene_build_shading_cache(lj_scene *s, const lj_scene_desc *desc, const int *d_tri_shape) {
    DScene &d = s->d;
    d.tri_shade = nullptr; d.texels0_rgb = nullptr; d.texels0_r = nullptr; d.tex0_off3 = nullptr; d.tex0_off1 = nullptr;
    if (getenv("LJ_NO_SHADE_CACHE")) return LJ_OK;
    const int nt = (int)desc->num_indices;
    if (nt > 0) {
        double *ts = nullptr;
        LJ_CUDA(lj_malloc(&ts, sizeof(double) * 24 * (size_t)nt));
        s->allocs.push_back(ts);
        k_tri_shade<<<(nt + 127) / 128, 128>>>(nt, d_tri_shape, d.shapes, d.indices, d.positions, d.normals, d.uvs, ts);
        LJ_CUDA(cudaGetLastError());
        d.tri_shade = ts;
    }
    for (int c = 0; c < 2; c++) {
        const lj_image *imgs = c ? desc->images3 : desc->images1;
        const int ni = (int)(c ? desc->num_images3 : desc->num_images1);
        if (ni == 0) continue;
        std::vector<Tex0Image> tab(ni);
        unsigned long long total = 0;
        for (int i = 0; i < ni; i++) {
            tab[i].src = imgs[i].offset[0]; tab[i].start = (unsigned)total; tab[i].count = (unsigned)((size_t)imgs[i].width[0] * imgs[i].height[0]);
            total += tab[i].count;
        }
        if (total == 0 || total >= 0xffffffffull) continue;
        Tex0Image *d_tab = nullptr; int *d_inexact = nullptr; unsigned *d_off = nullptr; void *d_tex = nullptr;
        LJ_CUDA(lj_malloc(&d_tab, sizeof(Tex0Image) * ni)); s->allocs.push_back(d_tab);
        LJ_CUDA(lj_malloc(&d_inexact, sizeof(int) * ni)); s->allocs.push_back(d_inexact);
        LJ_CUDA(lj_malloc(&d_off, sizeof(unsigned) * ni)); s->allocs.push_back(d_off);
        LJ_CUDA(lj_malloc(&d_tex, (c ? sizeof(float4) : sizeof(float)) * (size_t)total)); s->allocs.push_back(d_tex);
        LJ_CUDA(cudaMemcpyAsync(d_tab, tab.data(), sizeof(Tex0Image) * ni, cudaMemcpyHostToDevice));
        LJ_CUDA(cudaMemsetAsync(d_inexact, 0, sizeof(int) * ni));
        const unsigned grid = (unsigned)std::min<unsigned long long>((total + 255) / 256, (unsigned long long)s->ctx->sm_count * 16);
        if (c) {
            k_tex0_check<3><<<grid, 256>>>(d_tab, ni, (unsigned)total, d.texels, d_inexact);
            k_tex0_convert<3><<<grid, 256>>>(d_tab, ni, (unsigned)total, d.texels, (float4 *)d_tex, nullptr);
        } else {
            k_tex0_check<1><<<grid, 256>>>(d_tab, ni, (unsigned)total, d.texels, d_inexact);
            k_tex0_convert<1><<<grid, 256>>>(d_tab, ni, (unsigned)total, d.texels, nullptr, (float *)d_tex);
        }
        k_tex0_offsets<<<(ni + 63) / 64, 64>>>(d_tab, ni, d_inexact, d_off);
        LJ_CUDA(cudaGetLastError());
        if (c) { d.texels0_rgb = (const float4 *)d_tex; d.tex0_off3 = d_off; }
        else { d.texels0_r = (const float *)d_tex; d.tex0_off1 = d_off; }
    }
    return LJ_OK;
}
