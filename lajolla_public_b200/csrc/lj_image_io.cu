// Output step: what the reference does after render() returns -- imwrite(), image.cpp:135-173 --
// from the device buffer, plus save / resume of the f64 accumulation for progressive multi-pass renders.
//
//   PFM : "PF\n<w> <h>\n-1\n" followed by the f32 RGB rows, top to bottom (image.cpp:141-149; the reference
//         does not flip the rows, and neither do we).
//   EXR : what tinyexr's SaveEXR(data, w, h, 3, /*fp16*/1, ...) stores (image.cpp:156-172): scanline file, channels
//         B, G, R as HALF, float -> half with tinyexr's own rounding (tinyexr.h:889-924: denormal floats flush
//         to zero, round half up on mantissa bit 12, carry may run into the exponent / infinity).  The
//         reference deflates 16-line blocks (ZIP); this writer stores the lines uncompressed, which every EXR
//         reader decodes to the same pixels.
// The f64 sum -> f32 -> half conversion and the planar scanline layout are done on the device; the host
// only writes the ~300-byte header and the line offset table.
#include "lj_internal.h"
#include <mutex>
#include "lj_alloc.h"
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

// tinyexr.h:889-924 float_to_half_full, restated on the bit pattern
__device__ __forceinline__ unsigned short half_bits_tinyexr(float f) {
    const unsigned u = __float_as_uint(f);
    const unsigned sign = (u >> 16) & 0x8000u;
    const unsigned ex = (u >> 23) & 0xffu;
    const unsigned man = u & 0x7fffffu;
    unsigned o = 0;                                         // exponent << 10 | mantissa
    if (ex == 0) o = 0;                                     // zero / float denormal
    else if (ex == 255) o = (31u << 10) | (man ? 0x200u : 0u);
    else {
        const int newexp = (int)ex - 127 + 15;
        if (newexp >= 31) o = 31u << 10;                    // overflow -> infinity
        else if (newexp <= 0) {                             // half denormal or zero
            if (14 - newexp <= 24) {
                const unsigned mant = man | 0x800000u;
                o = mant >> (14 - newexp);
                if ((mant >> (13 - newexp)) & 1u) o++;
            }
        } else {
            o = ((unsigned)newexp << 10) | (man >> 13);
            if (man & 0x1000u) o++;
        }
    }
    return (unsigned short)(sign | (o & 0x7fffu));
}

// render.cpp:94 + image.cpp:146: sum / spp (multiply by the reciprocal), then narrowed to float
__global__ void k_accum_to_f32(long long n, const double *sum, double inv_spp, float *rgb) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) rgb[i] = (float)(sum[i] * inv_spp);
}

// One EXR scanline chunk per image row: {int32 y, int32 bytes, half B[w], half G[w], half R[w]}, written as
// 16-bit words because a chunk starts at a multiple of 2 bytes only.
__global__ void k_exr_lines(const float *rgb, int w, int h, unsigned short *body) {
    const long long words_per_line = 4 + 3LL * w;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= words_per_line * h) return;
    const int y = (int)(i / words_per_line);
    const int k = (int)(i % words_per_line);
    unsigned short v;
    if (k < 4) {
        const unsigned word = k < 2 ? (unsigned)y : (unsigned)(6 * w);
        v = (unsigned short)((k & 1) ? (word >> 16) : (word & 0xffffu));
    } else {
        const int c = (k - 4) / w, x = (k - 4) % w;         // c: 0 = B, 1 = G, 2 = R
        v = half_bits_tinyexr(rgb[((long long)y * w + x) * 3 + (2 - c)]);
    }
    body[i] = v;
}

void put(std::vector<unsigned char> &b, const void *p, size_t n) { const unsigned char *c = (const unsigned char *)p; b.insert(b.end(), c, c + n); }
void put_str(std::vector<unsigned char> &b, const char *s) { put(b, s, strlen(s) + 1); }
void put_i32(std::vector<unsigned char> &b, int v) { put(b, &v, 4); }
void put_f32(std::vector<unsigned char> &b, float v) { put(b, &v, 4); }
void put_attr(std::vector<unsigned char> &b, const char *name, const char *type, int size) { put_str(b, name); put_str(b, type); put_i32(b, size); }

// header of a single-part scanline file with the attributes (and their order) tinyexr writes
std::vector<unsigned char> exr_header(int w, int h) {
    std::vector<unsigned char> b;
    const unsigned char magic[8] = {0x76, 0x2f, 0x31, 0x01, 2, 0, 0, 0};
    put(b, magic, 8);
    put_attr(b, "channels", "chlist", 3 * 18 + 1);
    for (const char *name : {"B", "G", "R"}) {
        put_str(b, name);                                   // 2 bytes
        put_i32(b, 1);                                      // HALF
        const unsigned char lin[4] = {0, 0, 0, 0}; put(b, lin, 4);
        put_i32(b, 1); put_i32(b, 1);                       // x / y sampling
    }
    b.push_back(0);
    put_attr(b, "compression", "compression", 1); b.push_back(0);       // NO_COMPRESSION
    put_attr(b, "dataWindow", "box2i", 16); put_i32(b, 0); put_i32(b, 0); put_i32(b, w - 1); put_i32(b, h - 1);
    put_attr(b, "displayWindow", "box2i", 16); put_i32(b, 0); put_i32(b, 0); put_i32(b, w - 1); put_i32(b, h - 1);
    put_attr(b, "lineOrder", "lineOrder", 1); b.push_back(0);           // increasing y
    put_attr(b, "pixelAspectRatio", "float", 4); put_f32(b, 1.0f);
    put_attr(b, "screenWindowCenter", "v2f", 8); put_f32(b, 0.0f); put_f32(b, 0.0f);
    put_attr(b, "screenWindowWidth", "float", 4); put_f32(b, 1.0f);
    b.push_back(0);
    return b;
}

std::string pfm_header(int w, int h) {
    char buf[64];
    snprintf(buf, sizeof buf, "PF\n%d %d\n-1\n", w, h);
    return buf;
}

bool ends_with(const std::string &s, const char *suffix) {
    const size_t n = strlen(suffix);
    return s.size() >= n && s.compare(s.size() - n, n, suffix) == 0;
}

// rgb_dev: device f32 image.  Writes the encoded file into out.
int encode_from_device(int format, const float *rgb_dev, int w, int h, unsigned char *out, uint64_t capacity, uint64_t *bytes) {
    const uint64_t need = lj_image_encoded_size(format, w, h);
    if (need == 0) { lj_set_error("lj_image_encode: unknown format %d or empty image", format); return LJ_EINVAL; }
    if (capacity < need) { lj_set_error("lj_image_encode: buffer of %llu bytes, %llu needed", (unsigned long long)capacity, (unsigned long long)need); return LJ_EINVAL; }
    const size_t npix = (size_t)w * h;
    if (format == LJ_IMAGE_PFM) {
        const std::string hd = pfm_header(w, h);
        memcpy(out, hd.data(), hd.size());
        LJ_CUDA(cudaMemcpy(out + hd.size(), rgb_dev, npix * 3 * sizeof(float), cudaMemcpyDeviceToHost));
        *bytes = hd.size() + npix * 3 * sizeof(float);
        return LJ_OK;
    }
    const std::vector<unsigned char> hd = exr_header(w, h);
    const uint64_t line_bytes = 8 + 6ull * w;
    const uint64_t table = 8ull * h, body0 = hd.size() + table;
    memcpy(out, hd.data(), hd.size());
    for (int y = 0; y < h; y++) { const uint64_t off = body0 + line_bytes * y; memcpy(out + hd.size() + 8ull * y, &off, 8); }
    unsigned short *body = nullptr;
    LJ_CUDA(lj_malloc(&body, line_bytes * h));
    const long long words = (long long)(line_bytes / 2) * h;
    k_exr_lines<<<(unsigned)((words + 255) / 256), 256>>>(rgb_dev, w, h, body);
    cudaError_t e = cudaMemcpy(out + body0, body, line_bytes * h, cudaMemcpyDeviceToHost);
    lj_free(body);
    if (e != cudaSuccess) { lj_set_error("lj_image_encode: %s", cudaGetErrorString(e)); return LJ_ECUDA; }
    *bytes = body0 + line_bytes * h;
    return LJ_OK;
}

int write_file(const char *filename, const void *data, size_t n) {
    FILE *f = fopen(filename, "wb");
    if (!f) { lj_set_error("cannot open %s for writing", filename); return LJ_EIO; }
    const size_t wr = fwrite(data, 1, n, f);
    fclose(f);
    if (wr != n) { lj_set_error("short write to %s", filename); return LJ_EIO; }
    return LJ_OK;
}

int format_of(const char *filename, int *format) {
    const std::string s = filename ? filename : "";
    if (ends_with(s, ".pfm")) { *format = LJ_IMAGE_PFM; return LJ_OK; }
    if (ends_with(s, ".exr")) { *format = LJ_IMAGE_EXR; return LJ_OK; }
    lj_set_error("Unsupported image format: %s", s.c_str());           // image.cpp:130 wording
    return LJ_EINVAL;
}

struct AccumHeader { char magic[8]; int32_t width, height, sample_begin, sample_end; };
const char ACCUM_MAGIC[8] = {'L', 'J', 'A', 'C', 'C', 'U', 'M', '1'};

}  // namespace

extern "C" uint64_t lj_image_encoded_size(int32_t format, int32_t width, int32_t height) {
    if (width <= 0 || height <= 0) return 0;
    const uint64_t npix = (uint64_t)width * height;
    if (format == LJ_IMAGE_PFM) return pfm_header(width, height).size() + npix * 12;
    if (format == LJ_IMAGE_EXR) return exr_header(width, height).size() + 8ull * height + (8 + 6ull * width) * height;
    return 0;
}

extern "C" int lj_image_encode(lj_ctx *ctx, int32_t format, const float *rgb, int32_t width, int32_t height,
                               void *out, uint64_t capacity, uint64_t *bytes) {
    if (!ctx || !rgb || !out || !bytes) { lj_set_error("lj_image_encode: null argument"); return LJ_EINVAL; }
    if (width <= 0 || height <= 0) { lj_set_error("lj_image_encode: empty image"); return LJ_EINVAL; }
    LJ_CUDA(cudaSetDevice(ctx->device));
    const size_t n = (size_t)width * height * 3;
    float *dev = nullptr;
    LJ_CUDA(lj_malloc(&dev, n * sizeof(float)));
    cudaError_t e = cudaMemcpy(dev, rgb, n * sizeof(float), cudaMemcpyHostToDevice);
    int rc = LJ_OK;
    if (e != cudaSuccess) { lj_set_error("lj_image_encode: %s", cudaGetErrorString(e)); rc = LJ_ECUDA; }
    else rc = encode_from_device(format, dev, width, height, (unsigned char *)out, capacity, bytes);
    lj_free(dev);
    return rc;
}

extern "C" int lj_accum_encode(lj_ctx *ctx, int32_t format, const double *dev_sum, int32_t width, int32_t height, int32_t spp,
                               void *out, uint64_t capacity, uint64_t *bytes) {
    if (!ctx || !dev_sum || !out || !bytes) { lj_set_error("lj_accum_encode: null argument"); return LJ_EINVAL; }
    if (width <= 0 || height <= 0 || spp <= 0) { lj_set_error("lj_accum_encode: empty image or spp <= 0"); return LJ_EINVAL; }
    LJ_CUDA(cudaSetDevice(ctx->device));
    const long long n = (long long)width * height * 3;
    float *dev = nullptr;
    LJ_CUDA(lj_malloc(&dev, n * sizeof(float)));
    k_accum_to_f32<<<(unsigned)((n + 255) / 256), 256>>>(n, dev_sum, 1.0 / (double)spp, dev);
    const int rc = encode_from_device(format, dev, width, height, (unsigned char *)out, capacity, bytes);
    lj_free(dev);
    return rc;
}

extern "C" int lj_accum_finalize(lj_ctx *ctx, const double *dev_sum, int32_t width, int32_t height, int32_t spp, float *rgb_out) {
    if (!ctx || !dev_sum || !rgb_out) { lj_set_error("lj_accum_finalize: null argument"); return LJ_EINVAL; }
    if (width <= 0 || height <= 0 || spp <= 0) { lj_set_error("lj_accum_finalize: empty image or spp <= 0"); return LJ_EINVAL; }
    LJ_CUDA(cudaSetDevice(ctx->device));
    const long long n = (long long)width * height * 3;
    float *dev = nullptr;
    LJ_CUDA(lj_malloc(&dev, n * sizeof(float)));
    k_accum_to_f32<<<(unsigned)((n + 255) / 256), 256>>>(n, dev_sum, 1.0 / (double)spp, dev);
    cudaError_t e = cudaMemcpy(rgb_out, dev, n * sizeof(float), cudaMemcpyDeviceToHost);
    lj_free(dev);
    if (e != cudaSuccess) { lj_set_error("lj_accum_finalize: %s", cudaGetErrorString(e)); return LJ_ECUDA; }
    return LJ_OK;
}

// ---- allocation-free variant of lj_accum_finalize (no counterpart in the reference; see include/lajolla_gpu.h) ----
// The f32 staging image lives in a file-scope device array instead of a cudaMallocAsync block, so the call allocates
// nothing: usable under CUDA graph capture of the surrounding code, in serving loops and in harnesses that forbid
// cudaMalloc.  Every GPU that loads the library holds the array (LJ_STATIC_FINALIZE_MAX_PIXELS * 3 floats = 12 MiB)
// until the process exits.  In host code `g_finalize_scratch` is only the host-side stand-in of the symbol, never a
// device address: the address is looked up once per device with cudaGetSymbolAddress (that device current) and the
// kernel is given that pointer.
static __device__ __align__(256) float g_finalize_scratch[(size_t)LJ_STATIC_FINALIZE_MAX_PIXELS * 3];
enum { LJ_STATIC_MAX_DEVICES = 64 };
static float *g_finalize_scratch_ptr[LJ_STATIC_MAX_DEVICES];      // per device ordinal, null until looked up
static std::mutex g_finalize_scratch_mu;

static int finalize_scratch_for(int device, float **out) {
    if (device < 0 || device >= LJ_STATIC_MAX_DEVICES) { lj_set_error("lj_accum_finalize_static: device ordinal %d not supported (max %d)", device, LJ_STATIC_MAX_DEVICES - 1); return LJ_EUNSUPPORTED; }
    std::lock_guard<std::mutex> lock(g_finalize_scratch_mu);
    if (!g_finalize_scratch_ptr[device]) {
        float *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_finalize_scratch);     // the calling thread's current device = `device`
        if (e != cudaSuccess || !p) { lj_set_error("lj_accum_finalize_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e)); return LJ_ECUDA; }
        g_finalize_scratch_ptr[device] = p;
    }
    *out = g_finalize_scratch_ptr[device];
    return LJ_OK;
}

extern "C" int lj_accum_finalize_static(lj_ctx *ctx, const double *dev_sum, int32_t width, int32_t height, int32_t spp, float *rgb_out) {
    if (!ctx || !dev_sum || !rgb_out) { lj_set_error("lj_accum_finalize_static: null argument"); return LJ_EINVAL; }
    if (width <= 0 || height <= 0 || spp <= 0) { lj_set_error("lj_accum_finalize_static: empty image or spp <= 0"); return LJ_EINVAL; }
    const long long npix = (long long)width * height;
    if (npix > (long long)LJ_STATIC_FINALIZE_MAX_PIXELS) {
        lj_set_error("lj_accum_finalize_static: %d x %d = %lld pixels exceed the static scratch of %d pixels; use lj_accum_finalize",
                     width, height, npix, (int)LJ_STATIC_FINALIZE_MAX_PIXELS);
        return LJ_EINVAL;
    }
    LJ_CUDA(cudaSetDevice(ctx->device));
    float *dev = nullptr;
    if (int rc = finalize_scratch_for(ctx->device, &dev)) return rc;
    const long long n = npix * 3;
    // every value the copy below reads is written by this launch first: nothing has to be cleared between calls
    k_accum_to_f32<<<(unsigned)((n + 255) / 256), 256>>>(n, dev_sum, 1.0 / (double)spp, dev);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy(rgb_out, dev, n * sizeof(float), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { lj_set_error("lj_accum_finalize_static: %s", cudaGetErrorString(e)); return LJ_ECUDA; }
    return LJ_OK;
}

extern "C" int lj_image_write(lj_ctx *ctx, const char *filename, const float *rgb, int32_t width, int32_t height) {
    int format = 0;
    if (int rc = format_of(filename, &format)) return rc;
    std::vector<unsigned char> buf(lj_image_encoded_size(format, width, height));
    uint64_t n = 0;
    if (int rc = lj_image_encode(ctx, format, rgb, width, height, buf.data(), buf.size(), &n)) return rc;
    return write_file(filename, buf.data(), n);
}

extern "C" int lj_accum_write(lj_ctx *ctx, const char *filename, const double *dev_sum, int32_t width, int32_t height, int32_t spp) {
    int format = 0;
    if (int rc = format_of(filename, &format)) return rc;
    std::vector<unsigned char> buf(lj_image_encoded_size(format, width, height));
    uint64_t n = 0;
    if (int rc = lj_accum_encode(ctx, format, dev_sum, width, height, spp, buf.data(), buf.size(), &n)) return rc;
    return write_file(filename, buf.data(), n);
}

// ---- checkpoint of a progressive render: the f64 sums and the sample range [begin, end) they cover ----
extern "C" int lj_accum_save(lj_ctx *ctx, const char *filename, const double *dev_sum, int32_t width, int32_t height,
                             int32_t sample_begin, int32_t sample_end) {
    if (!ctx || !filename || !dev_sum) { lj_set_error("lj_accum_save: null argument"); return LJ_EINVAL; }
    if (width <= 0 || height <= 0 || sample_end < sample_begin) { lj_set_error("lj_accum_save: bad size or sample range"); return LJ_EINVAL; }
    LJ_CUDA(cudaSetDevice(ctx->device));
    const size_t n = (size_t)width * height * 3;
    std::vector<unsigned char> buf(sizeof(AccumHeader) + n * sizeof(double));
    AccumHeader hd; memcpy(hd.magic, ACCUM_MAGIC, 8);
    hd.width = width; hd.height = height; hd.sample_begin = sample_begin; hd.sample_end = sample_end;
    memcpy(buf.data(), &hd, sizeof hd);
    LJ_CUDA(cudaMemcpy(buf.data() + sizeof hd, dev_sum, n * sizeof(double), cudaMemcpyDeviceToHost));
    return write_file(filename, buf.data(), buf.size());
}

extern "C" int lj_accum_load(lj_ctx *ctx, const char *filename, double *dev_sum, int32_t width, int32_t height,
                             int32_t *sample_begin, int32_t *sample_end) {
    if (!ctx || !filename || !dev_sum) { lj_set_error("lj_accum_load: null argument"); return LJ_EINVAL; }
    LJ_CUDA(cudaSetDevice(ctx->device));
    FILE *f = fopen(filename, "rb");
    if (!f) { lj_set_error("cannot open %s", filename); return LJ_EIO; }
    AccumHeader hd;
    const size_t n = (size_t)(width > 0 ? width : 0) * (size_t)(height > 0 ? height : 0) * 3;
    std::vector<double> data(n);
    const bool ok = fread(&hd, sizeof hd, 1, f) == 1 && memcmp(hd.magic, ACCUM_MAGIC, 8) == 0 &&
                    hd.width == width && hd.height == height && fread(data.data(), sizeof(double), n, f) == n;
    fclose(f);
    if (!ok) { lj_set_error("%s is not a %dx%d accumulation checkpoint", filename, width, height); return LJ_EIO; }
    LJ_CUDA(cudaMemcpy(dev_sum, data.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    if (sample_begin) *sample_begin = hd.sample_begin;
    if (sample_end) *sample_end = hd.sample_end;
    return LJ_OK;
}
