// TEST INFRASTRUCTURE (oracle) -- not part of the product path.
//
// Oracle driver: our own main() around the UNMODIFIED reference code.  The
// reference's render.cpp is included as a translation unit (path_tracing.h /
// vol_path_tracing.h define non-inline functions and may live in one TU only,
// SURVEY.md appendix A.10), so this file can call path_tracing(),
// vol_path_tracing*() and render() directly.
//
//   ref_driver render <scene.xml> -o out.pfm [--spp N] [--mode A|B] [-t T]
//                     [--tiles i/n] [--stats stats.json]
//       mode A : the reference's own render() (render.cpp:155) -- one pcg32
//                stream per 16x16 tile, init_pcg32(ty*ntx+tx) (render.cpp:82).
//       mode B : same integrators, but one pcg32 stream per (pixel, sample):
//                init_pcg32(s*w*h + y*w + x).  This is the stream layout of
//                the GPU wavefront renderer; it lets the GPU be compared with
//                the reference path by path without serialising a tile.
//       --res W H overrides the film size only (camera matrices untouched); used to
//       mint small goldens.  The same override exists in lj_flatten.
//   ref_driver hits <scene.xml> -o hits.bin [--res W H]
//       ray-cast seam dump: per pixel a primary, a bounce and a shadow ray as f32 (what the
//       reference hands to Embree) with the results of the reference's intersect() / occluded()
//       (format: see cmd_hits).
//   ref_driver imwrite <in.f32img> <out.pfm|out.exr>
//       the reference's imwrite() (image.cpp:135-173) applied to a raw image
//       {int32 w, h; float rgb[h][w][3]} (widened to the Image3 doubles imwrite takes):
//       the golden generator for the output-encode step.
//   ref_driver mesh <file.obj|file.serialized> <shape_index> <out.bin> [16 doubles: to_world, row major]
//       the reference's parse_obj() (parse_obj.cpp:137-234) / load_serialized()
//       (load_serialized.cpp:174-256) applied to an asset file; dumps
//       {u64 nv, nt, n_normals, n_uvs; f64 to_world[16], inverse(to_world)[16];
//        f64 positions[nv][3], normals[..][3], uvs[..][2]; i32 indices[nt][3]}.
//   ref_driver vol <file.vol> <1|3> <out.bin>
//       the reference's load_volume_from_file<Real|Spectrum>() (volume.cpp:6-115); dumps
//       {i32 res[3], channels; f64 p_min[3], p_max[3], max_data[channels], data[nvox][channels]}.
//   ref_driver rng <n>
//       first n next_pcg32_real<double> draws of init_pcg32(7) (known-answer
//       vector for pcg.h:32-67).
#include "render.cpp"   // found via -I$(REF)/src : /root/reference/src/render.cpp

#include "parse_scene.h"
#include "parse_obj.h"
#include "load_serialized.h"
#include "volume.h"
#include "timer.h"
#include <chrono>
#include <unistd.h>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>

extern "C" void lj_shim_ray_counts(uint64_t *isect, uint64_t *occl);
extern "C" void lj_shim_reset_ray_counts();

typedef Spectrum (*SampleFn)(const Scene &, int, int, pcg32_state &);

static SampleFn pick_integrator(const Scene &scene, bool &is_vol) {
    is_vol = false;
    if (scene.options.integrator == Integrator::Path) return path_tracing;
    if (scene.options.integrator == Integrator::VolPath) {
        is_vol = true;
        switch (scene.options.vol_path_version) { // render.cpp:111-124
            case 1: return vol_path_tracing_1;
            case 2: return vol_path_tracing_2;
            case 3: return vol_path_tracing_3;
            case 4: return vol_path_tracing_4;
            case 5: return vol_path_tracing_5;
            default: return vol_path_tracing;
        }
    }
    return nullptr;
}

// Mode B: per-(pixel,sample) streams.  Accumulation order and the NaN filter
// follow render.cpp:88-95 / :135-146.
static Image3 render_mode_b(const Scene &scene, int s_begin, int s_end) {
    int w = scene.camera.width, h = scene.camera.height;
    Image3 img(w, h);
    bool is_vol; SampleFn f = pick_integrator(scene, is_vol);
    if (!f) { fprintf(stderr, "mode B needs a path/volpath integrator\n"); exit(2); }
    constexpr int tile_size = 16;
    int ntx = (w + tile_size - 1) / tile_size, nty = (h + tile_size - 1) / tile_size;
    parallel_for([&](const Vector2i &tile) {
        int x0 = tile[0] * tile_size, x1 = min(x0 + tile_size, w);
        int y0 = tile[1] * tile_size, y1 = min(y0 + tile_size, h);
        for (int y = y0; y < y1; y++) for (int x = x0; x < x1; x++) {
            Spectrum radiance = make_zero_spectrum();
            for (int s = s_begin; s < s_end; s++) {
                pcg32_state rng = init_pcg32((uint64_t)s * (uint64_t)(w * h) + (uint64_t)(y * w + x));
                Spectrum L = f(scene, x, y, rng);
                if (!is_vol || isfinite(L)) radiance += L;
            }
            img(x, y) = radiance / Real(s_end - s_begin);
        }
    }, Vector2i(ntx, nty));
    return img;
}

static int cmd_render(int argc, char **argv) {
    std::string scene_file, out = "out.pfm", mode = "A", stats_file;
    int spp = -1, threads = std::thread::hardware_concurrency();
    int s_begin = 0, s_end = -1, max_depth = -100, res_w = 0, res_h = 0;
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        if (a == "-o") out = argv[++i];
        else if (a == "--spp") spp = atoi(argv[++i]);
        else if (a == "--mode") mode = argv[++i];
        else if (a == "-t") threads = atoi(argv[++i]);
        else if (a == "--stats") stats_file = argv[++i];
        else if (a == "--sample-range") { s_begin = atoi(argv[++i]); s_end = atoi(argv[++i]); }
        else if (a == "--max-depth") max_depth = atoi(argv[++i]);
        else if (a == "--res") { res_w = atoi(argv[++i]); res_h = atoi(argv[++i]); }
        else scene_file = a;
    }
    char cwd[4096]; if (!getcwd(cwd, sizeof cwd)) return 2;
    auto abs_path = [&](const std::string &p) { return (p.size() && p[0] == '/') ? p : std::string(cwd) + "/" + p; };
    out = abs_path(out); if (!stats_file.empty()) stats_file = abs_path(stats_file);

    RTCDevice dev = rtcNewDevice(nullptr);
    parallel_init(threads);
    auto t0 = std::chrono::steady_clock::now();
    Scene scene = parse_scene(scene_file, dev); // chdir()s into the scene folder (parse_scene.cpp:1143)
    if (spp > 0) scene.options.samples_per_pixel = spp;
    if (max_depth != -100) scene.options.max_depth = max_depth;
    if (res_w > 0) { scene.camera.width = res_w; scene.camera.height = res_h; }  // test-size override (film size only)
    auto t1 = std::chrono::steady_clock::now();
    lj_shim_reset_ray_counts();
    Image3 img;
    if (mode == "A") img = render(scene);
    else {
        if (s_end < 0) s_end = scene.options.samples_per_pixel;
        img = render_mode_b(scene, s_begin, s_end);
    }
    auto t2 = std::chrono::steady_clock::now();
    uint64_t ni, no; lj_shim_ray_counts(&ni, &no);
    imwrite(out, img);
    double parse_s = std::chrono::duration<double>(t1 - t0).count();
    double render_s = std::chrono::duration<double>(t2 - t1).count();
    int nspp = (mode == "A") ? scene.options.samples_per_pixel : (s_end - s_begin);
    double samples = double(scene.camera.width) * scene.camera.height * nspp;
    char buf[1024];
    snprintf(buf, sizeof buf,
             "{\"scene\": \"%s\", \"mode\": \"%s\", \"width\": %d, \"height\": %d, \"spp\": %d, \"threads\": %d, "
             "\"parse_s\": %.6f, \"render_s\": %.6f, \"samples\": %.0f, \"intersect_rays\": %llu, \"occluded_rays\": %llu, "
             "\"msamples_per_s\": %.6f, \"mrays_per_s\": %.6f}",
             scene_file.c_str(), mode.c_str(), scene.camera.width, scene.camera.height, nspp, threads,
             parse_s, render_s, samples, (unsigned long long)ni, (unsigned long long)no,
             samples / render_s * 1e-6, double(ni + no) / render_s * 1e-6);
    printf("%s\n", buf);
    if (!stats_file.empty()) { std::ofstream f(stats_file); f << buf << "\n"; }
    parallel_cleanup();
    return 0;
}

// Seam dump (rtcIntersect1 / rtcOccluded1 as the reference reaches them): three ray sets per pixel, each stored as
// the f32 ray the reference hands to Embree (the casts of intersection.cpp:15-24 / :71-78) plus what came back
// through the reference's own intersect() / occluded():
//   primary : the pixel-centre camera ray (aux_render's ray, render.cpp:27)
//   bounce  : from the primary hit, a direction drawn from a per-pixel pcg32 stream in the hemisphere of the
//             geometric normal, tnear = eps, tfar = inf (the shape of path_tracing.h:237's bsdf ray): incoherent
//   shadow  : from the primary hit towards a light sample drawn like path_tracing.h:97-127 (sample_light,
//             sample_point_on_light, tnear = eps, tfar = (1-eps) dist; envmap: tfar = inf), through occluded()
// File: {int32 w, h, version = 3, n_sets = 3} then per set w*h records
//   {float org[3], tnear, dir[3], tfar;  float t, u, v; int32 geomID, primID; float Ng[3];  int32 valid, occluded}
// t, u, v, Ng are the raw RTCRayHit values of that intersect() call (read back from the shim, and cross-checked here
// against the PathVertex the reference built from them).  valid = 0: no ray for this pixel (the primary ray
// missed).  For the shadow set only `occluded` is meaningful.
extern "C" void lj_shim_last_rayhit(RTCRayHit *out);
struct SeamRec {
    float org[3], tnear, dir[3], tfar;
    float t, u, v; int geom, prim; float ng[3];
    int valid, occl;
};
static long g_seam_mismatch = 0;
static void seam_set_ray(SeamRec &r, const Ray &ray) {
    r.org[0] = (float)ray.org[0]; r.org[1] = (float)ray.org[1]; r.org[2] = (float)ray.org[2];
    r.dir[0] = (float)ray.dir[0]; r.dir[1] = (float)ray.dir[1]; r.dir[2] = (float)ray.dir[2];
    r.tnear = (float)ray.tnear; r.tfar = (float)ray.tfar; r.valid = 1;
}
static void seam_set_hit(SeamRec &r, const Ray &ray, const std::optional<PathVertex> &v) {
    if (!v) return;
    RTCRayHit rh; lj_shim_last_rayhit(&rh);
    r.t = rh.ray.tfar; r.u = rh.hit.u; r.v = rh.hit.v; r.geom = (int)rh.hit.geomID; r.prim = (int)rh.hit.primID;
    r.ng[0] = rh.hit.Ng_x; r.ng[1] = rh.hit.Ng_y; r.ng[2] = rh.hit.Ng_z;
    // the vertex the reference built must come from exactly these values (intersection.cpp:38-48)
    Vector3 p = ray.org + ray.dir * Real(rh.ray.tfar);
    if (r.geom != v->shape_id || r.prim != v->primitive_id || Real(r.u) != v->st[0] || Real(r.v) != v->st[1] ||
        p[0] != v->position[0] || p[1] != v->position[1] || p[2] != v->position[2]) g_seam_mismatch++;
}

static int cmd_hits(int argc, char **argv) {
    std::string scene_file, out = "hits.bin";
    int res_w = 0, res_h = 0;
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        if (a == "-o") out = argv[++i];
        else if (a == "--res") { res_w = atoi(argv[++i]); res_h = atoi(argv[++i]); }
        else scene_file = a;
    }
    char cwd[4096]; if (!getcwd(cwd, sizeof cwd)) return 2;
    if (out[0] != '/') out = std::string(cwd) + "/" + out;
    RTCDevice dev = rtcNewDevice(nullptr);
    parallel_init(1);
    Scene scene = parse_scene(scene_file, dev);
    if (res_w > 0) { scene.camera.width = res_w; scene.camera.height = res_h; }
    int w = scene.camera.width, h = scene.camera.height;
    SeamRec blank; memset(&blank, 0, sizeof blank); blank.geom = blank.prim = -1;
    std::vector<SeamRec> prim(w * h, blank), bounce(w * h, blank), shadow(w * h, blank);
    const Real eps = get_shadow_epsilon(scene);
    long n_hit = 0, n_bhit = 0, n_occl = 0, n_sh = 0;
    for (int y = 0; y < h; y++) for (int x = 0; x < w; x++) {
        int i = y * w + x;
        Ray ray = sample_primary(scene.camera, Vector2((x + Real(0.5)) / w, (y + Real(0.5)) / h));
        seam_set_ray(prim[i], ray);
        std::optional<PathVertex> v = intersect(scene, ray, init_ray_differential(w, h));
        seam_set_hit(prim[i], ray, v);
        if (!v) continue;
        n_hit++;
        pcg32_state rng = init_pcg32((uint64_t)i);
        // bounce: uniform direction on the sphere, flipped into the hemisphere of the geometric normal
        Real z = 1 - 2 * next_pcg32_real<Real>(rng), phi = 2 * c_PI * next_pcg32_real<Real>(rng);
        Real rr = sqrt(max(Real(0), 1 - z * z));
        Vector3 d{rr * cos(phi), rr * sin(phi), z};
        if (dot(d, v->geometry_normal) < 0) d = -d;
        Ray bray{v->position, d, get_intersection_epsilon(scene), infinity<Real>()};
        seam_set_ray(bounce[i], bray);
        std::optional<PathVertex> bv = intersect(scene, bray);
        seam_set_hit(bounce[i], bray, bv);
        if (bv) n_bhit++;
        // shadow: path_tracing.h:97-127
        if (scene.lights.empty()) continue;
        Vector2 light_uv{next_pcg32_real<Real>(rng), next_pcg32_real<Real>(rng)};
        Real light_w = next_pcg32_real<Real>(rng);
        Real shape_w = next_pcg32_real<Real>(rng);
        int light_id = sample_light(scene, light_w);
        const Light &light = scene.lights[light_id];
        PointAndNormal pl = sample_point_on_light(light, v->position, light_uv, shape_w, scene);
        Ray sray;
        if (!is_envmap(light)) {
            Vector3 dir_light = normalize(pl.position - v->position);
            sray = Ray{v->position, dir_light, eps, (1 - eps) * distance(pl.position, v->position)};
        } else {
            sray = Ray{v->position, -pl.normal, eps, infinity<Real>()};
        }
        seam_set_ray(shadow[i], sray);
        shadow[i].occl = occluded(scene, sray) ? 1 : 0;
        n_sh++; n_occl += shadow[i].occl;
    }
    std::ofstream f(out, std::ios::binary);
    int hdr[4] = {w, h, 3, 3};
    f.write((const char *)hdr, sizeof hdr);
    f.write((const char *)prim.data(), prim.size() * sizeof(SeamRec));
    f.write((const char *)bounce.data(), bounce.size() * sizeof(SeamRec));
    f.write((const char *)shadow.data(), shadow.size() * sizeof(SeamRec));
    printf("seam %d x %d: %ld primary hits, %ld bounce hits, %ld of %ld shadow rays occluded -> %s\n", w, h, n_hit, n_bhit, n_occl, n_sh, out.c_str());
    parallel_cleanup();
    if (g_seam_mismatch) { fprintf(stderr, "seam dump: %ld records disagree with the PathVertex of intersect()\n", g_seam_mismatch); return 3; }
    return 0;
}

static int cmd_imwrite(int argc, char **argv) {
    if (argc < 4) { fprintf(stderr, "usage: ref_driver imwrite in.f32img out.pfm|out.exr\n"); return 2; }
    std::ifstream f(argv[2], std::ios::binary);
    int w = 0, h = 0;
    f.read((char *)&w, 4); f.read((char *)&h, 4);
    if (!f || w <= 0 || h <= 0) { fprintf(stderr, "bad input %s\n", argv[2]); return 2; }
    std::vector<float> rgb((size_t)w * h * 3);
    f.read((char *)rgb.data(), rgb.size() * sizeof(float));
    Image3 img(w, h);
    for (int i = 0; i < w * h; i++) img(i) = Vector3{rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]};
    imwrite(argv[3], img);
    return 0;
}

static int cmd_mesh(int argc, char **argv) {
    if (argc < 5) { fprintf(stderr, "usage: ref_driver mesh file shape_index out.bin [16 doubles]\n"); return 2; }
    std::string file = argv[2];
    int shape_index = atoi(argv[3]);
    Matrix4x4 to_world = Matrix4x4::identity();
    if (argc >= 5 + 16) for (int i = 0; i < 16; i++) to_world(i / 4, i % 4) = strtod(argv[5 + i], nullptr);
    bool serialized = file.size() > 11 && file.substr(file.size() - 11) == ".serialized";
    TriangleMesh mesh = serialized ? load_serialized(file, shape_index, to_world) : parse_obj(file, to_world);
    Matrix4x4 inv = inverse(to_world);
    std::ofstream f(argv[4], std::ios::binary);
    uint64_t hdr[4] = {mesh.positions.size(), mesh.indices.size(), mesh.normals.size(), mesh.uvs.size()};
    f.write((const char *)hdr, sizeof hdr);
    for (int i = 0; i < 16; i++) { double v = to_world(i / 4, i % 4); f.write((const char *)&v, 8); }
    for (int i = 0; i < 16; i++) { double v = inv(i / 4, i % 4); f.write((const char *)&v, 8); }
    for (auto &p : mesh.positions) { double v[3] = {p[0], p[1], p[2]}; f.write((const char *)v, 24); }
    for (auto &p : mesh.normals) { double v[3] = {p[0], p[1], p[2]}; f.write((const char *)v, 24); }
    for (auto &p : mesh.uvs) { double v[2] = {p[0], p[1]}; f.write((const char *)v, 16); }
    for (auto &t : mesh.indices) { int v[3] = {t[0], t[1], t[2]}; f.write((const char *)v, 12); }
    printf("mesh %s: %zu vertices, %zu triangles, %zu normals, %zu uvs\n", file.c_str(), mesh.positions.size(),
           mesh.indices.size(), mesh.normals.size(), mesh.uvs.size());
    return 0;
}

static int cmd_vol(int argc, char **argv) {
    if (argc < 5) { fprintf(stderr, "usage: ref_driver vol file.vol 1|3 out.bin\n"); return 2; }
    int ch = atoi(argv[3]);
    std::ofstream f(argv[4], std::ios::binary);
    auto head = [&](const Vector3i &res, const Vector3 &a, const Vector3 &b) {
        int r[4] = {res[0], res[1], res[2], ch}; f.write((const char *)r, 16);
        double v[6] = {a[0], a[1], a[2], b[0], b[1], b[2]}; f.write((const char *)v, 48);
    };
    if (ch == 1) {
        GridVolume<Real> g = load_volume_from_file<Real>(argv[2]);
        head(g.resolution, g.p_min, g.p_max);
        f.write((const char *)&g.max_data, 8);
        f.write((const char *)g.data.data(), g.data.size() * 8);
        printf("vol %s: %d x %d x %d, max %.17g\n", argv[2], g.resolution[0], g.resolution[1], g.resolution[2], g.max_data);
    } else {
        GridVolume<Spectrum> g = load_volume_from_file<Spectrum>(argv[2]);
        head(g.resolution, g.p_min, g.p_max);
        double m[3] = {g.max_data[0], g.max_data[1], g.max_data[2]}; f.write((const char *)m, 24);
        for (auto &p : g.data) { double v[3] = {p[0], p[1], p[2]}; f.write((const char *)v, 24); }
        printf("vol %s: %d x %d x %d x 3, max %.17g %.17g %.17g\n", argv[2], g.resolution[0], g.resolution[1], g.resolution[2], m[0], m[1], m[2]);
    }
    return 0;
}

static int cmd_rng(int argc, char **argv) {
    int n = argc > 2 ? atoi(argv[2]) : 8;
    pcg32_state rng = init_pcg32(7);
    for (int i = 0; i < n; i++) printf("%.17g\n", next_pcg32_real<double>(rng));
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: ref_driver render|hits|imwrite|mesh|vol|rng ...\n"); return 2; }
    std::string cmd = argv[1];
    if (cmd == "render") return cmd_render(argc, argv);
    if (cmd == "hits") return cmd_hits(argc, argv);
    if (cmd == "imwrite") return cmd_imwrite(argc, argv);
    if (cmd == "mesh") return cmd_mesh(argc, argv);
    if (cmd == "vol") return cmd_vol(argc, argv);
    if (cmd == "rng") return cmd_rng(argc, argv);
    fprintf(stderr, "unknown command %s\n", cmd.c_str());
    return 2;
}
