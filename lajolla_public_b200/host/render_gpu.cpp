// GPU drop-in for the reference's render driver:
//     Image3 render(const Scene &scene);      // src/render.h:9, src/render.cpp:155-170
// Linked (instead of src/render.cpp + src/intersection.cpp + Embree) with the
// reference's own main.cpp, parser and loaders, it keeps `./lajolla scene.xml`,
// the pugixml scene format, the Scene variant types and the image writer, and
// moves everything that runs per sample onto the B200 through the C ABI of
// include/lajolla_gpu.h.  There is no CPU fallback: any ABI failure is turned
// into the reference's Error() exception (flexception.h:24).
//
// Environment knobs (the reference has no config system beyond the XML):
//   LJ_RNG_MODE=A|B   A = per-tile pcg32 streams of render.cpp:82 (tile-serial, for
//                     path-level parity); B = per-(pixel,sample) streams (default).
//   LJ_ENGINE=wavefront|megakernel     (default wavefront)
//   LJ_DEVICE=<ordinal>[,<ordinal>...]  devices to render on (default: EVERY visible device -- the scene is
//                     replicated, the render is split by tiles, the sums are added with one NCCL reduce;
//                     the reference's counterpart is parallel_init(num_threads), main.cpp:31)
//   LJ_PARTITION=tiles|spp             multi-device split (default tiles: bit-identical to one device)
//   LJ_SPP=<n>                         override sampleCount
//   LJ_SAVE_SCENE=<file.ljs>           also write the flattened scene blob
//   LJ_STATS=1                         print lj_stats to stderr
#include "render.h"
#include "scene.h"
#include "flexception.h"
#include "scene_flatten.h"
#include "lajolla_gpu.h"
#include "host_ctx.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static lj_ctx *g_ctx = nullptr;

static void check(int rc, const char *what) {
    if (rc != LJ_OK) Error(std::string(what) + " failed: " + lj_last_error());
}

lj_ctx *lj_host_ctx() {
    if (!g_ctx) {
        std::vector<int> ids;
        if (const char *e = getenv("LJ_DEVICE")) {
            for (const char *p = e; *p;) {
                char *end; long v = strtol(p, &end, 10);
                if (end == p) break;
                ids.push_back((int)v);
                p = *end == ',' ? end + 1 : end;
            }
        }
        check(lj_init_devices(ids.empty() ? nullptr : ids.data(), (int)ids.size(), &g_ctx), "lj_init_devices");
    }
    return g_ctx;
}

Image3 render(const Scene &scene) {
    FlatScene fs;
    flatten_scene(scene, fs);
    if (const char *p = getenv("LJ_SAVE_SCENE")) check(lj_scene_desc_save(&fs.desc, p), "lj_scene_desc_save");

    lj_host_ctx();
    lj_scene *gs = nullptr;
    check(lj_scene_create(g_ctx, &fs.desc, &gs), "lj_scene_create");

    lj_render_params rp; memset(&rp, 0, sizeof rp);
    rp.max_depth_override = -100;
    rp.rng_mode = LJ_RNG_SAMPLE_STREAMS;
    if (const char *m = getenv("LJ_RNG_MODE")) rp.rng_mode = (m[0] == 'A' || m[0] == 'a') ? LJ_RNG_TILE_STREAMS : LJ_RNG_SAMPLE_STREAMS;
    if (const char *e = getenv("LJ_ENGINE")) rp.engine = (e[0] == 'm') ? LJ_ENGINE_MEGAKERNEL : LJ_ENGINE_WAVEFRONT;
    if (const char *s = getenv("LJ_SPP")) rp.spp = atoi(s);

    int w = scene.camera.width, h = scene.camera.height;
    std::vector<float> rgb((size_t)w * h * 3);
    lj_stats st; memset(&st, 0, sizeof st);
    int rc = lj_render(gs, &rp, rgb.data(), &st);
    if (rc != LJ_OK) { std::string msg = lj_last_error(); lj_scene_destroy(gs); Error("lj_render failed: " + msg); }
    if (getenv("LJ_STATS")) {
        uint64_t rays = st.rays_primary + st.rays_extension + st.rays_shadow;
        fprintf(stderr, "[lajolla_b200] %d device(s), render %.3f ms (reduce %.3f ms), bvh build %.3f ms, upload %.3f ms, %llu samples (%.2f Msamples/s), "
                        "%llu rays (%.2f Mrays/s), %llu launches\n",
                st.n_devices, st.render_ms, st.reduce_ms, st.bvh_build_ms, st.upload_ms, (unsigned long long)st.samples,
                st.samples / st.render_ms * 1e-3, (unsigned long long)rays, rays / st.render_ms * 1e-3,
                (unsigned long long)st.kernel_launches);
    }
    lj_scene_destroy(gs);

    Image3 img(w, h);
    for (int i = 0; i < w * h; i++) img(i) = Vector3{rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]};
    return img;
}
