// lj_flatten scene.xml out.ljs : parse a scene with the reference's own parser
// (src/parse_scene.cpp) and write the flattened lj_scene_desc blob, so GPU
// tests / bench.py can run where the XML assets are absent.
#include "parse_scene.h"
#include "scene_flatten.h"
#include "lajolla_gpu.h"
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <string>

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: lj_flatten scene.xml out.ljs [--res W H] [--spp N]\n"); return 2; }
    int res_w = 0, res_h = 0, spp = 0;
    for (int i = 3; i < argc; i++) {
        std::string a = argv[i];
        if (a == "--res" && i + 2 < argc) { res_w = atoi(argv[++i]); res_h = atoi(argv[++i]); }
        else if (a == "--spp" && i + 1 < argc) spp = atoi(argv[++i]);
    }
    char cwd[4096]; if (!getcwd(cwd, sizeof cwd)) return 2;
    std::string out = argv[2];
    if (out[0] != '/') out = std::string(cwd) + "/" + out;
    RTCDevice dev = rtcNewDevice(nullptr);
    Scene scene = parse_scene(argv[1], dev);
    if (res_w > 0) { scene.camera.width = res_w; scene.camera.height = res_h; }   // film size only, as ref_driver --res
    if (spp > 0) scene.options.samples_per_pixel = spp;
    FlatScene fs;
    flatten_scene(scene, fs);
    if (lj_scene_desc_save(&fs.desc, out.c_str()) != LJ_OK) { fprintf(stderr, "save failed: %s\n", lj_last_error()); return 1; }
    printf("%s: %llu shapes, %llu tris, %llu lights, %llu materials, %llu texels -> %s\n", argv[1],
           (unsigned long long)fs.desc.num_shapes, (unsigned long long)fs.desc.num_indices,
           (unsigned long long)fs.desc.num_lights, (unsigned long long)fs.desc.num_materials,
           (unsigned long long)fs.desc.num_texels, out.c_str());
    return 0;
}
