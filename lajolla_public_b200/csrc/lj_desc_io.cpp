// Flattened-scene blobs: lj_scene_desc <-> one binary file ("LJS1").
// Layout: magic, the lj_scene_desc struct (pointers ignored), then every array
// in a fixed order, each as {uint64 byte_count, bytes}.
#include "../../include/lajolla_gpu.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>

void lj_set_error(const char *fmt, ...);

namespace {
struct Arr { const void *ptr; uint64_t bytes; };
void arrays_of(const lj_scene_desc *d, Arr *a) {
    a[0] = {d->shapes, d->num_shapes * sizeof(lj_shape)};
    a[1] = {d->lights, d->num_lights * sizeof(lj_light)};
    a[2] = {d->materials, d->num_materials * sizeof(lj_material)};
    a[3] = {d->media, d->num_media * sizeof(lj_medium)};
    a[4] = {d->textures, d->num_textures * sizeof(lj_texture)};
    a[5] = {d->images1, d->num_images1 * sizeof(lj_image)};
    a[6] = {d->images3, d->num_images3 * sizeof(lj_image)};
    a[7] = {d->positions, d->num_positions * 3 * sizeof(double)};
    a[8] = {d->normals, d->num_positions * 3 * sizeof(double)};
    a[9] = {d->uvs, d->num_positions * 2 * sizeof(double)};
    a[10] = {d->indices, d->num_indices * 3 * sizeof(int32_t)};
    a[11] = {d->texels, d->num_texels * sizeof(double)};
    a[12] = {d->tables, d->num_tables * sizeof(double)};
    a[13] = {d->grid_data, d->num_grid_data * sizeof(double)};
}
const int NARR = 14;
} // namespace

extern "C" int lj_scene_desc_save(const lj_scene_desc *d, const char *path) {
    if (!d || !path) { lj_set_error("lj_scene_desc_save: null argument"); return LJ_EINVAL; }
    FILE *f = fopen(path, "wb");
    if (!f) { lj_set_error("lj_scene_desc_save: cannot open %s", path); return LJ_EIO; }
    Arr a[NARR]; arrays_of(d, a);
    lj_scene_desc hdr = *d;      // host pointers mean nothing in a file: store them as null so blobs are reproducible
    hdr.shapes = nullptr; hdr.lights = nullptr; hdr.materials = nullptr; hdr.media = nullptr; hdr.textures = nullptr;
    hdr.images1 = nullptr; hdr.images3 = nullptr; hdr.positions = nullptr; hdr.normals = nullptr; hdr.uvs = nullptr;
    hdr.indices = nullptr; hdr.texels = nullptr; hdr.tables = nullptr; hdr.grid_data = nullptr;
    bool ok = fwrite("LJS1", 1, 4, f) == 4 && fwrite(&hdr, sizeof hdr, 1, f) == 1;
    for (int i = 0; ok && i < NARR; i++) {
        ok = fwrite(&a[i].bytes, 8, 1, f) == 1 && (a[i].bytes == 0 || fwrite(a[i].ptr, 1, a[i].bytes, f) == a[i].bytes);
    }
    fclose(f);
    if (!ok) { lj_set_error("lj_scene_desc_save: short write to %s", path); return LJ_EIO; }
    return LJ_OK;
}

extern "C" int lj_scene_desc_load(const char *path, lj_scene_desc **out) {
    if (!path || !out) { lj_set_error("lj_scene_desc_load: null argument"); return LJ_EINVAL; }
    FILE *f = fopen(path, "rb");
    if (!f) { lj_set_error("lj_scene_desc_load: cannot open %s", path); return LJ_EIO; }
    char magic[4];
    lj_scene_desc *d = (lj_scene_desc *)calloc(1, sizeof(lj_scene_desc));
    void *bufs[NARR] = {0};
    bool ok = fread(magic, 1, 4, f) == 4 && memcmp(magic, "LJS1", 4) == 0 && fread(d, sizeof *d, 1, f) == 1;
    Arr expect[NARR];
    if (ok) {   // counts come from the file: keep count * record size far from wrapping around 2^64
        const uint64_t counts[] = {d->num_shapes, d->num_lights, d->num_materials, d->num_media, d->num_textures, d->num_images1,
                                   d->num_images3, d->num_positions, d->num_indices, d->num_texels, d->num_tables, d->num_grid_data};
        for (uint64_t c : counts) ok = ok && c < (1ull << 40);
    }
    if (ok) arrays_of(d, expect);
    for (int i = 0; ok && i < NARR; i++) {
        uint64_t bytes = 0;
        ok = fread(&bytes, 8, 1, f) == 1 && bytes == expect[i].bytes;
        if (ok) {
            bufs[i] = malloc(bytes ? bytes : 1);
            ok = bufs[i] && (bytes == 0 || fread(bufs[i], 1, bytes, f) == bytes);
        }
    }
    fclose(f);
    if (!ok) {
        for (void *b : bufs) free(b);
        free(d);
        lj_set_error("lj_scene_desc_load: %s is not a valid LJS1 blob", path);
        return LJ_EIO;
    }
    d->shapes = (const lj_shape *)bufs[0]; d->lights = (const lj_light *)bufs[1];
    d->materials = (const lj_material *)bufs[2]; d->media = (const lj_medium *)bufs[3];
    d->textures = (const lj_texture *)bufs[4]; d->images1 = (const lj_image *)bufs[5]; d->images3 = (const lj_image *)bufs[6];
    d->positions = (const double *)bufs[7]; d->normals = (const double *)bufs[8]; d->uvs = (const double *)bufs[9];
    d->indices = (const int32_t *)bufs[10]; d->texels = (const double *)bufs[11];
    d->tables = (const double *)bufs[12]; d->grid_data = (const double *)bufs[13];
    *out = d;
    return LJ_OK;
}

extern "C" void lj_scene_desc_free(lj_scene_desc *d) {
    if (!d) return;
    free((void *)d->shapes); free((void *)d->lights); free((void *)d->materials); free((void *)d->media);
    free((void *)d->textures); free((void *)d->images1); free((void *)d->images3);
    free((void *)d->positions); free((void *)d->normals); free((void *)d->uvs); free((void *)d->indices);
    free((void *)d->texels); free((void *)d->tables); free((void *)d->grid_data);
    free(d);
}
