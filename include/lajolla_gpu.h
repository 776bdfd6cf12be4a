/* lajolla_gpu.h -- C ABI of the B200-native render hot path.
 *
 * Drop-in boundary: the reference's
 *     Image3 render(const Scene &scene);            // src/render.h:9, called from src/main.cpp:40
 * and, below it, the Embree seam
 *     rtcIntersect1 / rtcOccluded1                  // src/intersection.cpp:32, :83
 * are replaced by the entry points below.  The reference has no FFI of its own
 * ("plugins" are std::variant alternatives), so the host shim
 * (lajolla_public_b200/host/render_gpu.cpp) flattens `Scene` into the POD
 * `lj_scene_desc` and calls lj_scene_create + lj_render.
 *
 * Conventions: every call returns 0 or a negative LJ_E* code; the message is in
 * lj_last_error() (thread local).  No C++ type, exception or torch type crosses
 * this boundary.  The caller owns every host buffer it passes in; the library
 * owns all device memory behind the opaque handles.  There is NO CPU fallback:
 * without a CUDA device lj_init fails with LJ_ENODEVICE.
 */
#ifndef LAJOLLA_GPU_H
#define LAJOLLA_GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LJ_OK 0
#define LJ_EINVAL (-1)
#define LJ_ENODEVICE (-2)
#define LJ_ECUDA (-3)
#define LJ_EIO (-4)
#define LJ_EUNSUPPORTED (-5)

/* ---- enumerations (values mirror the reference's variant order) ---- */
enum { LJ_TEX_CONSTANT = 0, LJ_TEX_IMAGE = 1, LJ_TEX_CHECKERBOARD = 2 };        /* texture.h:134 */
enum { LJ_SHAPE_SPHERE = 0, LJ_SHAPE_MESH = 1 };                                /* shape.h:54 */
enum { LJ_LIGHT_AREA = 0, LJ_LIGHT_ENVMAP = 1 };                                /* light.h:34 */
enum {                                                                          /* material.h:102-110 */
    LJ_MAT_LAMBERTIAN = 0, LJ_MAT_ROUGHPLASTIC = 1, LJ_MAT_ROUGHDIELECTRIC = 2,
    LJ_MAT_DISNEY_DIFFUSE = 3, LJ_MAT_DISNEY_METAL = 4, LJ_MAT_DISNEY_GLASS = 5,
    LJ_MAT_DISNEY_CLEARCOAT = 6, LJ_MAT_DISNEY_SHEEN = 7, LJ_MAT_DISNEY_BSDF = 8,
    LJ_MAT_KIND_COUNT = 9
};
enum { LJ_FILTER_BOX = 0, LJ_FILTER_TENT = 1, LJ_FILTER_GAUSSIAN = 2 };         /* filter.h:45 */
enum { LJ_MEDIUM_HOMOGENEOUS = 0, LJ_MEDIUM_HETEROGENEOUS = 1 };                /* medium.h:22 */
enum { LJ_PHASE_ISOTROPIC = 0, LJ_PHASE_HG = 1 };                               /* phase_function.h:16 */
enum { LJ_VOLUME_CONSTANT = 0, LJ_VOLUME_GRID = 1 };                            /* volume.h:27 */
enum {                                                                          /* scene.h:14-22 */
    LJ_INTEGRATOR_DEPTH = 0, LJ_INTEGRATOR_SHADING_NORMAL = 1, LJ_INTEGRATOR_MEAN_CURVATURE = 2,
    LJ_INTEGRATOR_RAY_DIFFERENTIAL = 3, LJ_INTEGRATOR_MIPMAP_LEVEL = 4,
    LJ_INTEGRATOR_PATH = 5, LJ_INTEGRATOR_VOLPATH = 6
};

/* Material texture slots (index into lj_material.tex[]). */
enum {
    /* Lambertian */           LJ_SLOT_REFLECTANCE = 0,
    /* RoughPlastic */         LJ_SLOT_DIFFUSE = 0, LJ_SLOT_SPECULAR = 1, LJ_SLOT_ROUGHNESS = 2,
    /* RoughDielectric */      LJ_SLOT_SPEC_REFL = 0, LJ_SLOT_SPEC_TRANS = 1, /* roughness = 2 */
    /* Disney* */              LJ_SLOT_BASE_COLOR = 0, LJ_SLOT_SPEC_TRANSMISSION = 1, /* roughness = 2 */
    LJ_SLOT_METALLIC = 3, LJ_SLOT_SUBSURFACE = 4, LJ_SLOT_SPECULAR_D = 5, LJ_SLOT_SPECULAR_TINT = 6,
    LJ_SLOT_ANISOTROPIC = 7, LJ_SLOT_SHEEN = 8, LJ_SLOT_SHEEN_TINT = 9, LJ_SLOT_CLEARCOAT = 10,
    LJ_SLOT_CLEARCOAT_GLOSS = 11, LJ_SLOT_COUNT = 12
};

/* ---- POD scene description (host memory, caller owned) ---- */
typedef struct lj_texture {        /* texture.h:75-92; scalar textures use value[0]/color1[0] */
    int32_t kind, image_id;        /* image_id indexes images1 (scalar) or images3 (spectrum) */
    int32_t channels, _pad;        /* 1 or 3 */
    double value[3];               /* constant value, or checkerboard color0 */
    double color1[3];              /* checkerboard color1 */
    double uscale, vscale, uoffset, voffset;
} lj_texture;

typedef struct lj_image {          /* mipmap.h:8-48: up to 8 box-filtered levels, f64 texels */
    int32_t channels, num_levels;
    int32_t width[8], height[8];
    uint64_t offset[8];            /* start of each level in `texels`, in doubles */
} lj_image;

typedef struct lj_shape {          /* shape.h:26-50 */
    int32_t kind, material_id, area_light_id, interior_medium_id, exterior_medium_id;
    int32_t has_normals, has_uvs, _pad;
    double center[3], radius;                  /* sphere */
    uint32_t vertex_offset, num_vertices;      /* mesh: ranges in positions/normals/uvs */
    uint32_t triangle_offset, num_triangles;   /* mesh: range in indices */
    double total_area;                         /* mesh, shapes/triangle_mesh.inl:48-63 */
    uint64_t tri_cdf_offset, tri_pmf_offset;   /* TableDist1D of the mesh, in `tables` (doubles) */
} lj_shape;

typedef struct lj_light {          /* light.h:15-30 */
    int32_t kind, shape_id;
    double intensity[3];
    int32_t values_texture, _pad;  /* envmap: index into textures */
    double to_world[16], to_local[16], scale;  /* row-major 4x4 */
    int32_t dist_width, dist_height;           /* TableDist2D, table_dist.h:28-37 */
    uint64_t cdf_rows_offset, pdf_rows_offset, cdf_marginals_offset, pdf_marginals_offset;
    double total_values;
} lj_light;

typedef struct lj_material {       /* material.h:11-100 */
    int32_t kind, _pad;
    double eta;
    int32_t tex[LJ_SLOT_COUNT];    /* indices into textures, -1 if unused */
} lj_material;

typedef struct lj_volume {         /* volume.h:13-27 */
    int32_t kind, _pad;
    double value[3];
    int32_t resolution[3], _pad2;
    double p_min[3], p_max[3], max_data[3], scale;
    uint64_t data_offset;          /* in `grid_data`, 3 doubles per voxel */
} lj_volume;

typedef struct lj_medium {         /* medium.h:10-22 */
    int32_t kind, phase_kind;
    double g;
    double sigma_a[3], sigma_s[3];
    lj_volume albedo, density;
} lj_medium;

typedef struct lj_camera {         /* camera.h:10-25 */
    double sample_to_cam[16], cam_to_world[16];
    int32_t width, height, filter_kind, medium_id;
    double filter_param;           /* Box/Tent width, Gaussian stddev */
} lj_camera;

typedef struct lj_options {        /* scene.h:24-31 */
    int32_t integrator, samples_per_pixel, max_depth, rr_depth, vol_path_version, max_null_collisions;
} lj_options;

typedef struct lj_scene_desc {
    lj_camera camera;
    lj_options options;
    double bounds_radius;          /* scene.cpp:30-34, from the f32 AABB */
    double bounds_center[3];
    int32_t envmap_light_id, _pad;

    uint64_t num_shapes, num_lights, num_materials, num_media, num_textures, num_images1, num_images3;
    const lj_shape *shapes;
    const lj_light *lights;
    const lj_material *materials;
    const lj_medium *media;
    const lj_texture *textures;
    const lj_image *images1, *images3;

    uint64_t num_positions, num_indices, num_texels, num_tables, num_grid_data;
    const double *positions;       /* 3 per vertex, all meshes concatenated */
    const double *normals;         /* 3 per vertex (zero for meshes without normals) */
    const double *uvs;             /* 2 per vertex (zero for meshes without uvs) */
    const int32_t *indices;        /* 3 per triangle, mesh-local vertex indices */
    const double *texels;          /* all mip levels of all images */
    const double *tables;          /* all pmf/cdf tables */
    const double *grid_data;       /* all volume grids */

    uint64_t light_pmf_offset, light_cdf_offset;   /* scene.light_dist in `tables` */
} lj_scene_desc;

/* ---- render parameters / statistics ---- */
enum { LJ_RNG_TILE_STREAMS = 0,    /* mode A: init_pcg32(ty*ntx+tx), render.cpp:82, tile-serial */
       LJ_RNG_SAMPLE_STREAMS = 1 };/* mode B: init_pcg32(s*w*h + y*w + x), fully parallel */
enum { LJ_ENGINE_WAVEFRONT = 0, LJ_ENGINE_MEGAKERNEL = 1 };

typedef struct lj_render_params {
    int32_t spp;                   /* 0 => scene options */
    int32_t rng_mode;              /* LJ_RNG_* */
    int32_t engine;                /* LJ_ENGINE_*; mode A always runs tile-serial */
    int32_t sample_begin, sample_end;   /* spp slice [begin,end) for mode B; 0,0 => [0,spp) */
    int32_t tile_rank, tile_world;      /* this call renders the 16x16 tiles dealt to `rank` of `world` (tile (tx, ty) belongs to rank
                                           (ty*ntx + (tx + 3 ty) % ntx) % world: row-rotated round robin); world 0 => all tiles */
    int32_t max_depth_override;    /* -100 => scene options */
    int32_t paths_in_flight;       /* wavefront pool size, 0 => auto */
    int32_t flags;                 /* bit0: count traversal steps (nodes / prims per ray); bit1: per-stage CUDA-event times */
    int32_t partition;             /* LJ_PARTITION_*: how a multi-device context (lj_init_devices) splits one render */
} lj_render_params;

/* Work split of ONE render over the devices of a multi-device context; the scene is replicated, every
 * device renders its share into its own f64 sums, and the sums are added with one NCCL reduce to the first device.
 *   TILES  the 16x16 tiles of render.cpp:75-82 dealt round-robin, each tile row rotated against the one above (see tile_rank): disjoint pixels, so the reduced
 *          image is bit-identical to the one-device image for every device count and every integrator / RNG mode;
 *   SPP    device i renders a contiguous slice of the samples of EVERY pixel (balanced when cost is spatially skewed);
 *          needs per-sample streams (LJ_RNG_SAMPLE_STREAMS) and a sampling integrator, otherwise TILES is used.
 * AUTO = TILES.  Ignored by one-device contexts and when the caller sets tile_world / a sample slice itself. */
enum { LJ_PARTITION_AUTO = 0, LJ_PARTITION_TILES = 1, LJ_PARTITION_SPP = 2 };

typedef struct lj_stats {
    double render_ms;              /* CUDA-event time of the whole render on the device */
    double upload_ms, bvh_build_ms;
    uint64_t samples;
    uint64_t rays_primary, rays_extension, rays_shadow;
    uint64_t nodes_visited, prims_tested;   /* only with flags bit0: closest-hit rays */
    uint64_t kernel_launches;
    double extend_ms, shadow_ms, shade_ms, generate_ms;   /* per-stage CUDA-event sums (flags bit1) */
    uint64_t bvh_nodes, bvh_prims;
    uint64_t shadow_nodes_visited, shadow_prims_tested;   /* flags bit0, any-hit rays */
    uint64_t iterations;                  /* wavefront iterations */
    double tables_build_ms;               /* lj_scene_create_ex with LJ_SCENE_BUILD_TABLES: device time of the scene build step */
    /* multi-device contexts: render_ms = slowest device's render + the reduce; the counters above are sums over devices */
    int32_t n_devices, partition;         /* devices that rendered; LJ_PARTITION_* actually used */
    double reduce_ms;                     /* NCCL reduce of the f64 sums, CUDA-event time on the root device */
    double device_ms_min, device_ms_max;  /* fastest / slowest device's render (CUDA events on each device) */
} lj_stats;

typedef struct lj_ray { float org[3], tnear, dir[3], tfar; } lj_ray;           /* RTCRay subset, f32 seam */
typedef struct lj_hit { float t, u, v; int32_t geom_id, prim_id; float ng[3]; } lj_hit;

typedef struct lj_ctx lj_ctx;
typedef struct lj_scene lj_scene;

/* ---- entry points ---- */
int lj_init(int device_id, lj_ctx **out);                      /* replaces rtcNewDevice + parallel_init (main.cpp:30-31) */
/* The same for several devices of one box: replaces parallel_init(num_threads) + the tile
 * parallel_for of render.cpp:75-98 as the unit that fans one render out.  device_ids NULL / n_devices <= 0 = every
 * visible device.  lj_scene_create on such a context uploads the scene and builds the BVH on every device (one host
 * thread per device); lj_render / lj_render_device split the work (lj_render_params.partition), and add the per-device
 * sums with ncclReduce (NCCL is loaded with dlopen("libnccl.so.2") the first time it is needed; without it a
 * multi-device render fails with LJ_EUNSUPPORTED -- there is no host-side fallback for the reduction).  Every other
 * entry point that takes the context or one of its scenes uses the first device. */
int lj_init_devices(const int *device_ids, int n_devices, lj_ctx **out);
int lj_ctx_device_count(lj_ctx *ctx);                          /* devices behind a context (1 for lj_init) */
void lj_shutdown(lj_ctx *ctx);                                 /* main.cpp:47-48 */
int lj_scene_create(lj_ctx *ctx, const lj_scene_desc *desc, lj_scene **out);   /* Scene::Scene's Embree build, scene.cpp:20-34 */
/* Scene build step on the device (scene.cpp:36-52 minus the Embree build).  With
 * LJ_SCENE_BUILD_TABLES the derived data of `desc` -- mip levels >= 1 (make_mipmap, mipmap.h:25-48), per-mesh
 * area tables and total_area (triangle_mesh.inl:48-63, table_dist.cpp:3-25), the envmap TableDist2D and
 * total_values (envmap.inl:75-98, table_dist.cpp:40-114) and the light distribution (scene.cpp:46-51) -- are NOT
 * read: their ranges in `texels` / `tables` only reserve space, and kernels derive them from the level-0 texels,
 * positions / indices and light records.  Bit-identical to the host-built values, except that the envmap rows go
 * through the device's sin().  LJ_SCENE_POISON_DERIVED (tests) overwrites those ranges with NaN first.
 * lj_scene_create reads LJ_BUILD_TABLES=1 from the environment as the same switch. */
#define LJ_SCENE_BUILD_TABLES 1u
#define LJ_SCENE_POISON_DERIVED 2u
int lj_scene_create_ex(lj_ctx *ctx, const lj_scene_desc *desc, uint32_t flags, lj_scene **out);
/* reads the device copies of the derived arrays back (verification of the step above) */
int lj_scene_read_derived(lj_scene *scene, double *tables, uint64_t num_tables, double *texels, uint64_t num_texels,
                          lj_shape *shapes, uint64_t num_shapes, lj_light *lights, uint64_t num_lights);
void lj_scene_destroy(lj_scene *scene);                        /* scene.cpp:67-71 */
/* render(): rgb_out is w*h*3 f32, row-major, y=0 top (what imwrite keeps, image.cpp:146-149) */
int lj_render(lj_scene *scene, const lj_render_params *params, float *rgb_out, lj_stats *stats);
/* same, but leaves per-pixel radiance SUMS (f64, w*h*3) in caller-provided DEVICE memory
 * (e.g. a torch tensor that is then all-reduced over NCCL); no division by spp. */
int lj_render_device(lj_scene *scene, const lj_render_params *params, double *dev_sum, lj_stats *stats);
/* rtcIntersect1 / rtcOccluded1 for a batch of f32 rays (host buffers) */
int lj_intersect_batch(lj_scene *scene, const lj_ray *rays, int64_t n, lj_hit *hits, int any_hit);
int lj_scene_info(lj_scene *scene, int32_t *width, int32_t *height, int32_t *spp, lj_stats *build_stats);
const char *lj_last_error(void);

/* flattened-scene blobs (so tests / bench can run where the XML assets are absent) */
int lj_scene_desc_save(const lj_scene_desc *desc, const char *path);
int lj_scene_desc_load(const char *path, lj_scene_desc **out);    /* allocates; free with lj_scene_desc_free */
void lj_scene_desc_free(lj_scene_desc *desc);
/* page-lock (pin != 0) or unlock the arrays of a loaded desc so uploads run from pinned host memory */
int lj_scene_desc_pin(lj_scene_desc *desc, int pin);

/* ---- output step: imwrite(), image.cpp:135-173 ----
 * Encodes an f32 RGB image (rows top to bottom, as lj_render returns it) the way the reference writes it.
 *   LJ_IMAGE_PFM: "PF\n<w> <h>\n-1\n" + the f32 rows top to bottom (image.cpp:141-149), byte-identical.
 *   LJ_IMAGE_EXR: scanline OpenEXR, channels B G R as HALF with tinyexr's float_to_half_full rounding
 *                 (tinyexr.h:889-924, reached through SaveEXR(..., fp16) at image.cpp:161-166); lines are stored
 *                 uncompressed where the reference deflates 16-line blocks: same decoded pixels.
 * The conversion and the scanline layout run on the device. */
enum { LJ_IMAGE_PFM = 0, LJ_IMAGE_EXR = 1 };
uint64_t lj_image_encoded_size(int32_t format, int32_t width, int32_t height);      /* exact size; 0 = bad arguments */
int lj_image_encode(lj_ctx *ctx, int32_t format, const float *rgb, int32_t width, int32_t height,
                    void *out, uint64_t capacity, uint64_t *bytes);
/* format from the extension (.pfm / .exr) like imwrite; anything else is LJ_EINVAL ("Unsupported image format") */
int lj_image_write(lj_ctx *ctx, const char *filename, const float *rgb, int32_t width, int32_t height);

/* The same from the f64 radiance sums lj_render_device leaves on the device (sum / spp, render.cpp:94, then
 * narrowed to f32, image.cpp:146), without a host round trip of the sums; and the progressive-render
 * checkpoint: the sums plus the sample range [sample_begin, sample_end) they cover, to be resumed with
 * lj_render_params.sample_begin / sample_end. */
int lj_accum_finalize(lj_ctx *ctx, const double *dev_sum, int32_t width, int32_t height, int32_t spp, float *rgb_out);
/* lj_accum_finalize without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  The f32 staging image is a file-scope `static __device__` array of
 * LJ_STATIC_FINALIZE_MAX_PIXELS * 3 floats (12 MiB, held by every GPU that loads the library until the process
 * exits); larger images are refused with LJ_EINVAL.  It calls no cudaMalloc, cudaFree or other allocating routine; same result
 * as lj_accum_finalize bit for bit.  All calls on a device share the array (it keeps the last call's contents; every
 * value a call reads it writes first), so calls that overlap on the GPU race: serialise them. */
#define LJ_STATIC_FINALIZE_MAX_PIXELS (1 << 20)
int lj_accum_finalize_static(lj_ctx *ctx, const double *dev_sum, int32_t width, int32_t height, int32_t spp, float *rgb_out);
int lj_accum_encode(lj_ctx *ctx, int32_t format, const double *dev_sum, int32_t width, int32_t height, int32_t spp,
                    void *out, uint64_t capacity, uint64_t *bytes);
int lj_accum_write(lj_ctx *ctx, const char *filename, const double *dev_sum, int32_t width, int32_t height, int32_t spp);
int lj_accum_save(lj_ctx *ctx, const char *filename, const double *dev_sum, int32_t width, int32_t height,
                  int32_t sample_begin, int32_t sample_end);
int lj_accum_load(lj_ctx *ctx, const char *filename, double *dev_sum, int32_t width, int32_t height,
                  int32_t *sample_begin, int32_t *sample_end);

/* ---- on-disk asset formats decoded on the device ----
 * The loaders the reference's parser calls for every <shape type="obj|serialized"> and <volume type="gridvolume">
 * (parse_scene.cpp), taking the file's bytes and returning what the reference's TriangleMesh / GridVolume hold:
 *   lj_mesh_load_obj         parse_obj(filename, to_world)                    parse_obj.cpp:137-234
 *   lj_mesh_load_serialized  load_serialized(filename, shape_index, to_world) load_serialized.cpp:174-256
 *   lj_mesh_compute_normals  compute_normal(vertices, indices)                parse_obj.cpp:57-92
 *   lj_volume_load           load_volume_from_file<Real|Spectrum>(filename)   volume.cpp:6-115
 * `to_world` / `to_local` are row-major 4x4 (to_local = the reference's inverse(to_world), matrix.h:80).
 * Output arrays are host memory owned by the library until lj_mesh_free / lj_volume_free.  Results are bit-identical
 * to the reference's loaders, except normals produced by compute_normal (the device's asin(), last ulp).  OBJ numbers
 * outside the exactly convertible range (see csrc/lj_assets.cu) are refused with LJ_EUNSUPPORTED, n-gons with
 * LJ_EINVAL (the reference's Error()). */
typedef struct lj_mesh_data {      /* TriangleMesh, shape.h:39-49 */
    uint64_t num_vertices, num_triangles;
    int32_t has_normals, has_uvs;
    double *positions;             /* 3 per vertex */
    double *normals;               /* 3 per vertex, NULL without normals */
    double *uvs;                   /* 2 per vertex, NULL without uvs */
    int32_t *indices;              /* 3 per triangle */
    double decode_ms;              /* device time of the decode kernels */
} lj_mesh_data;
typedef struct lj_volume_data {    /* GridVolume<Real|Spectrum>, volume.h:19-26 (scale = 1) */
    int32_t resolution[3], channels;
    double p_min[3], p_max[3], max_data[3];
    double *data;                  /* channels doubles per voxel, x fastest */
    uint64_t num_values;
    double decode_ms;
} lj_volume_data;
int lj_mesh_load_obj(lj_ctx *ctx, const void *text, uint64_t text_bytes, const double *to_world, const double *to_local,
                     lj_mesh_data *out);
int lj_mesh_load_serialized(lj_ctx *ctx, const void *file, uint64_t file_bytes, int32_t shape_index,
                            const double *to_world, const double *to_local, lj_mesh_data *out);
int lj_mesh_compute_normals(lj_ctx *ctx, const double *positions, uint64_t num_vertices, const int32_t *indices,
                            uint64_t num_triangles, double *normals_out);
void lj_mesh_free(lj_mesh_data *mesh);
int lj_volume_load(lj_ctx *ctx, const void *file, uint64_t file_bytes, int32_t target_channels, lj_volume_data *out);
void lj_volume_free(lj_volume_data *vol);

#ifdef __cplusplus
}
#endif
#endif
