// Device-resident scene: the lj_scene_desc arrays copied to HBM plus the GPU-built BVH.
#pragma once
#include "lj_math.cuh"
#include "../../include/lajolla_gpu.h"

// ---- BVH ----
// Binary LBVH node (build intermediate and the simple traversal path), 32 bytes.
struct __align__(16) Bvh2Node {
    float lo[3]; int left;     // left  >= 0: inner child index; < 0: leaf, prim = ~left
    float hi[3]; int right;
};

// Compressed 8-wide node, 80 bytes = five 16-byte loads (layout after Ylitie, Karras,
// Laine 2017 "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs").
struct __align__(16) Bvh8Node {
    float px, py, pz;          // quantisation origin
    uint8_t ex, ey, ez, imask; // per-axis exponent, bit i set => child i is an inner node
    uint32_t child_base;       // index of first inner child (children are contiguous)
    uint32_t prim_base;        // index of first primitive referenced by this node
    uint8_t meta[8];           // inner: (1<<5)|(24+slot) ; leaf: (unary count << 5) | offset ; empty: 0
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(Bvh8Node) == 80, "Bvh8Node must be 80 bytes");

// A sphere as a BVH primitive (shapes/sphere.inl: one user primitive per sphere)
struct SpherePrim { double cx, cy, cz, r; int geom; int _pad; };

struct DScene {
    lj_camera camera;
    lj_options options;
    double bounds_radius;
    double eps;                 // get_shadow_epsilon == get_intersection_epsilon, scene.h:99-105
    int envmap_light_id;
    int num_lights, num_shapes, num_materials, num_media;

    const lj_shape *shapes;
    const lj_light *lights;
    const lj_material *materials;
    const lj_medium *media;
    const lj_texture *textures;
    const lj_image *images1, *images3;
    const double *positions, *normals, *uvs;
    const int *indices;
    const double *texels, *tables, *grid_data;
    // Volume grids at their native width (volume.cpp:6-104 reads f32 .vol files and widens): a grid whose values are
    // all exactly representable as f32 is kept as f32 on the device, with one channel when the three of the Spectrum
    // grid are equal (smoke.vol: 19.7 MB -> 3.3 MB).  Widening back in registers is exact, so lookups are
    // bit-identical.  The DEVICE copy of lj_volume marks the storage in `_pad`: 0 = f64 x 3 in grid_data (fallback),
    // 1 / 3 = f32 with that many channels in grid_f32; data_offset then counts floats.
    const float *grid_f32;
    const double *light_pmf, *light_cdf;

    // Shading caches derived on the device at scene creation (lj_scene_build.cu, lj_scene_build_shading_cache):
    //  * tri_shade: one 192-byte record per triangle (global triangle index = lj_shape.triangle_offset + primID):
    //    the f64 positions, normals and uvs of its three vertices, contiguous and 64-byte aligned -- one dependent
    //    fetch of six aligned sectors where the indexed layout needs the index triple first and then nine
    //    unaligned 16/24-byte gathers;
    //  * level 0 of every mip chain whose texels are exactly representable as f32 (image.cpp:94-108 widens
    //    stbi_loadf floats, so they are) as float4 (RGB) / float (scalar): tex0_off*[image] is the image's first
    //    texel in texels0_*, or 0xffffffff when the image keeps its f64 level 0.  Bounce rays carry a zero ray
    //    differential (intersection.h:41), so all but the camera-ray lookups read level 0.
    // Widening back to f64 in registers is exact: shading is bit-identical with and without the caches.
    const double *tri_shade;
    const float4 *texels0_rgb; const float *texels0_r;
    const unsigned *tex0_off3, *tex0_off1;

    // ray casting structures (f32, the Embree seam of intersection.cpp:15-24)
    int num_tris, num_spheres, num_prims;
    const float4 *tri_verts;    // 3 float4 per primitive slot in BVH leaf order: (v0, geom) (v1, prim) (v2, kind)
    const SpherePrim *spheres;
    const Bvh2Node *bvh2; int bvh2_root;
    const Bvh8Node *bvh8;
    // 0x47000000 (the bits of 2^15) as a kernel-parameter value: the node test splices plane bytes into its mantissa with
    // PRMT, and PRMT has one immediate operand -- with this one opaque to the compiler the per-child selector becomes the
    // immediate and the constant comes from the constant bank, instead of four selector registers re-materialised per use
    unsigned m15_bits;
};

// Deal of the 16x16 tiles (render.cpp:75-82) over `world` ranks: tile (tx, ty) gets the dealt index
// v = ty * ntx + (tx + 3 ty) mod ntx -- each tile row rotated by three more columns than the one above, so that a
// rank's tiles form diagonals instead of the vertical stripes `tile id mod world` gives whenever world divides the
// tiles per row (sponza: 48 tiles per row, 8 GPUs) -- and belongs to rank v mod world; v / world numbers a rank's
// tiles densely.  Any fixed deal keeps the partition bit-identical to one device; this one balances it.
LJ_HD int tile_dealt_index(int tx, int ty, int ntx) { return ty * ntx + (tx + 3 * ty) % ntx; }
LJ_HD void tile_from_dealt_index(int v, int ntx, int &tx, int &ty) { ty = v / ntx; tx = (v % ntx + ntx - (3 * ty) % ntx) % ntx; }

// intersection.h:15-35
struct PathVertex {
    d3 position, geometry_normal;
    Frame shading_frame;
    d2 st, uv;
    double uv_screen_size, mean_curvature, ray_radius;
    int shape_id, primitive_id, material_id, interior_medium_id, exterior_medium_id;
};

struct Ray { d3 org, dir; double tnear, tfar; };                 // ray.h:9-12
struct RayDiff { double radius, spread; };                       // ray.h:27-31
struct PointAndNormal { d3 position, normal; };                  // point_and_normal.h:9-12

// What crosses the ray-cast seam back (RTCRay.tfar + RTCHit, all f32 / u32)
struct Hit { float t, u, v; int geom, prim; float ngx, ngy, ngz; };

LJ_D Ray make_ray(d3 org, d3 dir, double tnear, double tfar) { Ray r; r.org = org; r.dir = dir; r.tnear = tnear; r.tfar = tfar; return r; }
LJ_D RayDiff init_ray_differential(int w, int h) { RayDiff r; r.radius = 0; r.spread = 0.25 / (w > h ? w : h); return r; }
LJ_D double rd_transfer(const RayDiff &r, double dist) { return r.radius + r.spread * dist; }
LJ_D double rd_reflect(const RayDiff &r, double mean_curvature, double roughness) {
    double spec_spread = r.spread + 2 * mean_curvature * r.radius;
    double diff_spread = 0.2;
    return fmax(spec_spread * (1 - roughness) + diff_spread * roughness, 0.0);
}
LJ_D double rd_refract(const RayDiff &r, double mean_curvature, double eta, double roughness) {
    double spec_spread = (r.spread + 2 * mean_curvature * r.radius) / eta;
    double diff_spread = 0.2;
    return fmax(spec_spread * (1 - roughness) + diff_spread * roughness, 0.0);
}
