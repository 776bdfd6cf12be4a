// GPU BVH builder interface (host side).  See lj_bvh_build.cu / lj_bvh8_build.cu.
#pragma once
#include "lj_scene.cuh"

struct BvhBuildInput {
    int num_tris, num_spheres;
    const double *positions;    // device
    const int *indices;         // device, 3 per triangle, mesh-local
    const int *tri_shape;       // device, shape id of each global triangle
    const lj_shape *shapes;     // device
    const SpherePrim *spheres;  // device
};

struct BvhBuildOutput {
    int num_prims;
    float lo[3], hi[3];         // f32 scene AABB (what rtcGetSceneBounds returned)
    float4 *tri_verts;          // primitive slots in binary-tree leaf order
    Bvh2Node *bvh2; int bvh2_root, num_bvh2;
    Bvh8Node *bvh8; int num_bvh8;
    float4 *tri_verts8;         // primitive slots in 8-wide-tree order
    int num_slots8;
    int levels8;                // depth of the 8-wide tree (levels of inner nodes), checked against the traversal stack
};

int lj_bvh_build(const BvhBuildInput &in, BvhBuildOutput &out, char *err /* >= 256 bytes */);
void lj_bvh_free(BvhBuildOutput &out);

// 8-wide collapse + quantisation of the binary tree (device arrays of the builder); d_root2 and d_counters are device
// pointers (root id in, 6 counters out: the caller reads them back), see lj_bvh8_build.cu
int lj_bvh8_collapse(int n, const int *d_root2, const float4 *tri_verts, const int *left, const int *right, const int *count,
                     const float4 *node_lo, const float4 *node_hi, const int *sorted_prim,
                     const float4 *prim_lo, const float4 *prim_hi, BvhBuildOutput &out, int *d_counters, char *err);
