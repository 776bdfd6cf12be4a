#include "lj_bvh_build.h"
int lj_bvh8_collapse(int, const Bvh2Node *, const float4 *, const int *, const int *, const float4 *, const float4 *,
                     const int *, const float4 *, const float4 *, BvhBuildOutput &, char *) { return 0; }
