// Entry points of the ray-cast seam used by the integrators.
#pragma once
#include "lj_traverse.cuh"

LJ_D bool trace_closest(const DScene &sc, const RayF &r, Hit *h) { return traverse_bvh2<false>(sc, r, h); }
LJ_D bool trace_any(const DScene &sc, const RayF &r) { return traverse_bvh2<true>(sc, r, nullptr); }
