// Volumetric path tracing (vol_path_tracing.h) -- placeholder until the media code lands.
#pragma once
#include "lj_path.cuh"
LJ_D d3 vol_path_tracing_dispatch(const DScene &sc, int x, int y, Pcg32 &rng, RayCounters &rc) {
    return mk3(0.0);
}
