// The drop-in binary's one lj_ctx (created on first use; LJ_DEVICE picks the CUDA ordinal), shared by
// render() (render_gpu.cpp) and the asset loaders (assets_gpu.cpp).
#pragma once
#include "lajolla_gpu.h"
lj_ctx *lj_host_ctx();
