// Device memory for everything the library allocates: the stream-ordered allocator on the legacy default
// stream (all kernels of the library run there).  lj_init raises the release threshold of the device's default
// memory pool, so that the allocate / free pairs of a scene build and of a render are served from memory the
// pool already holds instead of going to the driver each time (which showed up as 100-800 ms stalls in
// repeated scene_create / render / destroy cycles).
#pragma once
#include <cuda_runtime.h>

template <typename T>
static inline cudaError_t lj_malloc(T **p, size_t bytes) { return cudaMallocAsync((void **)p, bytes ? bytes : 1, (cudaStream_t)0); }
static inline cudaError_t lj_free(void *p) { return p ? cudaFreeAsync(p, (cudaStream_t)0) : cudaSuccess; }
