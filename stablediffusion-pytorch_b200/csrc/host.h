// Host-side plumbing shared by every translation unit: error text, launch counter, SM count.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/b200sd.h"

namespace b200 {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int sm_count();

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// Programmatic dependent launch (opt-in: B200SD_PDL=1). With it every kernel of this library is launched with
// programmatic stream serialisation:
// the next kernel in the stream (or captured graph) may become resident as soon as every CTA of this one has started,
// sets up its barriers / TMEM / descriptor prefetches, and blocks in pdl_wait() until this grid has completed and its
// memory is visible. The gap between two dependent kernels shrinks from a full drain + launch to the wait itself.
// Rule for kernels: no global-memory access (read OR write) before pdl_wait().
bool pdl_enabled();

template <typename... P, typename... A>
inline cudaError_t launch_k(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<P>(args)...);
}

// Same, as thread-block clusters of `cluster` CTAs along x (grid.x must be a multiple of it).
template <typename... P, typename... A>
inline cudaError_t launch_k_cluster(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                    unsigned cluster, A&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = cluster;
    at[1].val.clusterDim.y = 1;
    at[1].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<P>(args)...);
}

#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() {
    pdl_trigger();
    pdl_wait();
}
#endif

#define B200_REQUIRE(cond, ...)              \
    do {                                     \
        if (!(cond)) {                       \
            b200::set_error(__VA_ARGS__);    \
            return B200SD_ERR_ARG;           \
        }                                    \
    } while (0)

#define B200_CHECK_LAUNCH(name)                                                      \
    do {                                                                             \
        cudaError_t e__ = cudaGetLastError();                                        \
        if (e__ != cudaSuccess) {                                                    \
            b200::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__)); \
            return B200SD_ERR_CUDA;                                                  \
        }                                                                            \
        b200::count_launch();                                                        \
    } while (0)

}  // namespace b200
