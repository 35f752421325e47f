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
