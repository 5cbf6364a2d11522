// Host-side plumbing shared by the .cu files of libpsb200_kernels.so: error reporting,
// launch accounting, and a cache of TMA tensor maps.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace psb {

enum : int {
    PSB_OK = 0,
    PSB_ERR_ARG = -1,
    PSB_ERR_CUDA = -2,
    PSB_ERR_DRIVER = -3,
    PSB_ERR_LIMIT = -4,
};

// printf-style; returns `code` so call sites can `return fail(...)`.
int fail(int code, const char* fmt, ...);
void count_launch(uint64_t n = 1);

// Checks the launch (cudaGetLastError) and counts it.
int after_launch(const char* what);

// 2-D bf16 row-major tensor map with 128-byte swizzle; box = [64 elements, box_rows].
// Cached per (base, rows, cols, ld, box_rows).  Returns nullptr on failure (error text set).
const CUtensorMap* tensor_map_bf16_2d(const void* base, uint64_t rows, uint64_t cols,
                                      uint64_t ld_elems, uint32_t box_rows);

#define PSB_CUDA_OK(expr)                                                               \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess)                                                          \
            return ::psb::fail(::psb::PSB_ERR_CUDA, "%s failed: %s", #expr,             \
                               cudaGetErrorString(_e));                                 \
    } while (0)

}  // namespace psb
