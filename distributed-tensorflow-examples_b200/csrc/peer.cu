// Peer-memory substrate: one cudaMalloc'd "symmetric heap" per rank, exported to the other
// ranks of the box through CUDA IPC so that kernels can dereference peer pointers over
// NVLink/NVSwitch, plus the device-side noise generator of the GAN.
//
// ref: replaces the gRPC Send/Recv edges TF inserts between /job:ps and /job:worker
//      (encoder:96-98,157 are the only touch points in the reference); noise: gan:189.
#include "../../include/psb200.h"
#include "host_common.h"
#include "ptx.cuh"
#include <string.h>

extern "C" int psb_heap_alloc(size_t bytes, void** out) {
    using namespace psb;
    if (!out || bytes == 0) return fail(PSB_ERR_ARG, "psb_heap_alloc: bad argument");
    void* p = nullptr;
    PSB_CUDA_OK(cudaMalloc(&p, bytes));
    PSB_CUDA_OK(cudaMemset(p, 0, bytes));
    PSB_CUDA_OK(cudaDeviceSynchronize());
    *out = p;
    return PSB_OK;
}

extern "C" int psb_heap_free(void* p) {
    using namespace psb;
    if (p) PSB_CUDA_OK(cudaFree(p));
    return PSB_OK;
}

extern "C" int psb_ipc_export(void* p, unsigned char* handle64) {
    using namespace psb;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!p || !handle64) return fail(PSB_ERR_ARG, "psb_ipc_export: bad argument");
    cudaIpcMemHandle_t h;
    PSB_CUDA_OK(cudaIpcGetMemHandle(&h, p));
    memcpy(handle64, &h, 64);
    return PSB_OK;
}

extern "C" int psb_ipc_open(const unsigned char* handle64, void** out) {
    using namespace psb;
    if (!handle64 || !out) return fail(PSB_ERR_ARG, "psb_ipc_open: bad argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* p = nullptr;
    PSB_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *out = p;
    return PSB_OK;
}

extern "C" int psb_ipc_close(void* p) {
    using namespace psb;
    if (p) PSB_CUDA_OK(cudaIpcCloseMemHandle(p));
    return PSB_OK;
}

namespace psb {

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// out[i] = U(lo, hi), counter-based: (seed, *step_counter, i) -> value; replayable in a graph
// because the step counter lives in device memory.
__global__ void uniform_noise_kernel(float* out, int n, int ld, int cols, float lo, float hi, uint32_t seed,
                                     uint32_t* step_counter, __nv_bfloat16* out_bf16, int ld_bf16,
                                     __nv_bfloat16* out_t, int ld_t) {
    // single block: every thread reads the counter, then thread 0 advances it
    const uint32_t step = step_counter ? *step_counter : 0u;
    __syncthreads();
    if (step_counter && threadIdx.x == 0) *step_counter = step + 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint32_t h = mix32(static_cast<uint32_t>(i) * 0x9E3779B9u + mix32(seed ^ (step * 0x85EBCA6Bu + 0xC2B2AE35u)));
        const float u = (h >> 8) * (1.0f / 16777216.0f);
        const float v = lo + (hi - lo) * u;
        const int r = i / cols, c = i % cols;
        if (out) out[static_cast<size_t>(r) * ld + c] = v;
        if (out_bf16) out_bf16[static_cast<size_t>(r) * ld_bf16 + c] = __float2bfloat16_rn(v);
        if (out_t) out_t[static_cast<size_t>(c) * ld_t + r] = __float2bfloat16_rn(v);
    }
}

}  // namespace psb

extern "C" int psb_uniform_noise(float* out, int rows, int cols, int ld, float lo, float hi, uint32_t seed,
                                 uint32_t* step_counter, void* out_bf16, int ld_bf16, void* out_t,
                                 int ld_t, void* stream) {
    using namespace psb;
    if (rows < 1 || cols < 1 || (!out && !out_bf16 && !out_t)) return fail(PSB_ERR_ARG, "psb_uniform_noise: bad argument");
    const int n = rows * cols;
    uniform_noise_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(
        out, n, ld, cols, lo, hi, seed, step_counter, static_cast<__nv_bfloat16*>(out_bf16), ld_bf16,
        static_cast<__nv_bfloat16*>(out_t), ld_t);
    return after_launch("uniform_noise");
}

// Clear a (possibly strided) 2-D region with a memset node: `rows` runs of `row_bytes` bytes, `pitch`
// bytes apart.  What the plans use instead of a fill kernel for state that must start at zero.
extern "C" int psb_memset_zero(void* p, size_t pitch, size_t row_bytes, size_t rows, void* stream) {
    using namespace psb;
    if (!p || row_bytes == 0 || rows == 0 || pitch < row_bytes) return fail(PSB_ERR_ARG, "psb_memset_zero: bad argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (rows == 1 || pitch == row_bytes) PSB_CUDA_OK(cudaMemsetAsync(p, 0, row_bytes * rows, st));
    else PSB_CUDA_OK(cudaMemset2DAsync(p, pitch, 0, row_bytes, rows, st));
    return PSB_OK;
}
