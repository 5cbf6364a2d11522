// Pointwise LSTM cell kernels and a bf16 transpose for the LARGE-hidden-size path
// (BASELINE.json: "scaled 2-layer LSTM hidden 4096 seq 512 bf16").  At H = 4096 the cell
// kernel (8192 x 16384 bf16 = 268 MB per layer) cannot be resident on one SM, so each timestep
// is one full-chip tcgen05 GEMM (csrc/gemm.cu) streaming the weights, followed by this
// pointwise kernel; the input projections of all timesteps and both weight gradients are
// batched into three large GEMMs per layer.  Gate order i, j, f, o (lstm:51 semantics).
#include "../../include/psb200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace psb {

__device__ __forceinline__ float sigm(float x) { return fmaf(0.5f, fast_tanh(0.5f * x), 0.5f); }

// gx, gh: fp32 [B, 4H] pre-activations (x-part and h-part), c: fp32 [B, H] in/out.
__global__ void __launch_bounds__(256) lstm_cell_fwd_kernel(const float* __restrict__ gx, int ld_gx,
                                                            const float* __restrict__ gh, int ld_gh,
                                                            const float* __restrict__ bias, float forget_bias,
                                                            float* c, __nv_bfloat16* h_out, int ld_h,
                                                            __nv_bfloat16* gates_save, float* c_save, int B, int H) {
    // programmatic dependent launch: our launch overlapped the producing GEMM's tail; the next
    // GEMM may start its prologue (barrier init, TMEM allocation) while we run
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (u >= H || b >= B) return;
    const size_t r4 = static_cast<size_t>(b) * 4 * H;
    float pre[4];
#pragma unroll
    for (int g = 0; g < 4; ++g)
        pre[g] = gx[static_cast<size_t>(b) * ld_gx + g * H + u] + gh[static_cast<size_t>(b) * ld_gh + g * H + u] + bias[g * H + u];
    const float ig = sigm(pre[0]), jg = fast_tanh(pre[1]), fg = sigm(pre[2] + forget_bias), og = sigm(pre[3]);
    const size_t ci = static_cast<size_t>(b) * H + u;
    const float cn = fmaf(c[ci], fg, ig * jg);
    c[ci] = cn;
    h_out[static_cast<size_t>(b) * ld_h + u] = __float2bfloat16_rn(fast_tanh(cn) * og);
    if (gates_save != nullptr) {
        gates_save[r4 + u] = __float2bfloat16_rn(ig);
        gates_save[r4 + H + u] = __float2bfloat16_rn(jg);
        gates_save[r4 + 2 * H + u] = __float2bfloat16_rn(fg);
        gates_save[r4 + 3 * H + u] = __float2bfloat16_rn(og);
        c_save[ci] = cn;
    }
}

// dh = dh_rec (fp32, from dG_{t+1} Wh^T; nullable at t = T-1) + dy (fp32 from the layer above /
// the head; nullable).  Writes dG_t bf16 [B, 4H], updates dc in place, atomically adds dbias.
__global__ void __launch_bounds__(256) lstm_cell_bwd_kernel(const float* __restrict__ dh_rec, int ld_dh, int n_parts, long long part_stride,
                                                            const float* __restrict__ dy, int ld_dy,
                                                            const __nv_bfloat16* __restrict__ gates, const float* __restrict__ c_new,
                                                            const float* __restrict__ c_prev, float* dc,
                                                            __nv_bfloat16* dg, int ld_dg, float* dbias, int B, int H) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= H) return;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    // each thread walks batch rows blockIdx.y, +gridDim.y, ...; rows are processed four at a time
    // with every load issued before the first use (the kernel is pure memory latency)
    constexpr int R = 4;
    for (int b0 = blockIdx.y; b0 < B; b0 += gridDim.y * R) {
        float dh[R], gi[R], gj[R], gf[R], go[R], cn[R], co[R], dcv_in[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int b = b0 + r * gridDim.y;
            dh[r] = gi[r] = gj[r] = gf[r] = go[r] = cn[r] = co[r] = dcv_in[r] = 0.0f;
            if (b < B) {
                const size_t r4 = static_cast<size_t>(b) * 4 * H, ci = static_cast<size_t>(b) * H + u;
                if (dh_rec != nullptr)
                    for (int p = 0; p < n_parts; ++p) dh[r] += dh_rec[p * part_stride + static_cast<size_t>(b) * ld_dh + u];
                if (dy != nullptr) dh[r] += dy[static_cast<size_t>(b) * ld_dy + u];
                gi[r] = __bfloat162float(gates[r4 + u]);
                gj[r] = __bfloat162float(gates[r4 + H + u]);
                gf[r] = __bfloat162float(gates[r4 + 2 * H + u]);
                go[r] = __bfloat162float(gates[r4 + 3 * H + u]);
                cn[r] = c_new[ci];
                co[r] = c_prev != nullptr ? c_prev[ci] : 0.0f;
                dcv_in[r] = dc[ci];
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int b = b0 + r * gridDim.y;
            if (b >= B) continue;
            const size_t ci = static_cast<size_t>(b) * H + u;
            const float tc = fast_tanh(cn[r]);
            const float dcv = fmaf(dh[r] * go[r], 1.0f - tc * tc, dcv_in[r]);
            const float d_o = dh[r] * tc * go[r] * (1.0f - go[r]);
            const float d_i = dcv * gj[r] * gi[r] * (1.0f - gi[r]);
            const float d_j = dcv * gi[r] * (1.0f - gj[r] * gj[r]);
            const float d_f = dcv * co[r] * gf[r] * (1.0f - gf[r]);
            dc[ci] = dcv * gf[r];
            __nv_bfloat16* o = dg + static_cast<size_t>(b) * ld_dg;
            o[u] = __float2bfloat16_rn(d_i);
            o[H + u] = __float2bfloat16_rn(d_j);
            o[2 * H + u] = __float2bfloat16_rn(d_f);
            o[3 * H + u] = __float2bfloat16_rn(d_o);
            acc[0] += d_i; acc[1] += d_j; acc[2] += d_f; acc[3] += d_o;
        }
    }
    if (dbias != nullptr) {
#pragma unroll
        for (int g = 0; g < 4; ++g) atomicAdd(dbias + g * H + u, acc[g]);
    }
}

// out[c][r] = in[r][c] for bf16 matrices (32 x 32 tiles through shared memory).
__global__ void __launch_bounds__(256) transpose_bf16_kernel(const __nv_bfloat16* __restrict__ in, int ld_in, int R, int C,
                                                             __nv_bfloat16* out, int ld_out) {
    __shared__ __nv_bfloat16 tile[32][34];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int r = ty; r < 32; r += 8)
        tile[r][tx] = (r0 + r < R && c0 + tx < C) ? in[static_cast<size_t>(r0 + r) * ld_in + c0 + tx] : __float2bfloat16_rn(0.0f);
    __syncthreads();
#pragma unroll
    for (int c = ty; c < 32; c += 8)
        if (c0 + c < C && r0 + tx < R) out[static_cast<size_t>(c0 + c) * ld_out + r0 + tx] = tile[tx][c];
}

}  // namespace psb

extern "C" int psb_lstm_cell_fwd(const float* gx, int ld_gx, const float* gh, int ld_gh, const float* bias,
                                 float forget_bias, float* c, void* h_out, int ld_h, void* gates_save,
                                 float* c_save, int B, int H, void* stream) {
    using namespace psb;
    if (!gx || !gh || !bias || !c || !h_out || B < 1 || H < 1) return fail(PSB_ERR_ARG, "psb_lstm_cell_fwd: bad argument");
    if ((gates_save != nullptr) != (c_save != nullptr)) return fail(PSB_ERR_ARG, "psb_lstm_cell_fwd: gates_save and c_save go together");
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((H + 255) / 256, B);
    cfg.blockDim = dim3(256);
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    PSB_CUDA_OK(cudaLaunchKernelEx(&cfg, lstm_cell_fwd_kernel, gx, ld_gx, gh, ld_gh, bias, forget_bias, c,
                                   static_cast<__nv_bfloat16*>(h_out), ld_h, static_cast<__nv_bfloat16*>(gates_save),
                                   c_save, B, H));
    return after_launch("lstm_cell_fwd");
}

extern "C" int psb_lstm_cell_bwd(const float* dh_rec, int ld_dh, int n_parts, long long part_stride, const float* dy, int ld_dy, const void* gates,
                                 const float* c_new, const float* c_prev, float* dc, void* dg, int ld_dg,
                                 float* dbias, int B, int H, void* stream) {
    using namespace psb;
    if (!gates || !c_new || !dc || !dg || B < 1 || H < 1) return fail(PSB_ERR_ARG, "psb_lstm_cell_bwd: bad argument");
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((H + 255) / 256, B < 16 ? B : 16);
    cfg.blockDim = dim3(256);
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    PSB_CUDA_OK(cudaLaunchKernelEx(&cfg, lstm_cell_bwd_kernel, dh_rec, ld_dh, n_parts < 1 ? 1 : n_parts, part_stride, dy,
                                   ld_dy, static_cast<const __nv_bfloat16*>(gates), c_new, c_prev, dc,
                                   static_cast<__nv_bfloat16*>(dg), ld_dg, dbias, B, H));
    return after_launch("lstm_cell_bwd");
}

extern "C" int psb_transpose_bf16(const void* in, int ld_in, int R, int C, void* out, int ld_out, void* stream) {
    using namespace psb;
    if (!in || !out || R < 1 || C < 1) return fail(PSB_ERR_ARG, "psb_transpose_bf16: bad argument");
    dim3 grid((C + 31) / 32, (R + 31) / 32);
    transpose_bf16_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(in), ld_in, R, C, static_cast<__nv_bfloat16*>(out), ld_out);
    return after_launch("transpose_bf16");
}
