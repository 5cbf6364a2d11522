// Fused persistent LSTM (BasicLSTMCell x static_rnn) for H = 128: forward and BPTT.
//
// "Swap-AB" formulation so that a small batch slice becomes the MMA N dimension:
//
//   forward, per timestep:   gates^T[4H, nb] = Wt[4H, I+H] * xh_t^T[I+H, nb]
//       A = Wt (the cell kernel transposed, bf16, resident in shared memory for the whole
//           sequence), B = [x_t, h_{t-1}] of this CTA's 16 batch rows, D in TMEM:
//           gate g occupies TMEM columns [16g, 16g+16), lane = hidden unit.
//       The 128 epilogue threads (thread = hidden unit) read i,j,f,o for their unit and 16
//       batch rows, do the sigmoid/tanh pointwise with c held in registers, and write h_t
//       straight back into the B-operand tile in shared memory for the next step.
//   backward, per timestep:  dh_{t-1}^T[H, nb] = Wh[H, 4H] * dG_t^T[4H, nb]
//       A = rows I.. of the kernel in its natural layout, B = the gate gradients the same
//       threads just produced.
//
// One CTA per 16 batch rows, no inter-CTA synchronisation; the sequence loop never leaves
// the SM.  The weight load at kernel start uses plain 16-byte global loads, so the weight
// pointer may be a peer GPU's HBM (parameter pull fused into the consuming kernel).
//
// ref: rnn.BasicLSTMCell(num_hidden, forget_bias=1.0) + rnn.static_rnn, lstm:50-52, and the
//      backward that optimizer.minimize builds for it, lstm:97.
#include "../../include/psb200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace psb {

constexpr int kH = 128;              // hidden units (UMMA M of one gate block)
constexpr int kNB = 16;              // batch rows per CTA (UMMA N)
constexpr int kLstmThreads = 160;    // 4 epilogue warps + 1 MMA warp
constexpr int kFwdKB = 3;            // k-blocks of 64 covering I+H <= 192
constexpr int kFwdASmem = kFwdKB * 4 * kH * 128;   // 196608
constexpr int kFwdBSmem = kFwdKB * kNB * 128;      // 6144
constexpr int kBwdKB = 8;            // 4H / 64
constexpr int kBwdASmem = kBwdKB * kH * 128;       // 131072
constexpr int kBwdBSmem = kBwdKB * kNB * 128;      // 16384

__device__ __forceinline__ float sigmoid_t(float x) { return fmaf(0.5f, fast_tanh(0.5f * x), 0.5f); }

// byte offset of element (row, k) inside a K-major 128B-swizzled operand made of k-blocks of
// `rows` rows x 128 bytes
__device__ __forceinline__ uint32_t swz_off(int rows, int row, int k) {
    const int kb = k >> 6, kk = k & 63;
    return static_cast<uint32_t>(kb * rows * 128 + row * 128 + ((((kk >> 3) ^ (row & 7)) << 4)) + ((kk & 7) << 1));
}

struct LstmFwdArgs {
    const float* x;              // [B, T, I] fp32
    const __nv_bfloat16* wt;     // [4H, ld_wt] bf16 = kernel^T (may be peer memory)
    int ld_wt;
    const float* bias;           // [4H]
    float forget_bias;
    int B, T, I;
    __nv_bfloat16* h_last;       // [B, ld_h] bf16 (nullable)
    int ld_h;
    __nv_bfloat16* h_last_t;     // [H, ld_ht] bf16 (nullable)
    int ld_ht;
    // saved for backward (all nullable together)
    __nv_bfloat16* gates;        // [T, 4, B, H] bf16, activated i, j, f, o
    float* c_all;                // [T, B, H] fp32, c after step t
    __nv_bfloat16* xh_t;         // [I+H, ld_xht] bf16: row k, column t*B + b = xh_t[b, k]
    int ld_xht;
};

__global__ void __launch_bounds__(kLstmThreads, 1) lstm_fwd_kernel(LstmFwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;
    uint8_t* sB = smem + kFwdASmem;
    __shared__ __align__(8) uint64_t xh_ready, mma_done;
    __shared__ uint32_t tmem_slot;

    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = lane_id();
    const int b0 = blockIdx.x * kNB;
    const int K = a.I + kH;

    if (tid == 0) {
        mbar_init(&xh_ready, 128);
        mbar_init(&mma_done, 1);
        fence_mbar_init();
    }
    if (warp == 4) tmem_alloc<64>(&tmem_slot);

    // ---- weights: global (possibly peer) -> swizzled shared, 16 bytes at a time
    {
        const int chunks_per_row = kFwdKB * 8;                 // 24 chunks of 8 bf16
        const int total = 4 * kH * chunks_per_row;
        for (int idx = tid; idx < total; idx += kLstmThreads) {
            const int r = idx / chunks_per_row, c = idx % chunks_per_row;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (c * 8 < a.ld_wt) v = *reinterpret_cast<const uint4*>(a.wt + static_cast<size_t>(r) * a.ld_wt + c * 8);
            *reinterpret_cast<uint4*>(sA + swz_off(4 * kH, r, c * 8)) = v;
        }
        for (int idx = tid; idx < kFwdBSmem / 16; idx += kLstmThreads)
            reinterpret_cast<uint4*>(sB)[idx] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();   // sB zeroed before x_0 goes in

    const int x_per_step = kNB * a.I;      // elements of x this CTA consumes per step
    if (warp < 4) {
        for (int e = tid; e < x_per_step; e += 128) {
            const int b = e / a.I, i = e % a.I;
            float xv = (b0 + b < a.B) ? a.x[(static_cast<size_t>(b0 + b) * a.T + 0) * a.I + i] : 0.0f;
            __nv_bfloat16 xb = __float2bfloat16_rn(xv);
            *reinterpret_cast<__nv_bfloat16*>(sB + swz_off(kNB, b, i)) = xb;
            if (a.xh_t != nullptr && b0 + b < a.B)
                a.xh_t[static_cast<size_t>(i) * a.ld_xht + 0 * a.B + b0 + b] = xb;
        }
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tmem_slot;

    if (warp == 4) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kH, kNB);
            const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
            const int ksteps = (K + 15) >> 4;
            for (int t = 0; t < a.T; ++t) {
                if (t > 0) {
                    mbar_wait(&xh_ready, (t - 1) & 1);
                    tc_fence_after_sync();
                }
                for (int ks = 0; ks < ksteps; ++ks) {
                    const int kb = ks >> 2, k4 = ks & 3;
                    const uint64_t db = umma_desc_k128(b_base + kb * (kNB * 128)) + 2 * k4;
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const uint64_t da = umma_desc_k128(a_base + kb * (4 * kH * 128) + g * (kH * 128)) + 2 * k4;
                        umma_bf16(tmem + g * kNB, da, db, idesc, ks != 0 ? 1u : 0u);
                    }
                }
                umma_commit(&mma_done);
            }
        }
    } else {
        // ===================== pointwise / state threads (thread = hidden unit) =====================
        const int u = static_cast<int>(tid);
        const uint32_t t_lane = tmem + ((warp * 32u) << 16);
        const float bi = a.bias[u], bj = a.bias[kH + u], bf = a.bias[2 * kH + u] + a.forget_bias,
                    bo = a.bias[3 * kH + u];
        float c[kNB];
#pragma unroll
        for (int b = 0; b < kNB; ++b) c[b] = 0.0f;

        for (int t = 0; t < a.T; ++t) {
            // prefetch x_{t+1} for this slice (independent of the recurrence)
            float xn[4];
            const bool more = t + 1 < a.T;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int e = u + q * 128;
                xn[q] = 0.0f;
                if (more && e < x_per_step) {
                    const int b = e / a.I, i = e % a.I;
                    if (b0 + b < a.B) xn[q] = a.x[(static_cast<size_t>(b0 + b) * a.T + (t + 1)) * a.I + i];
                }
            }
            mbar_wait(&mma_done, t & 1);
            tc_fence_after_sync();
            uint32_t ri[16], rj[16], rf[16], ro[16];
            tmem_ld_32x16(t_lane + 0 * kNB, ri);
            tmem_ld_32x16(t_lane + 1 * kNB, rj);
            tmem_ld_32x16(t_lane + 2 * kNB, rf);
            tmem_ld_32x16(t_lane + 3 * kNB, ro);
            tmem_ld_wait();

            __align__(16) __nv_bfloat16 hb[kNB];
#pragma unroll
            for (int b = 0; b < kNB; ++b) {
                const float ig = sigmoid_t(__uint_as_float(ri[b]) + bi);
                const float jg = fast_tanh(__uint_as_float(rj[b]) + bj);
                const float fg = sigmoid_t(__uint_as_float(rf[b]) + bf);
                const float og = sigmoid_t(__uint_as_float(ro[b]) + bo);
                c[b] = fmaf(c[b], fg, ig * jg);
                const float h = fast_tanh(c[b]) * og;
                hb[b] = __float2bfloat16_rn(h);
                if (a.gates != nullptr && b0 + b < a.B) {
                    const size_t base = (static_cast<size_t>(t) * 4 * a.B + (b0 + b)) * kH + u;
                    const size_t gs = static_cast<size_t>(a.B) * kH;
                    a.gates[base] = __float2bfloat16_rn(ig);
                    a.gates[base + gs] = __float2bfloat16_rn(jg);
                    a.gates[base + 2 * gs] = __float2bfloat16_rn(fg);
                    a.gates[base + 3 * gs] = __float2bfloat16_rn(og);
                    a.c_all[(static_cast<size_t>(t) * a.B + (b0 + b)) * kH + u] = c[b];
                }
            }
            if (more) {
                // h_t -> B operand (columns I..I+H) and x_{t+1} -> columns 0..I
#pragma unroll
                for (int b = 0; b < kNB; ++b)
                    *reinterpret_cast<__nv_bfloat16*>(sB + swz_off(kNB, b, a.I + u)) = hb[b];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int e = u + q * 128;
                    if (e < x_per_step) {
                        const int b = e / a.I, i = e % a.I;
                        __nv_bfloat16 xb = __float2bfloat16_rn(xn[q]);
                        *reinterpret_cast<__nv_bfloat16*>(sB + swz_off(kNB, b, i)) = xb;
                        if (a.xh_t != nullptr && b0 + b < a.B)
                            a.xh_t[static_cast<size_t>(i) * a.ld_xht + static_cast<size_t>(t + 1) * a.B + b0 + b] = xb;
                    }
                }
                if (a.xh_t != nullptr) {
                    // h_t is the recurrent input of step t+1: row I+u, columns (t+1)*B + b0 ..
                    __nv_bfloat16* dst = a.xh_t + static_cast<size_t>(a.I + u) * a.ld_xht +
                                         static_cast<size_t>(t + 1) * a.B + b0;
                    if (b0 + kNB <= a.B) {
                        const uint4* src = reinterpret_cast<const uint4*>(hb);
                        reinterpret_cast<uint4*>(dst)[0] = src[0];
                        reinterpret_cast<uint4*>(dst)[1] = src[1];
                    } else {
                        for (int b = 0; b < kNB && b0 + b < a.B; ++b) dst[b] = hb[b];
                    }
                }
                tc_fence_before_sync();
                fence_proxy_async_smem();
                mbar_arrive(&xh_ready);
            } else {
#pragma unroll
                for (int b = 0; b < kNB; ++b) {
                    if (b0 + b < a.B) {
                        if (a.h_last != nullptr) a.h_last[static_cast<size_t>(b0 + b) * a.ld_h + u] = hb[b];
                        if (a.h_last_t != nullptr) a.h_last_t[static_cast<size_t>(u) * a.ld_ht + b0 + b] = hb[b];
                    }
                }
                tc_fence_before_sync();
            }
        }
    }
    __syncthreads();
    if (warp == 4) {
        tc_fence_after_sync();
        tmem_dealloc<64>(tmem);
    }
}

struct LstmBwdArgs {
    const __nv_bfloat16* w;      // [I+H, ld_w] bf16, natural kernel layout (rows I.. used)
    int ld_w;
    const float* dh_last;        // [B, ld_dh] fp32
    int ld_dh;
    const __nv_bfloat16* gates;  // [T, 4, B, H]
    const float* c_all;          // [T, B, H]
    int B, T, I;
    __nv_bfloat16* dg_t;         // [4H, ld_dgt] bf16: row g*H+u, column t*B + b
    int ld_dgt;
    float* dbias;                // [4H] fp32, atomicAdd
};

__global__ void __launch_bounds__(kLstmThreads, 1) lstm_bwd_kernel(LstmBwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;
    uint8_t* sB = smem + kBwdASmem;
    __shared__ __align__(8) uint64_t dg_ready, mma_done;
    __shared__ uint32_t tmem_slot;

    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = lane_id();
    const int b0 = blockIdx.x * kNB;

    if (tid == 0) {
        mbar_init(&dg_ready, 128);
        mbar_init(&mma_done, 1);
        fence_mbar_init();
    }
    if (warp == 4) tmem_alloc<32>(&tmem_slot);
    {
        // Wh = kernel rows I..I+H, each 4H = 512 bf16 = 64 chunks of 16 bytes
        const int total = kH * 64;
        for (int idx = tid; idx < total; idx += kLstmThreads) {
            const int r = idx >> 6, c = idx & 63;
            uint4 v = *reinterpret_cast<const uint4*>(a.w + static_cast<size_t>(a.I + r) * a.ld_w + c * 8);
            *reinterpret_cast<uint4*>(sA + swz_off(kH, r, c * 8)) = v;
        }
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tmem_slot;

    if (warp == 4) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kH, kNB);
            const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
            for (int s = 0; s < a.T - 1; ++s) {     // no dh needed below t = 0
                mbar_wait(&dg_ready, s & 1);
                tc_fence_after_sync();
#pragma unroll 4
                for (int ks = 0; ks < 32; ++ks) {
                    const int kb = ks >> 2, k4 = ks & 3;
                    const uint64_t da = umma_desc_k128(a_base + kb * (kH * 128)) + 2 * k4;
                    const uint64_t db = umma_desc_k128(b_base + kb * (kNB * 128)) + 2 * k4;
                    umma_bf16(tmem, da, db, idesc, ks != 0 ? 1u : 0u);
                }
                umma_commit(&mma_done);
            }
        }
    } else {
        const int u = static_cast<int>(tid);
        const uint32_t t_lane = tmem + ((warp * 32u) << 16);
        float dh[kNB], dc[kNB];
        float dbs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int b = 0; b < kNB; ++b) {
            dh[b] = (b0 + b < a.B) ? a.dh_last[static_cast<size_t>(b0 + b) * a.ld_dh + u] : 0.0f;
            dc[b] = 0.0f;
        }
        const size_t gs = static_cast<size_t>(a.B) * kH;
        for (int t = a.T - 1, s = 0; t >= 0; --t, ++s) {
            __align__(16) __nv_bfloat16 dgi[kNB], dgj[kNB], dgf[kNB], dgo[kNB];
#pragma unroll
            for (int b = 0; b < kNB; ++b) {
                float di = 0.f, dj = 0.f, df = 0.f, dO = 0.f;
                if (b0 + b < a.B) {
                    const size_t base = (static_cast<size_t>(t) * 4 * a.B + (b0 + b)) * kH + u;
                    const float ig = __bfloat162float(a.gates[base]);
                    const float jg = __bfloat162float(a.gates[base + gs]);
                    const float fg = __bfloat162float(a.gates[base + 2 * gs]);
                    const float og = __bfloat162float(a.gates[base + 3 * gs]);
                    const float cn = a.c_all[(static_cast<size_t>(t) * a.B + (b0 + b)) * kH + u];
                    const float co = (t > 0) ? a.c_all[(static_cast<size_t>(t - 1) * a.B + (b0 + b)) * kH + u] : 0.0f;
                    const float tc = fast_tanh(cn);
                    const float dcv = fmaf(dh[b] * og, 1.0f - tc * tc, dc[b]);
                    dO = dh[b] * tc * og * (1.0f - og);
                    di = dcv * jg * ig * (1.0f - ig);
                    dj = dcv * ig * (1.0f - jg * jg);
                    df = dcv * co * fg * (1.0f - fg);
                    dc[b] = dcv * fg;
                }
                dbs[0] += di; dbs[1] += dj; dbs[2] += df; dbs[3] += dO;
                dgi[b] = __float2bfloat16_rn(di);
                dgj[b] = __float2bfloat16_rn(dj);
                dgf[b] = __float2bfloat16_rn(df);
                dgo[b] = __float2bfloat16_rn(dO);
            }
            // gate gradients for the weight-gradient GEMM: row g*H+u, 16 consecutive columns
            if (a.dg_t != nullptr) {
                const __nv_bfloat16* srcs[4] = {dgi, dgj, dgf, dgo};
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    __nv_bfloat16* dst = a.dg_t + static_cast<size_t>(g * kH + u) * a.ld_dgt +
                                         static_cast<size_t>(t) * a.B + b0;
                    if (b0 + kNB <= a.B) {
                        const uint4* src = reinterpret_cast<const uint4*>(srcs[g]);
                        reinterpret_cast<uint4*>(dst)[0] = src[0];
                        reinterpret_cast<uint4*>(dst)[1] = src[1];
                    } else {
                        for (int b = 0; b < kNB && b0 + b < a.B; ++b) dst[b] = srcs[g][b];
                    }
                }
            }
            if (t > 0) {
                // B operand of dh_{t-1}^T = Wh * dG_t^T : element (row b, col g*H+u)
#pragma unroll
                for (int b = 0; b < kNB; ++b) {
                    *reinterpret_cast<__nv_bfloat16*>(sB + swz_off(kNB, b, 0 * kH + u)) = dgi[b];
                    *reinterpret_cast<__nv_bfloat16*>(sB + swz_off(kNB, b, 1 * kH + u)) = dgj[b];
                    *reinterpret_cast<__nv_bfloat16*>(sB + swz_off(kNB, b, 2 * kH + u)) = dgf[b];
                    *reinterpret_cast<__nv_bfloat16*>(sB + swz_off(kNB, b, 3 * kH + u)) = dgo[b];
                }
                tc_fence_before_sync();
                fence_proxy_async_smem();
                mbar_arrive(&dg_ready);
                mbar_wait(&mma_done, s & 1);
                tc_fence_after_sync();
                uint32_t r[16];
                tmem_ld_32x16(t_lane, r);
                tmem_ld_wait();
#pragma unroll
                for (int b = 0; b < kNB; ++b) dh[b] = __uint_as_float(r[b]);
            }
        }
        if (a.dbias != nullptr) {
#pragma unroll
            for (int g = 0; g < 4; ++g) atomicAdd(a.dbias + g * kH + u, dbs[g]);
        }
        tc_fence_before_sync();
    }
    __syncthreads();
    if (warp == 4) {
        tc_fence_after_sync();
        tmem_dealloc<32>(tmem);
    }
}

}  // namespace psb

extern "C" int psb_lstm_fwd(const float* x, const void* wt, int ld_wt, const float* bias,
                            float forget_bias, int B, int T, int I, int H, void* h_last, int ld_h,
                            void* h_last_t, int ld_ht, void* gates, float* c_all, void* xh_t,
                            int ld_xht, void* stream) {
    using namespace psb;
    if (H != kH) return fail(PSB_ERR_ARG, "psb_lstm_fwd: this kernel is built for H=128 (got %d)", H);
    if (I < 1 || I > 32) return fail(PSB_ERR_ARG, "psb_lstm_fwd: input width must be 1..32 (got %d)", I);
    if (!x || !wt || !bias || B < 1 || T < 1) return fail(PSB_ERR_ARG, "psb_lstm_fwd: bad argument");
    if (ld_wt % 8 || ld_wt < I + H) return fail(PSB_ERR_ARG, "psb_lstm_fwd: ld_wt must be a multiple of 8 covering I+H");
    if ((gates != nullptr) != (c_all != nullptr) || (gates != nullptr) != (xh_t != nullptr))
        return fail(PSB_ERR_ARG, "psb_lstm_fwd: gates, c_all and xh_t must be given together");
    if (xh_t && (ld_xht % 8 || B % 8)) return fail(PSB_ERR_ARG, "psb_lstm_fwd: saving needs B%%8==0 and ld_xht%%8==0");
    LstmFwdArgs a;
    a.x = x; a.wt = static_cast<const __nv_bfloat16*>(wt); a.ld_wt = ld_wt; a.bias = bias;
    a.forget_bias = forget_bias; a.B = B; a.T = T; a.I = I;
    a.h_last = static_cast<__nv_bfloat16*>(h_last); a.ld_h = ld_h;
    a.h_last_t = static_cast<__nv_bfloat16*>(h_last_t); a.ld_ht = ld_ht;
    a.gates = static_cast<__nv_bfloat16*>(gates); a.c_all = c_all;
    a.xh_t = static_cast<__nv_bfloat16*>(xh_t); a.ld_xht = ld_xht;
    const size_t smem = kFwdASmem + kFwdBSmem + 1024;
    static bool attr = false;
    if (!attr) {
        PSB_CUDA_OK(cudaFuncSetAttribute(lstm_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    lstm_fwd_kernel<<<(B + kNB - 1) / kNB, kLstmThreads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    return after_launch("lstm_fwd");
}

extern "C" int psb_lstm_bwd(const void* w, int ld_w, const float* dh_last, int ld_dh, const void* gates,
                            const float* c_all, int B, int T, int I, int H, void* dg_t, int ld_dgt,
                            float* dbias, void* stream) {
    using namespace psb;
    if (H != kH) return fail(PSB_ERR_ARG, "psb_lstm_bwd: this kernel is built for H=128 (got %d)", H);
    if (!w || !dh_last || !gates || !c_all || B < 1 || T < 1) return fail(PSB_ERR_ARG, "psb_lstm_bwd: bad argument");
    if (ld_w % 8 || ld_w < 4 * H) return fail(PSB_ERR_ARG, "psb_lstm_bwd: ld_w must be a multiple of 8 covering 4H");
    if (dg_t && (ld_dgt % 8 || B % 8)) return fail(PSB_ERR_ARG, "psb_lstm_bwd: dg_t needs B%%8==0 and ld_dgt%%8==0");
    LstmBwdArgs a;
    a.w = static_cast<const __nv_bfloat16*>(w); a.ld_w = ld_w; a.dh_last = dh_last; a.ld_dh = ld_dh;
    a.gates = static_cast<const __nv_bfloat16*>(gates); a.c_all = c_all; a.B = B; a.T = T; a.I = I;
    a.dg_t = static_cast<__nv_bfloat16*>(dg_t); a.ld_dgt = ld_dgt; a.dbias = dbias;
    const size_t smem = kBwdASmem + kBwdBSmem + 1024;
    static bool attr = false;
    if (!attr) {
        PSB_CUDA_OK(cudaFuncSetAttribute(lstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    lstm_bwd_kernel<<<(B + kNB - 1) / kNB, kLstmThreads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    return after_launch("lstm_bwd");
}
