// Dense GEMM with fused epilogue on the 5th-generation tensor cores.
//
//   acc[m,n] = alpha * sum_k A[m,k] * B[n,k]      A:[M,K] bf16, B:[N,K] bf16 (both K-major)
//
// One CTA per 128 x BN output tile, 6 warps:
//   warp 0      TMA producer   (cp.async.bulk.tensor, 128B swizzle, kStages-deep ring)
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer (fp32 accum in TMEM)
//   warps 2..5  epilogue: tcgen05.ld -> bias / activation / activation-derivative ->
//               bf16 row-major, bf16 transposed, fp32, and column-sum outputs
//
// Serves every dense call site of the three reference models (forward, dgrad, wgrad):
// encoder:69-84,137  gan:77-92,158-159  lstm:53,97 — see include/psb200.h.
#include "../../include/psb200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace psb {

constexpr int kBM = 128;          // tile rows   (UMMA M)
constexpr int kBK = 64;           // K per stage (one 128-byte swizzle atom of bf16)
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 192;
constexpr int kMaxSmem = 200 * 1024;

struct EpiArgs {
    float alpha;
    int act;
    int dact;
    const float* bias;
    const __nv_bfloat16* yprev;
    int ld_yprev;
    __nv_bfloat16* out;
    int ld_out;
    __nv_bfloat16* out_t;
    int ld_out_t;
    float* out_f32;
    int ld_f32;
    float* colsum;
    int atomic_f32;      // split-K: partial sums are atomically added into out_f32 (caller zeroes it)
    long long split_stride;   // split-K: CTA z stores its partial at out_f32 + z*split_stride instead (deterministic)
    int b_static;             // B was NOT written by the preceding kernel: its first stages may be fetched before griddepcontrol.wait
    // ---- parameter pull with per-block readiness: B is a weight copy that the owners' push kernels
    // are refreshing row block by row block; a tile is fetched as soon as ITS block's flag shows this step
    const uint32_t* ready;    // flag words, one per row block of the variable behind B (nullptr = no flags)
    const uint32_t* ready_step;   // device word holding the step the flags must reach
    int ready_rows;           // variable rows per block
    int ready_dim;            // 0: the blocks partition K (B = W^T copy), 1: they partition N (B = W copy)
    int ready_nblk;
    int ready_first;          // dim 0: k-blocks are visited starting at this row block (own block first, fixed per rank)
    const uint32_t* bias_ready;   // flag word of the bias variable (nullptr = none)
    uint32_t* ready_status;   // set to PSB_PUSH_ST_TIMEOUT_PARAMS when a wait gives up
    unsigned long long ready_timeout_ns;
    // ---- fused large-H LSTM step (BN == 128 only): B tile = 4 gate blocks x 32 units gathered with four
    // 32-row TMA boxes, epilogue = the BasicLSTMCell pointwise.  Active when lstm_c != nullptr.
    const float* lstm_gx;     // fp32 [B, 4H] x-part pre-activations of this timestep
    int ld_gx;
    const float* lstm_bias;   // fp32 [4H]
    float forget_bias;
    float* lstm_c;            // fp32 [B, H] in/out
    __nv_bfloat16* lstm_h;    // bf16 [B, ld_lstm_h] out
    int ld_lstm_h;
    __nv_bfloat16* lstm_gates;   // bf16 [B, 4H] activated gates (nullable, with lstm_csave)
    float* lstm_csave;        // fp32 [B, H]
    int lstm_H;
};

__host__ __device__ constexpr int stage_bytes(int bn) { return kBM * 128 + bn * 128; }

__host__ inline int pick_stages(int bn, int num_kb) {
    int s = (kMaxSmem - 1024) / stage_bytes(bn);
    if (s > 8) s = 8;
    if (s > num_kb) s = num_kb;
    if (s < 1) s = 1;
    return s;
}

// After the call, lane l of the warp holds (in the return value) the sum over all 32 lanes of
// v[l].  31 shuffles instead of 32 x 5.
__device__ __forceinline__ float warp_transpose_reduce(float (&v)[32], uint32_t lane) {
#pragma unroll
    for (int off = 16, n = 32; off >= 1; off >>= 1, n >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < n / 2; ++i) {
            float send = upper ? v[i] : v[i + n / 2];
            float keep = upper ? v[i + n / 2] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    return v[0];
}

template <int BN>
__device__ __forceinline__ void epilogue_chunk(const EpiArgs& e, float (&v)[32], int row, int n_base,
                                               int M, int N, uint32_t lane, float* s_col_q) {
    const bool row_ok = row < M;
    // ---- bias + activation
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const int n = n_base + j;
        float x = v[j] * e.alpha;
        if (e.bias != nullptr && n < N) x += __ldg(e.bias + n);
        if (e.act == 1) x = fast_sigmoid(x);
        else if (e.act == 2) x = fmaxf(x, 0.0f);
        v[j] = x;
    }
    // ---- multiply by the derivative of the activation that produced yprev
    if (e.yprev != nullptr && row_ok) {
        const __nv_bfloat16* yp = e.yprev + static_cast<size_t>(row) * e.ld_yprev + n_base;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (n_base + 8 * g < e.ld_yprev) {
                uint4 raw = *reinterpret_cast<const uint4*>(yp + 8 * g);
                const __nv_bfloat16* y8 = reinterpret_cast<const __nv_bfloat16*>(&raw);
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    float y = __bfloat162float(y8[t]);
                    float d = (e.dact == 1) ? y * (1.0f - y) : ((y > 0.0f) ? 1.0f : 0.0f);
                    v[8 * g + t] *= d;
                }
            }
        }
    }
    // ---- zero everything outside the logical matrix (pad columns / pad rows)
#pragma unroll
    for (int j = 0; j < 32; ++j)
        if (!row_ok || n_base + j >= N) v[j] = 0.0f;

    // ---- bf16 row-major
    if (e.out != nullptr && row_ok) {
        __nv_bfloat16* o = e.out + static_cast<size_t>(row) * e.ld_out + n_base;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (n_base + 8 * g < e.ld_out && n_base + 8 * g < ((N + 7) & ~7)) {
                uint4 pk;
                __nv_bfloat162* p2 = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                for (int t = 0; t < 4; ++t)
                    p2[t] = __floats2bfloat162_rn(v[8 * g + 2 * t], v[8 * g + 2 * t + 1]);
                *reinterpret_cast<uint4*>(o + 8 * g) = pk;
            }
        }
    }
    // ---- bf16 transposed: for a fixed column the 32 lanes write 32 consecutive rows
    if (e.out_t != nullptr && row_ok) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const int n = n_base + j;
            if (n < N) e.out_t[static_cast<size_t>(n) * e.ld_out_t + row] = __float2bfloat16_rn(v[j]);
        }
    }
    // ---- fp32 row-major
    if (e.out_f32 != nullptr && row_ok && e.atomic_f32) {
        float* o = e.out_f32 + static_cast<size_t>(row) * e.ld_f32 + n_base;
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (n_base + j < N) atomicAdd(o + j, v[j]);
    } else if (e.out_f32 != nullptr && row_ok) {
        float* o = e.out_f32 + static_cast<size_t>(blockIdx.z) * e.split_stride + static_cast<size_t>(row) * e.ld_f32 + n_base;
        if ((N & 3) == 0 && (e.ld_f32 & 3) == 0) {
#pragma unroll
            for (int g = 0; g < 8; ++g)
                if (n_base + 4 * g < N)
                    *reinterpret_cast<float4*>(o + 4 * g) =
                        make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (n_base + j < N) o[j] = v[j];
        }
    }
    // ---- column sums (bias gradient)
    // (per-warp partials go to shared memory; the CTA adds its four row-quadrants in a fixed
    //  order and issues ONE atomic per column, so two M-tiles still sum deterministically)
    if (e.colsum != nullptr) s_col_q[lane] = warp_transpose_reduce(v, lane);
}

// kPullB: the B operand (a weight matrix) is NOT staged by TMA from local memory but loaded by
// the producer warp with 128-bit ld.global from `b_pull` — typically the owner GPU's copy,
// reached over NVLink — and written into the same 128B-swizzled smem tile.  That is the
// parameter pull (all-gather) fused into the consuming GEMM: peer loads of stage s+1.. run
// under the MMAs of stage s.
// Pointwise half of a BasicLSTMCell step on one accumulator row: columns [g*32 + j] hold gate g of
// unit u0 + j.  Eight units at a time: four 8-column TMEM loads, vector loads of gx / bias / c.
__device__ __forceinline__ void lstm_step_epilogue(const EpiArgs& e, uint32_t t_row, int row, int M, int u0) {
    const int H = e.lstm_H;
    const bool ok = row < M;
    const size_t r_gx = static_cast<size_t>(row) * e.ld_gx, r_c = static_cast<size_t>(row) * H,
                 r_g = static_cast<size_t>(row) * 4 * H;
#pragma unroll 1
    for (int j8 = 0; j8 < 32; j8 += 8) {
        uint32_t acc[4][8];
#pragma unroll
        for (int g = 0; g < 4; ++g) tmem_ld_32x8(t_row + g * 32 + j8, acc[g]);
        const int u = u0 + j8;
        float pre[4][8], cv[8];
        if (ok) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const float4 x0 = *reinterpret_cast<const float4*>(e.lstm_gx + r_gx + g * H + u);
                const float4 x1 = *reinterpret_cast<const float4*>(e.lstm_gx + r_gx + g * H + u + 4);
                const float4 b0 = __ldg(reinterpret_cast<const float4*>(e.lstm_bias + g * H + u));
                const float4 b1 = __ldg(reinterpret_cast<const float4*>(e.lstm_bias + g * H + u + 4));
                pre[g][0] = x0.x + b0.x; pre[g][1] = x0.y + b0.y; pre[g][2] = x0.z + b0.z; pre[g][3] = x0.w + b0.w;
                pre[g][4] = x1.x + b1.x; pre[g][5] = x1.y + b1.y; pre[g][6] = x1.z + b1.z; pre[g][7] = x1.w + b1.w;
            }
            const float4 c0 = *reinterpret_cast<const float4*>(e.lstm_c + r_c + u);
            const float4 c1 = *reinterpret_cast<const float4*>(e.lstm_c + r_c + u + 4);
            cv[0] = c0.x; cv[1] = c0.y; cv[2] = c0.z; cv[3] = c0.w; cv[4] = c1.x; cv[5] = c1.y; cv[6] = c1.z; cv[7] = c1.w;
        }
        tmem_ld_wait();
        if (ok) {
            float ig[8], jg[8], fg[8], og[8], hv[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                ig[k] = fmaf(0.5f, fast_tanh(0.5f * (__uint_as_float(acc[0][k]) + pre[0][k])), 0.5f);
                jg[k] = fast_tanh(__uint_as_float(acc[1][k]) + pre[1][k]);
                fg[k] = fmaf(0.5f, fast_tanh(0.5f * (__uint_as_float(acc[2][k]) + pre[2][k] + e.forget_bias)), 0.5f);
                og[k] = fmaf(0.5f, fast_tanh(0.5f * (__uint_as_float(acc[3][k]) + pre[3][k])), 0.5f);
                cv[k] = fmaf(cv[k], fg[k], ig[k] * jg[k]);
                hv[k] = fast_tanh(cv[k]) * og[k];
            }
            *reinterpret_cast<float4*>(e.lstm_c + r_c + u) = make_float4(cv[0], cv[1], cv[2], cv[3]);
            *reinterpret_cast<float4*>(e.lstm_c + r_c + u + 4) = make_float4(cv[4], cv[5], cv[6], cv[7]);
            uint4 hp;
            hp.x = pack_bf16(hv[0], hv[1]); hp.y = pack_bf16(hv[2], hv[3]); hp.z = pack_bf16(hv[4], hv[5]); hp.w = pack_bf16(hv[6], hv[7]);
            *reinterpret_cast<uint4*>(e.lstm_h + static_cast<size_t>(row) * e.ld_lstm_h + u) = hp;
            if (e.lstm_gates != nullptr) {
                uint4 q;
                q.x = pack_bf16(ig[0], ig[1]); q.y = pack_bf16(ig[2], ig[3]); q.z = pack_bf16(ig[4], ig[5]); q.w = pack_bf16(ig[6], ig[7]);
                *reinterpret_cast<uint4*>(e.lstm_gates + r_g + u) = q;
                q.x = pack_bf16(jg[0], jg[1]); q.y = pack_bf16(jg[2], jg[3]); q.z = pack_bf16(jg[4], jg[5]); q.w = pack_bf16(jg[6], jg[7]);
                *reinterpret_cast<uint4*>(e.lstm_gates + r_g + H + u) = q;
                q.x = pack_bf16(fg[0], fg[1]); q.y = pack_bf16(fg[2], fg[3]); q.z = pack_bf16(fg[4], fg[5]); q.w = pack_bf16(fg[6], fg[7]);
                *reinterpret_cast<uint4*>(e.lstm_gates + r_g + 2 * H + u) = q;
                q.x = pack_bf16(og[0], og[1]); q.y = pack_bf16(og[2], og[3]); q.z = pack_bf16(og[4], og[5]); q.w = pack_bf16(og[6], og[7]);
                *reinterpret_cast<uint4*>(e.lstm_gates + r_g + 3 * H + u) = q;
                *reinterpret_cast<float4*>(e.lstm_csave + r_c + u) = make_float4(cv[0], cv[1], cv[2], cv[3]);
                *reinterpret_cast<float4*>(e.lstm_csave + r_c + u + 4) = make_float4(cv[4], cv[5], cv[6], cv[7]);
            }
        }
    }
}

// Block until the owner of row block `blk` has published this step's copy of it.  One thread.
__device__ __forceinline__ void wait_block_ready(const EpiArgs& e, const uint32_t* flag, uint32_t want) {
    unsigned long long t0 = 0;
    uint32_t it = 0;
    while (static_cast<int32_t>(ld_acquire_sys(flag) - want) < 0) {
        if ((++it & 1023u) == 0 && e.ready_timeout_ns != 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > e.ready_timeout_ns) {
                if (e.ready_status != nullptr) atomicMax(e.ready_status, 2u);   // PSB_PUSH_ST_TIMEOUT_PARAMS
                return;
            }
        }
    }
    // what the flag guards was written through the generic proxy (by another GPU); TMA reads it
    // through the async proxy
    asm volatile("fence.proxy.async.global;" ::: "memory");
}

template <int BN, bool kPullB>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                 int M, int N, int K, int stages, EpiArgs epi, const __nv_bfloat16* b_pull, int ldb_pull) {
    extern __shared__ uint8_t smem_raw[];
    // 128B-swizzled tiles need 1024-byte alignment
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int kStage = stage_bytes(BN);
    constexpr uint32_t kTmemCols = BN < 32 ? 32 : BN;

    __shared__ __align__(8) uint64_t full_bar[8];
    __shared__ __align__(8) uint64_t empty_bar[8];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ uint32_t tmem_base_slot;
    __shared__ float s_col[4][BN < 32 ? 32 : BN];

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = lane_id();
    const int m0 = blockIdx.y * kBM;
    const int n0 = blockIdx.x * BN;
    const int total_kb = (K + kBK - 1) / kBK;
    const int kb_per = (total_kb + static_cast<int>(gridDim.z) - 1) / static_cast<int>(gridDim.z);
    const int kb0 = static_cast<int>(blockIdx.z) * kb_per;                 // split-K: this CTA's k-block range
    const int kb1 = (kb0 + kb_per < total_kb) ? kb0 + kb_per : total_kb;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], kPullB ? 33 : 1);    // pull: lane 0's expect_tx arrive + 32 loader lanes
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&tmem_full_bar, 1);
        fence_mbar_init();
    }
    // fused LSTM step: this CTA's 128 B rows are four 32-row boxes, rows g*H + 32*blockIdx.x ..
    const bool lstm = (BN == 128) && !kPullB && epi.lstm_c != nullptr;
    auto load_b = [&](uint8_t* b_dst, uint64_t* bar, int kcoord) {
        if (lstm) {
#pragma unroll
            for (int g = 0; g < 4; ++g)
                tma_load_2d(b_dst + g * 32 * 128, &tma_b, bar, kcoord, g * epi.lstm_H + static_cast<int>(blockIdx.x) * 32);
        } else {
            tma_load_2d(b_dst, &tma_b, bar, kcoord, n0);
        }
    };
    // B operand that does not depend on the previous kernel (weights): start streaming its first
    // stages now, under the previous kernel's tail; A (activations) follows after the wait.
    int pre = 0;
    if (!kPullB && epi.b_static) {
        pre = (kb1 - kb0 < stages) ? (kb1 - kb0) : stages;
        if (warp == 0 && lane == 0) {
            for (int s = 0; s < pre; ++s) {
                mbar_expect_tx(&full_bar[s], kStage);
                load_b(smem + s * kStage + kBM * 128, &full_bar[s], (kb0 + s) * kBK);
            }
        }
    }
    if (warp == 1) tmem_alloc<kTmemCols>(&tmem_base_slot);
    // Programmatic dependent launch: everything above (barrier init, TMEM allocation, descriptor
    // prefetch) overlapped with the tail of the previous kernel in the stream; from here on we
    // touch memory it produced, so wait for it.  Let our own dependent start its prologue now.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ===================== producer =====================
        if constexpr (!kPullB) {
            if (lane == 0) {
                int s = 0;
                uint32_t ph = 0;
                // pull with readiness flags: visit the k-blocks in a rotated order that starts with this
                // rank's own row block (refreshed first, by the push kernel that just ran here) and wait for
                // each block's flag right before its first tile is requested -- the tensor cores work on the
                // blocks that have arrived while the others are still in flight over NVLink
                const bool flags = epi.ready != nullptr;
                uint32_t want = 0;
                int rot = 0;
                if (flags) {
                    want = *reinterpret_cast<const volatile uint32_t*>(epi.ready_step);
                    if (epi.ready_dim == 0) {
                        if (gridDim.z == 1) rot = (epi.ready_first * epi.ready_rows) / kBK;
                    } else {
                        const int nlast = (n0 + BN < N ? n0 + BN : N) - 1;
                        for (int b = n0 / epi.ready_rows; b <= nlast / epi.ready_rows && b < epi.ready_nblk; ++b)
                            wait_block_ready(epi, epi.ready + b, want);
                    }
                }
                int waited = -1;
                for (int kb = kb0; kb < kb1; ++kb) {
                    int kbr = kb;
                    if (rot != 0) { kbr = kb + rot; if (kbr >= kb1) kbr -= (kb1 - kb0); }
                    uint8_t* a_dst = smem + s * kStage;
                    uint8_t* b_dst = a_dst + kBM * 128;
                    if (flags && epi.ready_dim == 0) {
                        const int klast = (kbr * kBK + kBK < K ? kbr * kBK + kBK : K) - 1;
                        for (int b = (kbr * kBK) / epi.ready_rows; b <= klast / epi.ready_rows && b < epi.ready_nblk; ++b)
                            if (b != waited) { wait_block_ready(epi, epi.ready + b, want); waited = b; }
                    }
                    if (kb - kb0 < pre) {
                        // B of this stage is already in flight (and its expect_tx armed)
                        tma_load_2d(a_dst, &tma_a, &full_bar[s], kbr * kBK, m0);
                    } else {
                        mbar_wait(&empty_bar[s], ph ^ 1);
                        mbar_expect_tx(&full_bar[s], kStage);
                        tma_load_2d(a_dst, &tma_a, &full_bar[s], kbr * kBK, m0);
                        load_b(b_dst, &full_bar[s], kbr * kBK);
                    }
                    if (++s == stages) { s = 0; ph ^= 1; }
                }
            }
        } else {
            int s = 0;
            uint32_t ph = 0;
            for (int kb = kb0; kb < kb1; ++kb) {
                if (lane == 0) mbar_wait(&empty_bar[s], ph ^ 1);
                __syncwarp();
                uint8_t* a_dst = smem + s * kStage;
                uint8_t* b_dst = a_dst + kBM * 128;
                if (lane == 0) {
                    mbar_expect_tx(&full_bar[s], kBM * 128);     // A arrives by TMA (local activations)
                    tma_load_2d(a_dst, &tma_a, &full_bar[s], kb * kBK, m0);
                }
                // B: BN rows x 8 chunks of 16 bytes from the (peer) weight copy, 8 loads in flight per lane
                constexpr int kChunks = BN * 8;
#pragma unroll 1
                for (int base = 0; base < kChunks; base += 32 * 8) {
                    float4 v[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int idx = base + j * 32 + static_cast<int>(lane);
                        const int r = idx >> 3, c = idx & 7;
                        const int n = n0 + r, k = kb * kBK + c * 8;
                        v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (idx < kChunks && n < N && k < ldb_pull)
                            v[j] = ld_v4(reinterpret_cast<const float4*>(b_pull + static_cast<size_t>(n) * ldb_pull + k));
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int idx = base + j * 32 + static_cast<int>(lane);
                        const int r = idx >> 3, c = idx & 7;
                        if (idx < kChunks)
                            *reinterpret_cast<float4*>(b_dst + r * 128 + ((c ^ (r & 7)) << 4)) = v[j];
                    }
                }
                fence_proxy_async_smem();
                mbar_arrive(&full_bar[s]);                       // every lane: its part of B is in place
                if (++s == stages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN);
            int s = 0;
            uint32_t ph = 0;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&full_bar[s], ph);
                tc_fence_after_sync();
                const uint32_t a_addr = smem_u32(smem + s * kStage);
                const uint32_t b_addr = a_addr + kBM * 128;
                const uint64_t da = umma_desc_k128(a_addr);
                const uint64_t db = umma_desc_k128(b_addr);
#pragma unroll
                for (int k = 0; k < kBK / kUmmaK; ++k) {
                    // +32 bytes per K=16 step inside the swizzle atom (>>4 => +2)
                    umma_bf16(tmem_base, da + 2 * k, db + 2 * k, idesc, (kb > kb0 || k != 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[s]);   // frees the smem slot when these MMAs retire
                if (++s == stages) { s = 0; ph ^= 1; }
            }
            umma_commit(&tmem_full_bar);      // accumulator complete
        }
    } else {
        // ===================== epilogue =====================
        if (epi.bias_ready != nullptr) {
            // the bias vector is being refreshed by its owner too
            if (lane == 0) wait_block_ready(epi, epi.bias_ready, *reinterpret_cast<const volatile uint32_t*>(epi.ready_step));
            __syncwarp();
        }
        mbar_wait(&tmem_full_bar, 0);
        tc_fence_after_sync();
        const uint32_t quad = warp & 3;       // TMEM lane quadrant this warp may read
        const int row = m0 + static_cast<int>(quad * 32 + lane);
        const uint32_t t_row = tmem_base + ((quad * 32u) << 16);
        if (lstm) {
            lstm_step_epilogue(epi, t_row, row, M, static_cast<int>(blockIdx.x) * 32);
        } else if constexpr (BN >= 32) {
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                if (n0 + c >= ((N + 7) & ~7)) break;   // warp-uniform
                uint32_t r[32];
                tmem_ld_32x32(t_row + c, r);
                tmem_ld_wait();
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                epilogue_chunk<BN>(epi, v, row, n0 + c, M, N, lane, &s_col[quad][c]);
            }
        } else {
            uint32_t r[16];
            tmem_ld_32x16(t_row, r);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
#pragma unroll
            for (int j = 16; j < 32; ++j) v[j] = 0.0f;
            // columns 16..31 are beyond this tile: treat them as outside the matrix
            epilogue_chunk<BN>(epi, v, row, n0, M, (N < n0 + 16) ? N : n0 + 16, lane, &s_col[quad][0]);
        }
        if (epi.colsum != nullptr) {
            asm volatile("bar.sync 1, 128;" ::: "memory");      // the four epilogue warps only
            const int te = static_cast<int>(threadIdx.x) - 64;   // 0..127
            const int ncols = (((N + 7) & ~7) - n0) < BN ? (((N + 7) & ~7) - n0) : BN;
            for (int c = te; c < BN && c < ((ncols + 31) & ~31); c += 128) {
                const float s = ((s_col[0][c] + s_col[1][c]) + s_col[2][c]) + s_col[3][c];
                if (n0 + c < N) atomicAdd(epi.colsum + n0 + c, s);
            }
        }
        tc_fence_before_sync();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

template <int BN, bool kPullB>
static int launch_gemm(const CUtensorMap* ta, const CUtensorMap* tb, int M, int N, int K,
                       const EpiArgs& e, cudaStream_t st, const __nv_bfloat16* b_pull, int ldb_pull, int split_k = 1) {
    const int num_kb = ((K + kBK - 1) / kBK + split_k - 1) / split_k;
    const int stages = pick_stages(BN, num_kb);
    const size_t smem = static_cast<size_t>(stages) * stage_bytes(BN) + 1024;
    static bool attr_set = false;  // per instantiation
    if (!attr_set) {
        PSB_CUDA_OK(cudaFuncSetAttribute(gemm_bf16_kernel<BN, kPullB>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem));
        attr_set = true;
    }
    dim3 grid((N + BN - 1) / BN, (M + kBM - 1) / kBM, split_k);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (grid.x * grid.y * grid.z <= 148) ? 1 : 0;     // only small grids: a waiting dependent must not crowd out work
    PSB_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_bf16_kernel<BN, kPullB>, *ta, *tb, M, N, K, stages, e, b_pull, ldb_pull));
    return after_launch(kPullB ? "gemm_bf16_pull" : "gemm_bf16");
}

}  // namespace psb

static int gemm_entry(const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                      const psb_epilogue_t* epi, int block_n, void* stream, bool pull) {
    using namespace psb;
    const char* who = pull ? "psb_gemm_bf16_pull" : "psb_gemm_bf16";
    if (!A || !B || !epi) return fail(PSB_ERR_ARG, "%s: null pointer argument", who);
    if (M <= 0 || N <= 0 || K <= 0) return fail(PSB_ERR_ARG, "%s: M, N, K must be positive", who);
    if (lda < K || ldb < K) return fail(PSB_ERR_ARG, "%s: leading dimension smaller than K", who);
    if (epi->yprev && (epi->ld_yprev % 8 || epi->dact < 1 || epi->dact > 2))
        return fail(PSB_ERR_ARG, "%s: yprev needs ld%%8==0 and dact in {1,2}", who);
    if (epi->out && epi->ld_out % 8) return fail(PSB_ERR_ARG, "%s: ld_out must be a multiple of 8", who);
    if (epi->out && epi->ld_out < ((N + 7) & ~7))
        return fail(PSB_ERR_ARG, "%s: ld_out must cover N rounded up to 8", who);
    if (pull && (ldb % 8 || (reinterpret_cast<uintptr_t>(B) & 15)))
        return fail(PSB_ERR_ARG, "%s: B must be 16-byte aligned with ldb a multiple of 8", who);

    int bn = block_n;
    if (bn == 0) {
        const int mt = (M + kBM - 1) / kBM;
        bn = (N <= 16) ? 16 : 32;
        const int cands[4] = {256, 128, 64, 32};
        for (int c : cands) {
            if (c <= 16) break;
            if (mt * ((N + c - 1) / c) >= 148 && N >= c) { bn = c; break; }
        }
    }
    const CUtensorMap* ta = tensor_map_bf16_2d(A, M, K, lda, kBM);
    if (!ta) return PSB_ERR_DRIVER;
    // pull_mode 0 (default): the peer copy is read by TMA like a local one -- a tensor map over the
    // peer-mapped address; pull_mode 1: the producer warp's 128-bit ld.global variant
    if (pull && epi->pull_mode == 0) pull = false;
    const CUtensorMap* tb = ta;      // unused by the ld.global pull variant
    if (!pull) {
        tb = tensor_map_bf16_2d(B, N, K, ldb, bn);
        if (!tb) return PSB_ERR_DRIVER;
    }

    EpiArgs e;
    e.alpha = epi->alpha;
    e.act = epi->act;
    e.dact = epi->dact;
    e.bias = epi->bias;
    e.yprev = static_cast<const __nv_bfloat16*>(epi->yprev);
    e.ld_yprev = epi->ld_yprev;
    e.out = static_cast<__nv_bfloat16*>(epi->out);
    e.ld_out = epi->ld_out;
    e.out_t = static_cast<__nv_bfloat16*>(epi->out_t);
    e.ld_out_t = epi->ld_out_t;
    e.out_f32 = epi->out_f32;
    e.ld_f32 = epi->ld_f32;
    e.colsum = epi->colsum;
    e.atomic_f32 = 0;
    e.split_stride = 0;
    e.b_static = epi->b_static;
    e.ready = epi->ready; e.ready_step = epi->ready_step; e.ready_rows = epi->ready_rows; e.ready_dim = epi->ready_dim;
    e.ready_nblk = epi->ready_nblk; e.ready_first = epi->ready_first; e.bias_ready = epi->bias_ready;
    e.ready_status = epi->ready_status; e.ready_timeout_ns = epi->ready_timeout_ns;
    if (e.ready != nullptr || e.bias_ready != nullptr) {
        if (!e.ready_step) return fail(PSB_ERR_ARG, "%s: readiness flags need ready_step", who);
        if (e.ready != nullptr && (e.ready_rows < 1 || e.ready_nblk < 1 || e.ready_first < 0 || e.ready_first >= e.ready_nblk))
            return fail(PSB_ERR_ARG, "%s: bad readiness geometry", who);
        if (pull) return fail(PSB_ERR_ARG, "%s: readiness flags go with the TMA pull (pull_mode 0)", who);
        e.b_static = 0;          // B is exactly what is NOT static
    }
    e.lstm_gx = nullptr; e.ld_gx = 0; e.lstm_bias = nullptr; e.forget_bias = 0.f; e.lstm_c = nullptr; e.lstm_h = nullptr;
    e.ld_lstm_h = 0; e.lstm_gates = nullptr; e.lstm_csave = nullptr; e.lstm_H = 0;
    // split-K for skinny problems with a long reduction (e.g. dh = dG Wh^T of the large-H LSTM:
    // M=64, N=4096, K=16384): partial sums are atomically added into the fp32 output
    int split_k = 1;
    if (epi->split_k > 1) {
        if (!epi->out_f32 || epi->out || epi->out_t || epi->colsum || epi->bias || epi->yprev || epi->act)
            return fail(PSB_ERR_ARG, "%s: split_k needs a plain fp32 output (caller zeroes it)", who);
        split_k = epi->split_k;
        const int total_kb = (K + kBK - 1) / kBK;
        if (split_k > total_kb) split_k = total_kb;
        if (split_k > 32) split_k = 32;
        e.split_stride = epi->split_stride;
        e.atomic_f32 = (split_k > 1 && epi->split_stride == 0) ? 1 : 0;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const __nv_bfloat16* bp = static_cast<const __nv_bfloat16*>(B);
    if (pull) {
        switch (bn) {
            case 16: return launch_gemm<16, true>(ta, tb, M, N, K, e, st, bp, ldb, split_k);
            case 32: return launch_gemm<32, true>(ta, tb, M, N, K, e, st, bp, ldb, split_k);
            case 64: return launch_gemm<64, true>(ta, tb, M, N, K, e, st, bp, ldb, split_k);
            case 128: return launch_gemm<128, true>(ta, tb, M, N, K, e, st, bp, ldb, split_k);
            case 256: return launch_gemm<256, true>(ta, tb, M, N, K, e, st, bp, ldb, split_k);
            default: return fail(PSB_ERR_ARG, "%s: block_n must be 0, 16, 32, 64, 128 or 256", who);
        }
    }
    switch (bn) {
        case 16: return launch_gemm<16, false>(ta, tb, M, N, K, e, st, nullptr, 0, split_k);
        case 32: return launch_gemm<32, false>(ta, tb, M, N, K, e, st, nullptr, 0, split_k);
        case 64: return launch_gemm<64, false>(ta, tb, M, N, K, e, st, nullptr, 0, split_k);
        case 128: return launch_gemm<128, false>(ta, tb, M, N, K, e, st, nullptr, 0, split_k);
        case 256: return launch_gemm<256, false>(ta, tb, M, N, K, e, st, nullptr, 0, split_k);
        default: return fail(PSB_ERR_ARG, "%s: block_n must be 0, 16, 32, 64, 128 or 256", who);
    }
}

extern "C" int psb_gemm_bf16(const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                             const psb_epilogue_t* epi, int block_n, void* stream) {
    return gemm_entry(A, lda, B, ldb, M, N, K, epi, block_n, stream, false);
}

extern "C" int psb_gemm_bf16_pull(const void* A, int lda, const void* B_peer, int ldb, int M, int N, int K,
                                  const psb_epilogue_t* epi, int block_n, void* stream) {
    return gemm_entry(A, lda, B_peer, ldb, M, N, K, epi, block_n, stream, true);
}


extern "C" int psb_lstm_step_fwd(const void* h_prev, int ld_hp, const void* wh_t, int ld_w, int B, int H,
                                 const float* gx, int ld_gx, const float* bias, float forget_bias, float* c,
                                 void* h_out, int ld_h, void* gates_save, float* c_save, void* stream) {
    using namespace psb;
    if (!h_prev || !wh_t || !gx || !bias || !c || !h_out) return fail(PSB_ERR_ARG, "psb_lstm_step_fwd: null pointer argument");
    if (B < 1 || H < 32 || H % 32) return fail(PSB_ERR_ARG, "psb_lstm_step_fwd: H must be a positive multiple of 32");
    if (ld_hp % 8 || ld_w % 8 || ld_h % 8 || ld_gx % 4) return fail(PSB_ERR_ARG, "psb_lstm_step_fwd: leading dimensions must be 16-byte multiples");
    if ((gates_save != nullptr) != (c_save != nullptr)) return fail(PSB_ERR_ARG, "psb_lstm_step_fwd: gates_save and c_save go together");
    const CUtensorMap* ta = tensor_map_bf16_2d(h_prev, B, H, ld_hp, kBM);
    if (!ta) return PSB_ERR_DRIVER;
    const CUtensorMap* tb = tensor_map_bf16_2d(wh_t, 4 * static_cast<uint64_t>(H), H, ld_w, 32);   // 32-row boxes: one gate block each
    if (!tb) return PSB_ERR_DRIVER;
    EpiArgs e = {};
    e.alpha = 1.0f;
    e.b_static = 1;
    e.lstm_gx = gx; e.ld_gx = ld_gx; e.lstm_bias = bias; e.forget_bias = forget_bias; e.lstm_c = c;
    e.lstm_h = static_cast<__nv_bfloat16*>(h_out); e.ld_lstm_h = ld_h;
    e.lstm_gates = static_cast<__nv_bfloat16*>(gates_save); e.lstm_csave = c_save; e.lstm_H = H;
    // N = 128 * (H / 32): every CTA owns 32 hidden units = 4 x 32 gate columns
    return launch_gemm<128, false>(ta, tb, B, 4 * H, H, e, static_cast<cudaStream_t>(stream), nullptr, 0, 1);
}
