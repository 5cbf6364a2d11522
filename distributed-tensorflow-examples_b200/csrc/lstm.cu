// Fused persistent LSTM (BasicLSTMCell x static_rnn) for H = 128: forward and BPTT.
//
// "Swap-AB" formulation so that a small batch slice becomes the MMA N dimension:
//
//   forward, per timestep:   gates^T[4H, nb] = Wt[4H, I+H] * xh_t^T[I+H, nb]
//       A = Wt (the cell kernel transposed, bf16, resident in TENSOR MEMORY for the whole
//           sequence: 4 gates x 80 columns, so an MMA never re-reads the weights from shared
//           memory), B = [x_t, h_{t-1}] of this CTA's 16 batch rows, D in TMEM:
//           gate g occupies TMEM columns [16g, 16g+16), lane = hidden unit.
//       256 pointwise threads (thread = hidden unit x half of the batch slice) read i,j,f,o
//       for their unit from TMEM, do the sigmoid/tanh math with c held in registers, and
//       write h_t straight back into the B-operand tile in shared memory for the next step.
//   backward, per timestep:  dh_{t-1}^T[H, nb] = Wh[H, 4H] * dG_t^T[4H, nb]
//       A = rows I.. of the kernel in its natural layout, B = the gate gradients the same
//       threads just produced; the saved activations of step t-1 are prefetched while the
//       MMA of step t is in flight.
//
// One CTA per 16 batch rows, no inter-CTA synchronisation; the sequence loop never leaves
// the SM.  The weight load at kernel start is plain 128-bit ld.global + tcgen05.st, so the
// weight pointer may be a peer GPU's HBM (parameter pull fused into the consuming kernel).
//
// Saved-for-backward blobs (private layout shared by the two kernels; 16-byte vectors, unit
// index fastest so that a warp's store/load is one contiguous 512-byte run):
//   gates[t][cta][g][half][u][8] bf16 (activated i,j,f,o)     c[t][cta][half][2][u][4] fp32
//
// ref: rnn.BasicLSTMCell(num_hidden, forget_bias=1.0) + rnn.static_rnn, lstm:50-52, and the
//      backward that optimizer.minimize builds for it, lstm:97.
#include "../../include/psb200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace psb {

constexpr int kH = 128;              // hidden units (UMMA M of one gate block)
constexpr int kNB = 16;              // batch rows per CTA (UMMA N)
constexpr int kHB = 8;               // batch rows per pointwise thread
constexpr int kPW = 256;             // pointwise threads (8 warps)
constexpr int kMmaWarps = 4;           // backward: issuing warps (a thread sustains ~1 tcgen05.mma / 50 cycles)
constexpr int kLstmThreads = kPW + 32 * kMmaWarps;
constexpr int kFwdMmaWarps = 4;        // forward: one issuing warp per gate block
constexpr int kFwdThreads = kPW + 32 * kFwdMmaWarps;
constexpr int kFwdBSmem = 6144;      // >= 16 x 160 x 2 bytes: the [x_t, h_{t-1}] operand tile
constexpr int kFwdKMax = 160;        // I + H <= 160: 80 TMEM columns per gate block
// [x_t, h_{t-1}] operand, MN-major without swizzle: element (b, k) of the 16 x 160 tile at
//   (b>>3)*kXhSbo + (k>>3)*128 + (k&7)*16 + (b&7)*2   — a thread's 8 batch rows of one k are 16 bytes
constexpr uint32_t kXhSbo = (kFwdKMax / 8) * 128;
// One thread sustains about one tcgen05.mma per ~50 cycles (measured with clock64 around the
// issue loop), so the k-steps are spread over several issuing warps: one per gate block here,
// four interleaved accumulators in the backward kernel.
constexpr uint32_t kFwdAcc = 0;      // TMEM columns [0, 64): gate accumulators (columns 64..127 unused)
constexpr uint32_t kFwdW = 128;      // TMEM columns [128, 128 + 4*80): the cell kernel (A operand)
constexpr int kBwdBSmem = 16 * 4 * kH * 2;      // 16384: the dG_t operand tile (16 x 512 bf16)
constexpr int kStageRow = 144;                 // 128 data bytes + 16 pad: conflict-free 16-byte column writes
constexpr int kFwdStage = kH * kStageRow;      // h rows x (4 steps x 16 batch) bf16, double-buffered
constexpr int kFwdXStage = 32 * kStageRow;     // x rows (I <= 32), same slots, double-buffered
constexpr int kBwdStage = 4 * kH * kStageRow;  // dG rows x (4 steps x 16 batch) bf16, double-buffered
constexpr uint32_t kBwdW = 64;       // TMEM columns [64, 64 + 256): Wh (A operand); [0,64): four partial dh accumulators

__device__ __forceinline__ float sigmoid_t(float x) { return fmaf(0.5f, fast_tanh(0.5f * x), 0.5f); }

__device__ __forceinline__ uint32_t xh_off(int b, int k) {
    return static_cast<uint32_t>((b >> 3) * kXhSbo + (k >> 3) * 128 + (k & 7) * 16 + (b & 7) * 2);
}

// Column order shared by xh_t and dg_t (the two operands of the weight-gradient GEMM):
//   col(cta, t, b) = (cta * T + t) * 16 + b
// so that four consecutive timesteps of one CTA are 128 contiguous bytes of a row.  The kernels
// stage four steps in shared memory and flush them as full 128-byte lines.
__device__ __forceinline__ void flush_group(const uint8_t* stage, int nrows, __nv_bfloat16* dst, int ld,
                                            size_t col0, int valid_slots, uint32_t pw_warp, uint32_t lane) {
    const int per_warp = nrows / 8;                       // 8 pointwise warps; 4 rows per warp-instruction
    const int ch = static_cast<int>(lane) & 7;
    if ((ch >> 1) >= valid_slots) return;
    const int r0 = static_cast<int>(pw_warp) * per_warp + (static_cast<int>(lane) >> 3);
#pragma unroll 4
    for (int k = 0; k < per_warp; k += 16) {
        uint4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = *reinterpret_cast<const uint4*>(stage + (r0 + k + 4 * j) * kStageRow + ch * 16);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            *reinterpret_cast<uint4*>(dst + static_cast<size_t>(r0 + k + 4 * j) * ld + col0 + ch * 8) = v[j];
    }
}

// Same for the (at most 32) x rows of xh_t: one 16-byte chunk per pointwise thread.
__device__ __forceinline__ void flush_x(const uint8_t* stage, int nrows, __nv_bfloat16* dst, int ld, size_t col0,
                                        int valid_slots, uint32_t tid) {
    const int row = static_cast<int>(tid >> 3), ch = static_cast<int>(tid & 7);
    if (row < nrows && (ch >> 1) < valid_slots)
        *reinterpret_cast<uint4*>(dst + static_cast<size_t>(row) * ld + col0 + ch * 8) =
            *reinterpret_cast<const uint4*>(stage + row * kStageRow + ch * 16);
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
    uint4* gates;                // blob, see header comment
    float4* c_all;               // blob
    __nv_bfloat16* xh_t;         // [I+H, ld_xht] bf16: row k, column t*B + b = xh_t[b, k]
    int ld_xht;
};

__global__ void __launch_bounds__(kFwdThreads, 1) lstm_fwd_kernel(LstmFwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sB = smem;
    uint8_t* sStage = smem + kFwdBSmem;                  // 2 x kFwdStage
    uint8_t* sXStage = sStage + 2 * kFwdStage;           // 2 x kFwdXStage
    __shared__ __align__(8) uint64_t xh_ready, mma_done;
    __shared__ uint32_t tmem_slot;

    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = lane_id();
    const int b0 = blockIdx.x * kNB;
    const int K = a.I + kH;

    if (tid == 0) {
        mbar_init(&xh_ready, kPW / 32);       // one arrival per pointwise warp
        mbar_init(&mma_done, kFwdMmaWarps);   // one tcgen05.commit per issuing warp
        fence_mbar_init();
    }
    if (warp == 8) tmem_alloc<512>(&tmem_slot);
    for (int idx = tid; idx < (kFwdBSmem + 2 * kFwdStage + 2 * kFwdXStage) / 16; idx += blockDim.x)
        reinterpret_cast<uint4*>(sB)[idx] = make_uint4(0, 0, 0, 0);     // operand tile and h staging (h_{-1} = 0)
    tc_fence_before_sync();
    __syncthreads();   // sB zeroed before x_0 goes in; TMEM base address published
    tc_fence_after_sync();
    const uint32_t tmem = tmem_slot;

    // ---- weights: global (possibly a peer GPU) -> registers -> TMEM, where they stay for the
    // whole sequence as the A operand.  Thread (u, half) owns row g*H+u of gates 2*half, 2*half+1.
    if (warp < 8) {
        const int u = static_cast<int>(tid & 127), half = static_cast<int>(tid >> 7);
        const uint32_t lane_base = tmem + (((warp & 3u) * 32u) << 16);
#pragma unroll 1
        for (int gg = 0; gg < 2; ++gg) {
            const int g = 2 * half + gg;
            const __nv_bfloat16* row = a.wt + static_cast<size_t>(g * kH + u) * a.ld_wt;
#pragma unroll 1
            for (int c16 = 0; c16 < kFwdKMax / 32; ++c16) {       // 32 bf16 = 16 TMEM columns per batch
                uint32_t r[16];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int k = c16 * 32 + q * 8;
                    uint4 v = make_uint4(0, 0, 0, 0);
                    if (k < a.ld_wt) v = *reinterpret_cast<const uint4*>(row + k);
                    r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
                }
                tmem_st_32x16(lane_base + kFwdW + g * (kFwdKMax / 2) + c16 * 16, r);
            }
        }
        tmem_st_wait();
    }

    const int x_per_step = kNB * a.I;      // <= 512 elements of x per step for this CTA
    if (warp < 8) {
        for (int e = tid; e < x_per_step; e += kPW) {
            const int b = e / a.I, i = e % a.I;
            const float xv = (b0 + b < a.B) ? a.x[(static_cast<size_t>(b0 + b) * a.T + 0) * a.I + i] : 0.0f;
            const __nv_bfloat16 xb = __float2bfloat16_rn(xv);
            *reinterpret_cast<__nv_bfloat16*>(sB + xh_off(b, i)) = xb;
            if (a.gates != nullptr) *reinterpret_cast<__nv_bfloat16*>(sXStage + i * kStageRow + b * 2) = xb;   // column block 0
        }
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();

    if (warp >= 8) {
        // ===================== MMA issuers: warp 8+g owns gate block g =====================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16_bmn(kH, kNB);
            const uint32_t b_base = smem_u32(sB);
            const int g = static_cast<int>(warp) - 8;
            const int ksteps = (K + 15) >> 4;                    // <= 10
            uint64_t db[kFwdKMax / 16];                           // loop-invariant B descriptors
#pragma unroll
            for (int ks = 0; ks < kFwdKMax / 16; ++ks)
                db[ks] = umma_desc_mn_interleaved(b_base + ks * 256, 128, kXhSbo);
            const uint32_t acc = tmem + kFwdAcc + g * kNB, wbase = tmem + kFwdW + g * (kFwdKMax / 2);
            for (int t = 0; t < a.T; ++t) {
                if (t > 0) {
                    mbar_wait(&xh_ready, (t - 1) & 1);
                    tc_fence_after_sync();
                }
#pragma unroll
                for (int ks = 0; ks < kFwdKMax / 16; ++ks)
                    if (ks < ksteps) umma_bf16_ts(acc, wbase + ks * 8, db[ks], idesc, ks != 0 ? 1u : 0u);
                umma_commit(&mma_done);
            }
        }
    } else {
        // ===================== pointwise / state threads =====================
        const int u = static_cast<int>(tid & 127);          // hidden unit == TMEM lane
        const int half = static_cast<int>(tid >> 7);        // which 8 batch rows of the slice
        const int bb = half * kHB;
        const uint32_t t_lane = tmem + (((warp & 3u) * 32u) << 16) + bb;
        const float bi = a.bias[u], bj = a.bias[kH + u], bf = a.bias[2 * kH + u] + a.forget_bias,
                    bo = a.bias[3 * kH + u];
        float c[kHB];
#pragma unroll
        for (int b = 0; b < kHB; ++b) c[b] = 0.0f;

        // loop-invariant addressing (no integer divisions inside the time loop)
        const uint32_t h_off = xh_off(bb, a.I + u);           // this thread's 8 batch rows of k = I+u: 16 bytes
        const float* xsrc[2];
        uint32_t x_off[2], xs_off[2];
        bool x_ok[2];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int e = static_cast<int>(tid) + q * kPW;
            const int b = e / a.I, i = e % a.I;
            x_ok[q] = e < x_per_step;
            x_off[q] = x_ok[q] ? xh_off(b, i) : 0u;
            xs_off[q] = x_ok[q] ? static_cast<uint32_t>(i * kStageRow + b * 2) : 0u;
            const bool in_batch = x_ok[q] && (b0 + b < a.B);
            xsrc[q] = in_batch ? a.x + (static_cast<size_t>(b0 + b) * a.T) * a.I + i : nullptr;
        }
        const bool saving = a.gates != nullptr;
        const size_t cta_col0 = static_cast<size_t>(blockIdx.x) * a.T * kNB;     // first xh_t column of this CTA

        for (int t = 0; t < a.T; ++t) {
            // prefetch x_{t+1} for this slice (independent of the recurrence)
            float xn[2];
            const bool more = t + 1 < a.T;
#pragma unroll
            for (int q = 0; q < 2; ++q)
                xn[q] = (more && xsrc[q] != nullptr) ? xsrc[q][static_cast<size_t>(t + 1) * a.I] : 0.0f;
            if (!more) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // head GEMM may set up now
            mbar_wait(&mma_done, t & 1);
            tc_fence_after_sync();
            uint32_t ri[8], rj[8], rf[8], ro[8];
            tmem_ld_32x8(t_lane + 0 * kNB, ri);
            tmem_ld_32x8(t_lane + 1 * kNB, rj);
            tmem_ld_32x8(t_lane + 2 * kNB, rf);
            tmem_ld_32x8(t_lane + 3 * kNB, ro);
            tmem_ld_wait();

            float ig[kHB], jg[kHB], fg[kHB], og[kHB], h[kHB];
#pragma unroll
            for (int b = 0; b < kHB; ++b) {
                ig[b] = sigmoid_t(__uint_as_float(ri[b]) + bi);
                jg[b] = fast_tanh(__uint_as_float(rj[b]) + bj);
                fg[b] = sigmoid_t(__uint_as_float(rf[b]) + bf);
                og[b] = sigmoid_t(__uint_as_float(ro[b]) + bo);
                c[b] = fmaf(c[b], fg[b], ig[b] * jg[b]);
                h[b] = fast_tanh(c[b]) * og[b];
            }
            uint4 hp;
            hp.x = pack_bf16(h[0], h[1]); hp.y = pack_bf16(h[2], h[3]);
            hp.z = pack_bf16(h[4], h[5]); hp.w = pack_bf16(h[6], h[7]);
            if (more) {
                // h_t -> B operand (columns I..I+H), rows bb..bb+7; x_{t+1} -> columns 0..I
                *reinterpret_cast<uint4*>(sB + h_off) = hp;
                // [x_{t+1}, h_t] is column block s = t+1 of xh_t: stage it (slot s&3 of buffer (s>>2)&1)
                const int sidx = t + 1;
                uint8_t* xs = sXStage + ((sidx >> 2) & 1) * kFwdXStage + (sidx & 3) * 32;
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    if (x_ok[q]) {
                        const __nv_bfloat16 xb = __float2bfloat16_rn(xn[q]);
                        *reinterpret_cast<__nv_bfloat16*>(sB + x_off[q]) = xb;
                        if (saving) *reinterpret_cast<__nv_bfloat16*>(xs + xs_off[q]) = xb;
                    }
                }
                if (saving)
                    *reinterpret_cast<uint4*>(sStage + ((sidx >> 2) & 1) * kFwdStage + u * kStageRow + (sidx & 3) * 32 + half * 16) = hp;
                tc_fence_before_sync();
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&xh_ready);
            }
            // ---- everything below is off the recurrence's critical path
            if (a.gates != nullptr) {
                const size_t blk = static_cast<size_t>(t) * gridDim.x + blockIdx.x;
                uint4* gp = a.gates + blk * (8 * kH) + half * kH + u;      // + g * 2 * kH per gate
                uint4 v;
                v.x = pack_bf16(ig[0], ig[1]); v.y = pack_bf16(ig[2], ig[3]); v.z = pack_bf16(ig[4], ig[5]); v.w = pack_bf16(ig[6], ig[7]);
                gp[0] = v;
                v.x = pack_bf16(jg[0], jg[1]); v.y = pack_bf16(jg[2], jg[3]); v.z = pack_bf16(jg[4], jg[5]); v.w = pack_bf16(jg[6], jg[7]);
                gp[2 * kH] = v;
                v.x = pack_bf16(fg[0], fg[1]); v.y = pack_bf16(fg[2], fg[3]); v.z = pack_bf16(fg[4], fg[5]); v.w = pack_bf16(fg[6], fg[7]);
                gp[4 * kH] = v;
                v.x = pack_bf16(og[0], og[1]); v.y = pack_bf16(og[2], og[3]); v.z = pack_bf16(og[4], og[5]); v.w = pack_bf16(og[6], og[7]);
                gp[6 * kH] = v;
                float4* cp = a.c_all + blk * (4 * kH) + half * (2 * kH) + u;
                cp[0] = make_float4(c[0], c[1], c[2], c[3]);
                cp[kH] = make_float4(c[4], c[5], c[6], c[7]);
                // group (t>>2) of h columns is complete and visible (every warp staged step t-1's
                // h before arriving on xh_ready, and this step's mma_done wait ordered after that)
                if ((t & 3) == 3) {
                    flush_group(sStage + ((t >> 2) & 1) * kFwdStage, kH, a.xh_t + static_cast<size_t>(a.I) * a.ld_xht,
                                a.ld_xht, cta_col0 + static_cast<size_t>(t >> 2) * 64, 4, warp, lane);
                    flush_x(sXStage + ((t >> 2) & 1) * kFwdXStage, a.I, a.xh_t, a.ld_xht,
                            cta_col0 + static_cast<size_t>(t >> 2) * 64, 4, tid);
                }
            }
            if (!more) {
                const uint32_t hv[4] = {hp.x, hp.y, hp.z, hp.w};
#pragma unroll
                for (int b = 0; b < kHB; ++b) {
                    if (b0 + bb + b < a.B) {
                        const uint16_t bits = static_cast<uint16_t>((b & 1) ? (hv[b >> 1] >> 16) : (hv[b >> 1] & 0xffffu));
                        if (a.h_last != nullptr)
                            *reinterpret_cast<uint16_t*>(a.h_last + static_cast<size_t>(b0 + bb + b) * a.ld_h + u) = bits;
                        if (a.h_last_t != nullptr)
                            *reinterpret_cast<uint16_t*>(a.h_last_t + static_cast<size_t>(u) * a.ld_ht + b0 + bb + b) = bits;
                    }
                }
                tc_fence_before_sync();
            }
        }
    }
    __syncthreads();
    if (warp == 8) {
        tc_fence_after_sync();
        tmem_dealloc<512>(tmem);
    }
    if (warp < 8 && a.gates != nullptr && (a.T & 3) != 0) {
        // tail: the last (partial) group of h columns, steps 4*(T>>2) .. T-1
        const int grp = (a.T - 1) >> 2;
        flush_group(sStage + (grp & 1) * kFwdStage, kH, a.xh_t + static_cast<size_t>(a.I) * a.ld_xht, a.ld_xht,
                    static_cast<size_t>(blockIdx.x) * a.T * kNB + static_cast<size_t>(grp) * 64, a.T & 3, warp, lane);
        flush_x(sXStage + (grp & 1) * kFwdXStage, a.I, a.xh_t, a.ld_xht,
                static_cast<size_t>(blockIdx.x) * a.T * kNB + static_cast<size_t>(grp) * 64, a.T & 3, tid);
    }
}

struct LstmBwdArgs {
    const __nv_bfloat16* w;      // [I+H, ld_w] bf16, natural kernel layout (rows I.. used)
    int ld_w;
    const float* dh_last;        // [B, ld_dh] fp32
    int ld_dh;
    const uint4* gates;          // blob written by lstm_fwd_kernel
    const float4* c_all;         // blob
    int B, T, I;
    __nv_bfloat16* dg_t;         // [4H, ld_dgt] bf16: row g*H+u, column t*B + b
    int ld_dgt;
    float* dbias;                // [4H] fp32, += sum over CTAs in CTA order
    float* dbias_part;           // scratch [gridDim.x][4H] fp32 + 1 counter word at the end
    psb_lstm_head_t head;        // head.w_out != nullptr: classifier head fused into the prologue
};

// Per-CTA partials of the fused head: dW_out [H][16], db_out [16], loss, correct.
constexpr int kHeadC = 16;
constexpr int kHeadPart = kH * kHeadC + kHeadC + 2;

__device__ __forceinline__ float grp16_max(float v) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float grp16_sum(float v) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int grp16_min(int v) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(kLstmThreads, 1) lstm_bwd_kernel(LstmBwdArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sB = smem;
    uint8_t* sStage = smem + kBwdBSmem;                  // 2 x kBwdStage
    __shared__ __align__(8) uint64_t dg_ready, mma_done;
    __shared__ uint32_t tmem_slot;

    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = lane_id();
    const int b0 = blockIdx.x * kNB;

    if (tid == 0) {
        mbar_init(&dg_ready, kPW / 32);
        mbar_init(&mma_done, kMmaWarps);
        fence_mbar_init();
    }
    if (warp == 8) tmem_alloc<512>(&tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (warp < 8) {
        // Wh = kernel rows I..I+H (4H = 512 bf16 each) -> TMEM columns [kBwdW, kBwdW+256);
        // thread (u, half) loads K elements [256*half, 256*half+256) of row I+u
        const int u = static_cast<int>(tid & 127), half = static_cast<int>(tid >> 7);
        const uint32_t lane_base = tmem + (((warp & 3u) * 32u) << 16);
        const __nv_bfloat16* row = a.w + static_cast<size_t>(a.I + u) * a.ld_w + half * 256;
#pragma unroll 1
        for (int c16 = 0; c16 < 8; ++c16) {
            uint32_t r[16];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint4 v = *reinterpret_cast<const uint4*>(row + c16 * 32 + q * 8);
                r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
            }
            tmem_st_32x16(lane_base + kBwdW + half * 128 + c16 * 16, r);
        }
        tmem_st_wait();
    }
    // Head operands that do not depend on the forward kernel (parameters, labels) are fetched here,
    // under the previous kernel's tail: W_out -> shared memory, this thread's label and bias.
    float head_y = 0.0f, head_b = 0.0f;
    if (warp < 8 && a.head.w_out != nullptr) {
        const psb_lstm_head_t& hd = a.head;
        float* sW = reinterpret_cast<float*>(sStage) + kNB * 129;
        const __nv_bfloat16* wo = static_cast<const __nv_bfloat16*>(hd.w_out);
        for (int e = static_cast<int>(tid); e < kH * kHeadC; e += kPW) {
            const int j = e >> 4, k = e & 15;
            sW[j * 17 + k] = (k < hd.C) ? __bfloat162float(wo[static_cast<size_t>(j) * hd.ld_wo + k]) : 0.0f;
        }
        const int b = static_cast<int>(tid >> 4), k = static_cast<int>(tid & 15);
        if (k < hd.C) {
            head_b = hd.b_out[k];
            if (b0 + b < a.B) head_y = hd.labels[static_cast<size_t>(b0 + b) * hd.ld_l + k];
        }
    }
    // Wh came from the parameter arena (not written by the kernels just before us): it was loaded
    // under the tail of the head dgrad GEMM.  From here on we read what those kernels produced.
    if (a.head.zero_ptr != nullptr) {
        // Clear the buffer the following split-K GEMM accumulates into (nobody reads or writes it
        // until this grid has finished), still under the previous kernel's tail.
        uint4* z = static_cast<uint4*>(a.head.zero_ptr);
        const size_t n16 = a.head.zero_bytes >> 4;
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + tid; i < n16; i += static_cast<size_t>(gridDim.x) * blockDim.x)
            z[i] = make_uint4(0, 0, 0, 0);
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();

    if (warp >= 8) {
        // a single thread sustains roughly one tcgen05.mma per ~50 cycles, so the 32 k-steps are
        // issued by four warps, each accumulating its quarter into its own TMEM accumulator
        if (lane == 0) {
            // dG_t is stored MN-major (the 16 batch rows of one k contiguous): element (b, k) at
            //   (b>>3)*8192 + (k>>3)*128 + (k&7)*16 + (b&7)*2
            // so a pointwise thread (= one k, eight b) writes its gate gradient with ONE 16-byte store
            constexpr uint32_t idesc = umma_idesc_bf16_bmn(kH, kNB);
            const uint32_t b_base = smem_u32(sB);
            const int j = static_cast<int>(warp) - 8;
            uint64_t db[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int ks = 4 * q + j;
                db[q] = umma_desc_mn_interleaved(b_base + ks * 256, 128, 8192);
            }
            for (int s = 0; s < a.T - 1; ++s) {     // no dh needed below t = 0
                mbar_wait(&dg_ready, s & 1);
                tc_fence_after_sync();
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    umma_bf16_ts(tmem + j * kNB, tmem + kBwdW + (4 * q + j) * 8, db[q], idesc, q != 0 ? 1u : 0u);
                umma_commit(&mma_done);
            }
        }
    } else {
        const int u = static_cast<int>(tid & 127);
        const int half = static_cast<int>(tid >> 7);
        const int bb = half * kHB;
        const uint32_t t_lane = tmem + (((warp & 3u) * 32u) << 16) + bb;
        const size_t cta_col0 = static_cast<size_t>(blockIdx.x) * a.T * kNB;
        float dh[kHB], dc[kHB], c_new[kHB], c_old[kHB];
        float dbs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int b = 0; b < kHB; ++b) dc[b] = 0.0f;
        if (a.head.w_out != nullptr) {
            // ---- fused classifier head for this CTA's 16 rows (lstm:53,92-94 and their gradients):
            // logits = h_T W + b, softmax cross-entropy, dlogits, dh_T = dlogits W^T, and this CTA's
            // partial dW = h_T^T dlogits / db / loss / correct (summed in CTA order by the last CTA).
            // The staging buffers are free until the first timestep: they hold the temporaries.
            const psb_lstm_head_t& hd = a.head;
            const int C = hd.C;
            float* sh = reinterpret_cast<float*>(sStage);        // [16][129] h_T
            float* sW = sh + kNB * 129;                          // [128][17] W_out
            float* sdl = sW + kH * 17;                           // [16][17]  dlogits
            float* sred = sdl + kNB * 17;                        // [16][2]   loss, correct per row
            const __nv_bfloat16* hl = static_cast<const __nv_bfloat16*>(hd.h_last);
            {   // h_T rows of this CTA: one 16-byte load (8 units) per thread
                const int b = static_cast<int>(tid >> 4), j0 = static_cast<int>(tid & 15) * 8;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (b0 + b < a.B) v = *reinterpret_cast<const uint4*>(hl + static_cast<size_t>(b0 + b) * hd.ld_h + j0);
                const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    sh[b * 129 + j0 + 2 * q] = bf16_lo(w4[q]);
                    sh[b * 129 + j0 + 2 * q + 1] = bf16_hi(w4[q]);
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            {
                const int b = static_cast<int>(tid >> 4), k = static_cast<int>(tid & 15);
                const bool in_c = k < C, valid = in_c && (b0 + b < a.B);
                float z = 0.0f;
#pragma unroll 8
                for (int j = 0; j < kH; ++j) z = fmaf(sh[b * 129 + j], sW[j * 17 + k], z);
                z += head_b;
                const float y = head_y;
                const float zz = in_c ? z : -INFINITY;
                const float mx = grp16_max(zz);
                const int am = grp16_min((in_c && zz == mx) ? k : 99);      // first maximum, like the loss kernel
                const float ymax = grp16_max(in_c ? y : -INFINITY);
                const int ay = grp16_min((in_c && y == ymax) ? k : 99);
                const float ysum = grp16_sum(y);
                const float se = grp16_sum(in_c ? __expf(zz - mx) : 0.0f);
                const float lse = mx + __logf(se);
                const float g = valid ? (__expf(zz - lse) * ysum - y) / static_cast<float>(a.B) : 0.0f;
                const float lrow = grp16_sum(valid ? -y * (zz - lse) : 0.0f);
                sdl[b * 17 + k] = g;
                if (valid && hd.logits != nullptr) hd.logits[static_cast<size_t>(b0 + b) * hd.ld_logits + k] = z;
                if (k == 0) {
                    const bool row = b0 + b < a.B;
                    sred[2 * b] = row ? lrow : 0.0f;
                    sred[2 * b + 1] = (row && am == ay) ? 1.0f : 0.0f;
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
#pragma unroll
            for (int b = 0; b < kHB; ++b) {
                float acc = 0.0f;
#pragma unroll
                for (int k = 0; k < kHeadC; ++k) acc = fmaf(sdl[(bb + b) * 17 + k], sW[u * 17 + k], acc);
                dh[b] = acc;
            }
            float* part = hd.scratch + static_cast<size_t>(blockIdx.x) * kHeadPart;
#pragma unroll
            for (int kk = 0; kk < kHeadC / 2; ++kk) {
                const int k = half * (kHeadC / 2) + kk;
                float acc = 0.0f;
#pragma unroll
                for (int b = 0; b < kNB; ++b) acc = fmaf(sh[b * 129 + u], sdl[b * 17 + k], acc);
                part[u * kHeadC + k] = acc;
            }
            if (tid < kHeadC) {
                float acc = 0.0f;
                for (int b = 0; b < kNB; ++b) acc += sdl[b * 17 + tid];
                part[kH * kHeadC + tid] = acc;
            } else if (tid < kHeadC + 2) {
                float acc = 0.0f;
                for (int b = 0; b < kNB; ++b) acc += sred[2 * b + (tid - kHeadC)];
                part[kH * kHeadC + tid] = acc;
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");      // temporaries dead: staging may be written
        } else {
#pragma unroll
            for (int b = 0; b < kHB; ++b)
                dh[b] = (b0 + bb + b < a.B) ? a.dh_last[static_cast<size_t>(b0 + bb + b) * a.ld_dh + u] : 0.0f;
        }
        // saved activations of the last step
        uint4 gq[4];
        {
            const size_t blk = static_cast<size_t>(a.T - 1) * gridDim.x + blockIdx.x;
            const uint4* gp = a.gates + blk * (8 * kH) + half * kH + u;
#pragma unroll
            for (int g = 0; g < 4; ++g) gq[g] = gp[g * 2 * kH];
            const float4* cp = a.c_all + blk * (4 * kH) + half * (2 * kH) + u;
            const float4 c0 = cp[0], c1 = cp[kH];
            c_new[0] = c0.x; c_new[1] = c0.y; c_new[2] = c0.z; c_new[3] = c0.w;
            c_new[4] = c1.x; c_new[5] = c1.y; c_new[6] = c1.z; c_new[7] = c1.w;
            if (a.T > 1) {
                const float4* pp = cp - static_cast<size_t>(gridDim.x) * (4 * kH);
                const float4 p0 = pp[0], p1 = pp[kH];
                c_old[0] = p0.x; c_old[1] = p0.y; c_old[2] = p0.z; c_old[3] = p0.w;
                c_old[4] = p1.x; c_old[5] = p1.y; c_old[6] = p1.z; c_old[7] = p1.w;
            } else {
#pragma unroll
                for (int b = 0; b < kHB; ++b) c_old[b] = 0.0f;
            }
        }
        // this thread's 16-byte slot in the MN-major dG operand (gate 0; +2048 per gate)
        const uint32_t dg_off = static_cast<uint32_t>(half * 8192 + (u >> 3) * 128 + (u & 7) * 16);
        for (int t = a.T - 1, s = 0; t >= 0; --t, ++s) {
            if (t == 0) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // wgrad GEMM may set up now
            // prefetch the saved activations of step t-1 now: they have the whole step to arrive
            uint4 gq_n[4];
            float4 cp0 = make_float4(0.f, 0.f, 0.f, 0.f), cp1 = cp0;
            if (t > 0) {
                const size_t blk = static_cast<size_t>(t - 1) * gridDim.x + blockIdx.x;
                const uint4* gp = a.gates + blk * (8 * kH) + half * kH + u;
#pragma unroll
                for (int g = 0; g < 4; ++g) gq_n[g] = gp[g * 2 * kH];
                if (t > 1) {
                    const float4* pp = a.c_all + (blk - gridDim.x) * (4 * kH) + half * (2 * kH) + u;
                    cp0 = pp[0];
                    cp1 = pp[kH];
                }
            }
            float di[kHB], dj[kHB], df[kHB], dO[kHB];
            const uint32_t gi[4] = {gq[0].x, gq[0].y, gq[0].z, gq[0].w};
            const uint32_t gj[4] = {gq[1].x, gq[1].y, gq[1].z, gq[1].w};
            const uint32_t gf[4] = {gq[2].x, gq[2].y, gq[2].z, gq[2].w};
            const uint32_t go[4] = {gq[3].x, gq[3].y, gq[3].z, gq[3].w};
#pragma unroll
            for (int b = 0; b < kHB; ++b) {
                const float ig = (b & 1) ? bf16_hi(gi[b >> 1]) : bf16_lo(gi[b >> 1]);
                const float jg = (b & 1) ? bf16_hi(gj[b >> 1]) : bf16_lo(gj[b >> 1]);
                const float fg = (b & 1) ? bf16_hi(gf[b >> 1]) : bf16_lo(gf[b >> 1]);
                const float og = (b & 1) ? bf16_hi(go[b >> 1]) : bf16_lo(go[b >> 1]);
                const float tc = fast_tanh(c_new[b]);
                const float dcv = fmaf(dh[b] * og, 1.0f - tc * tc, dc[b]);
                dO[b] = dh[b] * tc * og * (1.0f - og);
                di[b] = dcv * jg * ig * (1.0f - ig);
                dj[b] = dcv * ig * (1.0f - jg * jg);
                df[b] = dcv * c_old[b] * fg * (1.0f - fg);
                dc[b] = dcv * fg;
                dbs[0] += di[b]; dbs[1] += dj[b]; dbs[2] += df[b]; dbs[3] += dO[b];
            }
            uint4 pk[4];
            pk[0].x = pack_bf16(di[0], di[1]); pk[0].y = pack_bf16(di[2], di[3]); pk[0].z = pack_bf16(di[4], di[5]); pk[0].w = pack_bf16(di[6], di[7]);
            pk[1].x = pack_bf16(dj[0], dj[1]); pk[1].y = pack_bf16(dj[2], dj[3]); pk[1].z = pack_bf16(dj[4], dj[5]); pk[1].w = pack_bf16(dj[6], dj[7]);
            pk[2].x = pack_bf16(df[0], df[1]); pk[2].y = pack_bf16(df[2], df[3]); pk[2].z = pack_bf16(df[4], df[5]); pk[2].w = pack_bf16(df[6], df[7]);
            pk[3].x = pack_bf16(dO[0], dO[1]); pk[3].y = pack_bf16(dO[2], dO[3]); pk[3].z = pack_bf16(dO[4], dO[5]); pk[3].w = pack_bf16(dO[6], dO[7]);
            if (a.dg_t != nullptr) {
                // gate gradients of step t -> staging (row g*H+u, slot t&3 of buffer (t>>2)&1)
                uint8_t* st = sStage + ((t >> 2) & 1) * kBwdStage + (t & 3) * 32 + half * 16;
#pragma unroll
                for (int g = 0; g < 4; ++g) *reinterpret_cast<uint4*>(st + (g * kH + u) * kStageRow) = pk[g];
            }
            if (t > 0) {
                // B operand of dh_{t-1}^T = Wh * dG_t^T : element (row b, col g*H+u)
#pragma unroll
                for (int g = 0; g < 4; ++g) *reinterpret_cast<uint4*>(sB + dg_off + g * (kH / 8) * 128) = pk[g];
                tc_fence_before_sync();
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&dg_ready);
            }
            // (the stage write below happened before the arrive; the flush of a completed group of
            //  four timesteps is issued here, off the critical path)
            if (a.dg_t != nullptr && t + 1 < a.T && ((t + 1) & 3) == 0)
                flush_group(sStage + (((t + 1) >> 2) & 1) * kBwdStage, 4 * kH, a.dg_t, a.ld_dgt,
                            cta_col0 + static_cast<size_t>((t + 1) >> 2) * 64,
                            (a.T - (t + 1)) < 4 ? (a.T - (t + 1)) : 4, warp, lane);
            if (t > 0) {
#pragma unroll
                for (int g = 0; g < 4; ++g) gq[g] = gq_n[g];
#pragma unroll
                for (int b = 0; b < kHB; ++b) c_new[b] = c_old[b];
                c_old[0] = cp0.x; c_old[1] = cp0.y; c_old[2] = cp0.z; c_old[3] = cp0.w;
                c_old[4] = cp1.x; c_old[5] = cp1.y; c_old[6] = cp1.z; c_old[7] = cp1.w;
                mbar_wait(&mma_done, s & 1);
                tc_fence_after_sync();
                uint32_t r0[8], r1[8], r2[8], r3[8];
                tmem_ld_32x8(t_lane, r0);
                tmem_ld_32x8(t_lane + kNB, r1);
                tmem_ld_32x8(t_lane + 2 * kNB, r2);
                tmem_ld_32x8(t_lane + 3 * kNB, r3);
                tmem_ld_wait();
#pragma unroll
                for (int b = 0; b < kHB; ++b)
                    dh[b] = (__uint_as_float(r0[b]) + __uint_as_float(r1[b])) + (__uint_as_float(r2[b]) + __uint_as_float(r3[b]));
            }
        }
        if (a.dbias != nullptr) {
            // both halves of the batch slice -> one partial per (cta, gate, unit), summed in a
            // fixed order so the bias gradient is bit-reproducible
            float* part = a.dbias_part + static_cast<size_t>(blockIdx.x) * 4 * kH;
            if (half == 1) {
#pragma unroll
                for (int g = 0; g < 4; ++g) part[g * kH + u] = dbs[g];
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (half == 0) {
#pragma unroll
                for (int g = 0; g < 4; ++g) part[g * kH + u] += dbs[g];
            }
        }
        tc_fence_before_sync();
    }
    __syncthreads();
    if (warp == 8) {
        tc_fence_after_sync();
        tmem_dealloc<512>(tmem);
    }
    if (warp < 8 && a.dg_t != nullptr) {
        // group 0 (steps 0..3) completes with the last iteration: flush it after the CTA-wide sync
        flush_group(sStage, 4 * kH, a.dg_t, a.ld_dgt, static_cast<size_t>(blockIdx.x) * a.T * kNB,
                    a.T < 4 ? a.T : 4, warp, lane);
    }
    if (a.dbias_part != nullptr) {
        // the last CTA to get here adds the per-CTA partials in CTA order
        __shared__ bool last;
        unsigned int* counter = reinterpret_cast<unsigned int*>(a.dbias_part + static_cast<size_t>(gridDim.x) * 4 * kH);
        if (tid == 0) {
            __threadfence();
            last = (atomicAdd(counter, 1u) == gridDim.x - 1);
        }
        __syncthreads();
        if (last) {
            __threadfence();
            const bool keep = a.head.w_out == nullptr || a.head.accumulate != 0;     // += or =
            for (int i = tid; a.dbias != nullptr && i < 4 * kH; i += blockDim.x) {
                float s = 0.0f;
                for (unsigned int cta = 0; cta < gridDim.x; ++cta)
                    s += *reinterpret_cast<volatile float*>(a.dbias_part + static_cast<size_t>(cta) * 4 * kH + i);
                a.dbias[i] = keep ? a.dbias[i] + s : s;
            }
            if (a.head.w_out != nullptr) {
                const psb_lstm_head_t& hd = a.head;
                for (int i = tid; i < kHeadPart; i += blockDim.x) {
                    const int k = i & (kHeadC - 1);
                    if (i < kH * kHeadC + kHeadC && k >= hd.C) continue;
                    float s = 0.0f;
                    for (unsigned int cta = 0; cta < gridDim.x; ++cta)
                        s += *reinterpret_cast<volatile float*>(hd.scratch + static_cast<size_t>(cta) * kHeadPart + i);
                    if (i < kH * kHeadC) {
                        float* d = hd.d_w_out + static_cast<size_t>(i >> 4) * hd.ld_dwo + k;
                        if (hd.d_w_out != nullptr) *d = keep ? *d + s : s;
                    } else if (i < kH * kHeadC + kHeadC) {
                        if (hd.d_b_out != nullptr) hd.d_b_out[k] = keep ? hd.d_b_out[k] + s : s;
                    }
                    else if (i == kH * kHeadC + kHeadC) { if (hd.loss != nullptr) *hd.loss = s / static_cast<float>(a.B); }
                    else if (hd.correct != nullptr) *hd.correct = s;
                }
            }
            if (tid == 0) *counter = 0;
        }
    }
}

}  // namespace psb

extern "C" size_t psb_lstm_saved_bytes(int B, int T, int H, int which) {
    // which: 0 = gates blob (8 uint4 per (t, cta, unit)), 1 = c blob (4 float4 per (t, cta, unit))
    const size_t ctas = (static_cast<size_t>(B) + psb::kNB - 1) / psb::kNB;
    const size_t rows = static_cast<size_t>(T) * ctas * static_cast<size_t>(H);
    return which == 0 ? rows * 8 * 16 : rows * 4 * 16;
}

extern "C" int psb_lstm_fwd(const float* x, const void* wt, int ld_wt, const float* bias,
                            float forget_bias, int B, int T, int I, int H, void* h_last, int ld_h,
                            void* h_last_t, int ld_ht, void* gates, float* c_all, void* xh_t,
                            int ld_xht, void* stream) {
    using namespace psb;
    if (H != kH) return fail(PSB_ERR_ARG, "psb_lstm_fwd: this kernel is built for H=128 (got %d)", H);
    if (I < 1 || I > 32) return fail(PSB_ERR_ARG, "psb_lstm_fwd: input width must be 1..32 (got %d)", I);
    if (!x || !wt || !bias || B < 1 || T < 1) return fail(PSB_ERR_ARG, "psb_lstm_fwd: bad argument");
    if (ld_wt % 8 || ld_wt < I + H || ld_wt > kFwdKMax)
        return fail(PSB_ERR_ARG, "psb_lstm_fwd: ld_wt must be a multiple of 8 with I+H <= ld_wt <= 160");
    if ((gates != nullptr) != (c_all != nullptr) || (gates != nullptr) != (xh_t != nullptr))
        return fail(PSB_ERR_ARG, "psb_lstm_fwd: gates, c_all and xh_t must be given together");
    if (xh_t && (ld_xht % 8)) return fail(PSB_ERR_ARG, "psb_lstm_fwd: saving needs ld_xht%%8==0");
    if (gates && ((reinterpret_cast<uintptr_t>(gates) | reinterpret_cast<uintptr_t>(c_all)) & 15))
        return fail(PSB_ERR_ARG, "psb_lstm_fwd: gates and c_all must be 16-byte aligned");
    LstmFwdArgs a;
    a.x = x; a.wt = static_cast<const __nv_bfloat16*>(wt); a.ld_wt = ld_wt; a.bias = bias;
    a.forget_bias = forget_bias; a.B = B; a.T = T; a.I = I;
    a.h_last = static_cast<__nv_bfloat16*>(h_last); a.ld_h = ld_h;
    a.h_last_t = static_cast<__nv_bfloat16*>(h_last_t); a.ld_ht = ld_ht;
    a.gates = static_cast<uint4*>(gates); a.c_all = reinterpret_cast<float4*>(c_all);
    a.xh_t = static_cast<__nv_bfloat16*>(xh_t); a.ld_xht = ld_xht;
    const size_t smem = kFwdBSmem + 2 * kFwdStage + 2 * kFwdXStage + 1024;
    static bool attr = false;
    if (!attr) {
        PSB_CUDA_OK(cudaFuncSetAttribute(lstm_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    const int ctas = (B + kNB - 1) / kNB;
    if (xh_t != nullptr && ld_xht < ctas * T * kNB)
        return fail(PSB_ERR_ARG, "psb_lstm_fwd: xh_t needs %d columns (ceil(B/16)*T*16)", ctas * T * kNB);
    lstm_fwd_kernel<<<ctas, kFwdThreads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    return after_launch("lstm_fwd");
}

static int lstm_bwd_launch(const void* w, int ld_w, const float* dh_last, int ld_dh, const psb_lstm_head_t* head,
                           const void* gates, const float* c_all, int B, int T, int I, int H, void* dg_t,
                           int ld_dgt, float* dbias, float* dbias_scratch, void* stream) {
    using namespace psb;
    if (H != kH) return fail(PSB_ERR_ARG, "psb_lstm_bwd: this kernel is built for H=128 (got %d)", H);
    if (!w || (!dh_last && !head) || !gates || !c_all || B < 1 || T < 1) return fail(PSB_ERR_ARG, "psb_lstm_bwd: bad argument");
    if (head != nullptr) {
        if (!head->h_last || !head->w_out || !head->b_out || !head->labels || !head->scratch || !dbias_scratch)
            return fail(PSB_ERR_ARG, "psb_lstm_bwd_head: h_last, w_out, b_out, labels, scratch and dbias_scratch are required");
        if (head->C < 1 || head->C > kHeadC || head->ld_wo < head->C || head->ld_h < kH)
            return fail(PSB_ERR_ARG, "psb_lstm_bwd_head: 1 <= C <= 16, ld_wo >= C, ld_h >= H");
        if ((head->ld_h % 8) || (reinterpret_cast<uintptr_t>(head->h_last) & 15))
            return fail(PSB_ERR_ARG, "psb_lstm_bwd_head: h_last must be 16-byte aligned with ld_h%%8==0");
        if (head->zero_ptr && ((reinterpret_cast<uintptr_t>(head->zero_ptr) | head->zero_bytes) & 15))
            return fail(PSB_ERR_ARG, "psb_lstm_bwd_head: zero_ptr and zero_bytes must be multiples of 16");
    }
    if (ld_w % 8 || ld_w < 4 * H) return fail(PSB_ERR_ARG, "psb_lstm_bwd: ld_w must be a multiple of 8 covering 4H");
    if (dg_t && (ld_dgt % 8 || ld_dgt < ((B + kNB - 1) / kNB) * T * kNB))
        return fail(PSB_ERR_ARG, "psb_lstm_bwd: dg_t needs ld_dgt%%8==0 and ceil(B/16)*T*16 columns");
    if (dbias && !dbias_scratch) return fail(PSB_ERR_ARG, "psb_lstm_bwd: dbias needs dbias_scratch (ceil(B/16)*4H+1 zeroed floats)");
    LstmBwdArgs a;
    a.w = static_cast<const __nv_bfloat16*>(w); a.ld_w = ld_w; a.dh_last = dh_last; a.ld_dh = ld_dh;
    a.gates = static_cast<const uint4*>(gates); a.c_all = reinterpret_cast<const float4*>(c_all);
    a.B = B; a.T = T; a.I = I;
    a.dg_t = static_cast<__nv_bfloat16*>(dg_t); a.ld_dgt = ld_dgt; a.dbias = dbias; a.dbias_part = dbias_scratch;
    if (head != nullptr) a.head = *head;
    else memset(&a.head, 0, sizeof(a.head));
    const size_t smem = kBwdBSmem + 2 * kBwdStage + 1024;
    static bool attr = false;
    if (!attr) {
        PSB_CUDA_OK(cudaFuncSetAttribute(lstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((B + kNB - 1) / kNB);
    cfg.blockDim = dim3(kLstmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute pdl[1];
    pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pdl[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = pdl;
    cfg.numAttrs = 1;
    PSB_CUDA_OK(cudaLaunchKernelEx(&cfg, lstm_bwd_kernel, a));
    return after_launch("lstm_bwd");
}

extern "C" int psb_lstm_bwd(const void* w, int ld_w, const float* dh_last, int ld_dh, const void* gates,
                            const float* c_all, int B, int T, int I, int H, void* dg_t, int ld_dgt,
                            float* dbias, float* dbias_scratch, void* stream) {
    return lstm_bwd_launch(w, ld_w, dh_last, ld_dh, nullptr, gates, c_all, B, T, I, H, dg_t, ld_dgt, dbias,
                           dbias_scratch, stream);
}

extern "C" size_t psb_lstm_head_scratch_floats(int B) {
    return static_cast<size_t>((B + psb::kNB - 1) / psb::kNB) * psb::kHeadPart;
}

extern "C" int psb_lstm_bwd_head(const void* w, int ld_w, const psb_lstm_head_t* head, const void* gates,
                                 const float* c_all, int B, int T, int I, int H, void* dg_t, int ld_dgt,
                                 float* dbias, float* dbias_scratch, void* stream) {
    if (!head) return psb::fail(psb::PSB_ERR_ARG, "psb_lstm_bwd_head: head is required");
    return lstm_bwd_launch(w, ld_w, nullptr, 0, head, gates, c_all, B, T, I, H, dg_t, ld_dgt, dbias,
                           dbias_scratch, stream);
}
