// Fused loss + loss-gradient kernels and the input cast/transpose kernel.
//
// Each loss kernel reads the fp32 output of the model's last GEMM, adds its share of the
// scalar loss with one atomicAdd per block, and writes the gradient w.r.t. the last
// pre-activation in the three forms the backward GEMMs want: bf16 row-major [M, ld]
// (dgrad A operand), bf16 transposed [N, ld_t] (wgrad B operand) and fp32 column sums
// (bias gradient).
//
// ref: encoder:131-135 (MSE on sigmoid outputs), gan:142-143 (log-losses on D's sigmoid),
//      lstm:92-94,99-100 (softmax cross-entropy, accuracy).
#include "../../include/psb200.h"
#include "host_common.h"
#include "ptx.cuh"

namespace psb {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float softplus(float x) {
    // log(1 + e^x), stable for both signs
    return fmaxf(x, 0.0f) + log1pf(__expf(-fabsf(x)));
}

// ---------------------------------------------------------------- cast + pad + transpose
// in fp32 [M, N] (ld_in) -> out bf16 [M, ld_out] and/or out_t bf16 [N, ld_t]; 32x32 tiles.
__global__ void __launch_bounds__(256) cast_transpose_kernel(const float* __restrict__ in, int ld_in, int M, int N,
                                                             __nv_bfloat16* out, int ld_out,
                                                             __nv_bfloat16* out_t, int ld_t,
                                                             float* clear, int n_clear) {
    __shared__ float tile[32][33];
    const int n0 = blockIdx.x * 32, m0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
    // first kernel of a step: also resets the accumulators later kernels add into (loss scalars)
    if (clear != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && static_cast<int>(threadIdx.x) < n_clear)
        clear[threadIdx.x] = 0.0f;
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int m = m0 + r, n = n0 + tx;
        float v = (m < M && n < N) ? in[static_cast<size_t>(m) * ld_in + n] : 0.0f;
        tile[r][tx] = v;
        if (out != nullptr && m < M && n < ld_out) out[static_cast<size_t>(m) * ld_out + n] = __float2bfloat16_rn(v);
    }
    if (out_t == nullptr) return;
    __syncthreads();
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int n = n0 + r, m = m0 + tx;
        if (n < N && m < M) out_t[static_cast<size_t>(n) * ld_t + m] = __float2bfloat16_rn(tile[tx][r]);
    }
}

// ---------------------------------------------------------------- MSE on sigmoid outputs
// y fp32 [M, ld_y] = sigmoid output, target fp32 [M, ld_x]:
//   loss += sum((y-x)^2) / (M*N);  dz = 2 (y-x)/(M*N) * y (1-y)
// One block owns a 32-column strip for ALL rows, so the bias gradient (column sums) is
// accumulated in a fixed order without atomics (deterministic).
__global__ void __launch_bounds__(256) mse_sigmoid_grad_kernel(const float* __restrict__ y, int ld_y,
                                                               const float* __restrict__ x, int ld_x, int M, int N,
                                                               float* loss, __nv_bfloat16* dz, int ld_dz,
                                                               __nv_bfloat16* dz_t, int ld_dzt, float* colsum) {
    __shared__ float tile[32][33];
    __shared__ float red[8];
    const int n0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const float inv = 1.0f / (static_cast<float>(M) * static_cast<float>(N));
    float part = 0.0f, csum = 0.0f;
    // software pipeline: the loads of row-tile k+1 are in flight while tile k is processed
    float yn[4], xn[4];
    const int n = n0 + tx;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int m = ty + 8 * q;
        const bool ok = m < M && n < N;
        yn[q] = ok ? y[static_cast<size_t>(m) * ld_y + n] : 0.0f;
        xn[q] = ok ? x[static_cast<size_t>(m) * ld_x + n] : 0.0f;
    }
    for (int m0 = 0; m0 < M; m0 += 32) {
        float yc[4], xc[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) { yc[q] = yn[q]; xc[q] = xn[q]; }
        if (m0 + 32 < M) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int m = m0 + 32 + ty + 8 * q;
                const bool ok = m < M && n < N;
                yn[q] = ok ? y[static_cast<size_t>(m) * ld_y + n] : 0.0f;
                xn[q] = ok ? x[static_cast<size_t>(m) * ld_x + n] : 0.0f;
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int r = ty + 8 * q, m = m0 + r;
            float g = 0.0f;
            if (m < M && n < N) {
                const float d = yc[q] - xc[q];
                part += d * d;
                g = 2.0f * d * inv * yc[q] * (1.0f - yc[q]);
            }
            tile[r][tx] = g;
            if (dz != nullptr && m < M && n < ld_dz) dz[static_cast<size_t>(m) * ld_dz + n] = __float2bfloat16_rn(g);
        }
        __syncthreads();
#pragma unroll
        for (int r = ty; r < 32; r += 8) {
            const int nn = n0 + r, m = m0 + tx;
            if (dz_t != nullptr && nn < N && m < M) dz_t[static_cast<size_t>(nn) * ld_dzt + m] = __float2bfloat16_rn(tile[tx][r]);
        }
        if (ty == 0) {
#pragma unroll
            for (int r = 0; r < 32; ++r) csum += tile[r][tx];
        }
        __syncthreads();
    }
    if (colsum != nullptr && ty == 0 && n0 + tx < N) colsum[n0 + tx] += csum;
    part = warp_sum(part);
    if (tx == 0) red[ty] = part;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += red[i];
        atomicAdd(loss, s * inv);
    }
}

// ---------------------------------------------------------------- GAN log-losses from logits
// logit fp32 [2B] (stride ld): rows 0..B-1 real, B..2B-1 fake.
//   gen_loss  = mean softplus(-l_fake)          d/dl_fake = -(1-sigmoid(l_fake))/B
//   disc_loss = mean softplus(-l_real)+softplus(l_fake)
//               d/dl_real = -(1-sigmoid(l_real))/B   d/dl_fake = sigmoid(l_fake)/B
// outputs: losses[0]=gen, losses[1]=disc (overwritten); d_disc bf16 [2B, ld_d] col 0 and its
// transpose [1, 2B]; d_gen bf16 [B, ld_g] col 0; db_disc_out (sum of d_disc) overwritten.
__global__ void __launch_bounds__(256) gan_loss_grad_kernel(const float* __restrict__ logit, int ld, int B,
                                                            float* losses, __nv_bfloat16* d_disc, int ld_d,
                                                            __nv_bfloat16* d_disc_t, __nv_bfloat16* d_gen, int ld_g,
                                                            float* db_disc_out) {
    __shared__ float red[3][8];
    const float invB = 1.0f / static_cast<float>(B);
    float gl = 0.0f, dl = 0.0f, dbs = 0.0f;
    for (int r = threadIdx.x; r < 2 * B; r += blockDim.x) {
        const float l = logit[static_cast<size_t>(r) * ld];
        const float s = fast_sigmoid(l);
        float gd;
        if (r < B) {
            dl += softplus(-l);
            gd = -(1.0f - s) * invB;
        } else {
            gl += softplus(-l);
            dl += softplus(l);
            gd = s * invB;
            const __nv_bfloat16 gg = __float2bfloat16_rn(-(1.0f - s) * invB);
            d_gen[static_cast<size_t>(r - B) * ld_g] = gg;
            for (int c = 1; c < ld_g && c < 8; ++c) d_gen[static_cast<size_t>(r - B) * ld_g + c] = __float2bfloat16_rn(0.0f);
        }
        const __nv_bfloat16 gb = __float2bfloat16_rn(gd);
        d_disc[static_cast<size_t>(r) * ld_d] = gb;
        for (int c = 1; c < ld_d && c < 8; ++c) d_disc[static_cast<size_t>(r) * ld_d + c] = __float2bfloat16_rn(0.0f);
        d_disc_t[r] = gb;
        dbs += gd;
    }
    gl = warp_sum(gl); dl = warp_sum(dl); dbs = warp_sum(dbs);
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { red[0][w] = gl; red[1][w] = dl; red[2][w] = dbs; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f, c = 0.f;
        for (int i = 0; i < static_cast<int>(blockDim.x >> 5); ++i) { a += red[0][i]; b += red[1][i]; c += red[2][i]; }
        losses[0] = a * invB;
        losses[1] = b * invB;
        if (db_disc_out != nullptr) *db_disc_out = c;
    }
}

// ---------------------------------------------------------------- softmax cross-entropy
// logits fp32 [B, C] (ld), labels fp32 [B, C] (ld_l) one-hot/soft.  One thread per row, C <= 32.
//   loss (overwritten) = mean_b( -sum_c y log softmax )   dz = (softmax * sum_c y - y)/B
//   correct (optional, overwritten) = #rows with argmax(logits) == argmax(labels)
__global__ void __launch_bounds__(1024) softmax_xent_grad_kernel(const float* __restrict__ logits, int ld,
                                                                 const float* __restrict__ labels, int ld_l,
                                                                 int B, int C, float* loss, __nv_bfloat16* dz,
                                                                 int ld_dz, __nv_bfloat16* dz_t, int ld_dzt,
                                                                 float* colsum, float* correct) {
    __shared__ float s_part[32][34];     // [warp][0..31 column sums, 32 loss, 33 correct]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const float invB = 1.0f / static_cast<float>(B);
    float col[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) col[c] = 0.0f;
    float l_acc = 0.0f, corr = 0.0f;
    for (int r = threadIdx.x; r < B; r += blockDim.x) {
        float z[32], y[32];
        float mx = -INFINITY, ysum = 0.0f, ymax = -INFINITY;
        int am = 0, ay = 0;
#pragma unroll
        for (int c = 0; c < 32; ++c) {
            if (c < C) {
                z[c] = logits[static_cast<size_t>(r) * ld + c];
                y[c] = labels[static_cast<size_t>(r) * ld_l + c];
                if (z[c] > mx) { mx = z[c]; am = c; }
                if (y[c] > ymax) { ymax = y[c]; ay = c; }
                ysum += y[c];
            }
        }
        float se = 0.0f;
#pragma unroll
        for (int c = 0; c < 32; ++c)
            if (c < C) se += __expf(z[c] - mx);
        const float lse = mx + __logf(se);
#pragma unroll
        for (int c = 0; c < 32; ++c) {
            if (c < C) {
                l_acc -= y[c] * (z[c] - lse);
                const float g = (__expf(z[c] - lse) * ysum - y[c]) * invB;
                col[c] += g;
                if (dz != nullptr) dz[static_cast<size_t>(r) * ld_dz + c] = __float2bfloat16_rn(g);
                if (dz_t != nullptr) dz_t[static_cast<size_t>(c) * ld_dzt + r] = __float2bfloat16_rn(g);
            }
        }
        if (dz != nullptr)
            for (int c = C; c < ld_dz && c < ((C + 7) & ~7); ++c) dz[static_cast<size_t>(r) * ld_dz + c] = __float2bfloat16_rn(0.0f);
        if (am == ay) corr += 1.0f;
    }
    // warp reductions, then one pass over the per-warp partials
#pragma unroll
    for (int c = 0; c < 32; ++c) {
        if (c < C) {
            const float s = warp_sum(col[c]);
            if (lane == 0) s_part[warp][c] = s;
        }
    }
    l_acc = warp_sum(l_acc);
    corr = warp_sum(corr);
    if (lane == 0) { s_part[warp][32] = l_acc; s_part[warp][33] = corr; }
    __syncthreads();
    if (warp == 0) {
        for (int k = lane; k < 34; k += 32) {
            if (k < C || k >= 32) {
                float s = 0.0f;
                for (int w = 0; w < nwarps; ++w) s += s_part[w][k];
                if (k < C) { if (colsum != nullptr) atomicAdd(colsum + k, s); }
                else if (k == 32) *loss = s * invB;
                else if (correct != nullptr) *correct = s;
            }
        }
    }
}

}  // namespace psb

extern "C" int psb_cast_transpose(const float* in, int ld_in, int M, int N, void* out, int ld_out,
                                  void* out_t, int ld_t, void* stream) {
    using namespace psb;
    if (!in || M < 1 || N < 1 || (!out && !out_t)) return fail(PSB_ERR_ARG, "psb_cast_transpose: bad argument");
    const int ncols = out ? (ld_out > N ? ld_out : N) : N;
    dim3 grid((ncols + 31) / 32, (M + 31) / 32);
    cast_transpose_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        in, ld_in, M, N, static_cast<__nv_bfloat16*>(out), ld_out, static_cast<__nv_bfloat16*>(out_t), ld_t, nullptr, 0);
    return after_launch("cast_transpose");
}

extern "C" int psb_cast_transpose_clear(const float* in, int ld_in, int M, int N, void* out, int ld_out,
                                        void* out_t, int ld_t, float* clear, int n_clear, void* stream) {
    using namespace psb;
    if (!in || M < 1 || N < 1 || (!out && !out_t)) return fail(PSB_ERR_ARG, "psb_cast_transpose_clear: bad argument");
    if (n_clear < 0 || n_clear > 256 || (n_clear > 0 && !clear))
        return fail(PSB_ERR_ARG, "psb_cast_transpose_clear: clear at most 256 floats");
    const int ncols = out ? (ld_out > N ? ld_out : N) : N;
    dim3 grid((ncols + 31) / 32, (M + 31) / 32);
    cast_transpose_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        in, ld_in, M, N, static_cast<__nv_bfloat16*>(out), ld_out, static_cast<__nv_bfloat16*>(out_t), ld_t, clear, n_clear);
    return after_launch("cast_transpose_clear");
}

extern "C" int psb_mse_sigmoid_grad(const float* y, int ld_y, const float* x, int ld_x, int M, int N,
                                    float* loss, void* dz, int ld_dz, void* dz_t, int ld_dzt,
                                    float* colsum, void* stream) {
    using namespace psb;
    if (!y || !x || !loss || M < 1 || N < 1) return fail(PSB_ERR_ARG, "psb_mse_sigmoid_grad: bad argument");
    const int ncols = dz ? (ld_dz > N ? ld_dz : N) : N;
    dim3 grid((ncols + 31) / 32, 1);
    mse_sigmoid_grad_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        y, ld_y, x, ld_x, M, N, loss, static_cast<__nv_bfloat16*>(dz), ld_dz,
        static_cast<__nv_bfloat16*>(dz_t), ld_dzt, colsum);
    return after_launch("mse_sigmoid_grad");
}

extern "C" int psb_gan_loss_grad(const float* logit, int ld, int B, float* losses, void* d_disc, int ld_d,
                                 void* d_disc_t, void* d_gen, int ld_g, float* db_disc_out, void* stream) {
    using namespace psb;
    if (!logit || !losses || !d_disc || !d_disc_t || !d_gen || B < 1)
        return fail(PSB_ERR_ARG, "psb_gan_loss_grad: bad argument");
    gan_loss_grad_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        logit, ld, B, losses, static_cast<__nv_bfloat16*>(d_disc), ld_d, static_cast<__nv_bfloat16*>(d_disc_t),
        static_cast<__nv_bfloat16*>(d_gen), ld_g, db_disc_out);
    return after_launch("gan_loss_grad");
}

extern "C" int psb_softmax_xent_grad(const float* logits, int ld, const float* labels, int ld_l, int B, int C,
                                     float* loss, void* dz, int ld_dz, void* dz_t, int ld_dzt, float* colsum,
                                     float* correct, void* stream) {
    using namespace psb;
    if (!logits || !labels || !loss || B < 1 || C < 1 || C > 32)
        return fail(PSB_ERR_ARG, "psb_softmax_xent_grad: need 1 <= C <= 32 and non-null logits/labels/loss");
    int threads = ((B + 31) / 32) * 32;
    if (threads > 1024) threads = 1024;
    softmax_xent_grad_kernel<<<1, threads, 0, static_cast<cudaStream_t>(stream)>>>(
        logits, ld, labels, ld_l, B, C, loss, static_cast<__nv_bfloat16*>(dz), ld_dz,
        static_cast<__nv_bfloat16*>(dz_t), ld_dzt, colsum, correct);
    return after_launch("softmax_xent_grad");
}
