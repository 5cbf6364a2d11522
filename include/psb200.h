/* psb200 — C ABI of the sm_100a kernel library (libpsb200_kernels.so) and of the native
 * parameter-server runtime (libpsb200_ps.so).
 *
 * Conventions
 *   - every function returns 0 on success and a negative code on failure; the text of
 *     the last failure on the calling thread is psb_last_error();
 *   - device pointers are plain `void*` / typed pointers, `stream` is a cudaStream_t
 *     passed as `void*` (NULL = legacy default stream);
 *   - bf16 matrices are row-major with a leading dimension in ELEMENTS that must be a
 *     multiple of 8 (16-byte rows, the TMA stride rule);
 *   - nothing here allocates device memory unless its comment says so.
 *
 * The "ref:" notes name the reference call site each entry point replaces
 * (encoder: = encoder/distributed_encoder.py, gan: = gan/distributed_gan.py,
 *  lstm: = lstm/distributed_lstm.py of Xingskcs/Distributed-TensorFlow-Examples).
 */
#ifndef PSB200_H_
#define PSB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ library state */
const char* psb_last_error(void);
/* Number of kernel launches issued through this library since load (the bench's
 * gpu_launches evidence). */
uint64_t psb_launch_count(void);
int psb_version(void);

/* ------------------------------------------------------------------ dense GEMM + epilogue
 * acc[m,n] = alpha * sum_k A[m,k] * B[n,k]          (A: [M,K] bf16, B: [N,K] bf16, both K-major)
 * then, in this order, per element:
 *   + bias[n]                       (bias != NULL)
 *   act: 0 none | 1 sigmoid | 2 relu
 *   * act'(yprev[m,n])              (yprev != NULL; dact 1: y(1-y)  2: y>0 ? 1 : 0)
 * and any subset of the outputs:
 *   out      bf16 [M, ld_out]   (pad columns N..round_up(N,8) are written as zero)
 *   out_t    bf16 [N, ld_out_t] (the transpose, for the weight-gradient GEMM)
 *   out_f32  fp32 [M, ld_f32]
 *   colsum   fp32 [N]           (atomicAdd of column sums: the bias gradient)
 * tcgen05.mma (M=128 x N=BN tiles, fp32 accumulators in TMEM), operands staged by TMA
 * with 128-byte swizzle through a multi-stage mbarrier ring.
 * ref: tf.matmul + tf.add + sigmoid/relu at encoder:69-84, gan:77-92, lstm:53, and their
 *      autodiff backward (dX = dZ W^T, dW = X^T dZ, db = sum dZ) from minimize():
 *      encoder:137, gan:158-159, lstm:97. */
typedef struct psb_epilogue {
    float alpha;
    int act;
    int dact;
    const float* bias;
    const void* yprev; /* bf16 [M, ld_yprev] */
    int ld_yprev;
    void* out;
    int ld_out;
    void* out_t;
    int ld_out_t;
    float* out_f32;
    int ld_f32;
    float* colsum;
} psb_epilogue_t;

int psb_gemm_bf16(const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                  const psb_epilogue_t* epi, int block_n /*0 = auto*/, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PSB200_H_ */
