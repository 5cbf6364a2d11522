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
    int split_k;       /* > 1: split the reduction over this many CTAs; requires out_f32 only */
    long long split_stride; /* elements between the split_k partial outputs (split z writes out_f32 +
                          z*split_stride, the consumer adds them: deterministic); 0 = atomically add all
                          partials into out_f32, which the caller has zeroed */
    int b_static;      /* 1: B (e.g. a weight copy) was not produced by the kernel launched just before this
                          one on the stream, so its first pipeline stages may be fetched while that kernel is
                          still finishing (programmatic dependent launch); 0 when in doubt */
    /* Parameter pull with per-block readiness.  B is a compute copy of a variable that the owners'
     * push kernels refresh row block by row block (psb_push_seg_t.ready_idx); `ready` points at this
     * rank's flag words for that variable, one per row block.  The TMA producer requests a tile only
     * after the flag of the row block it belongs to has reached *ready_step, visiting the k-blocks from
     * `ready_first` on (fixed order per rank: results stay bit-reproducible), so the tensor cores start
     * on the blocks that have arrived while the rest is still crossing NVLink.  All NULL/0 = off. */
    const uint32_t* ready;
    const uint32_t* ready_step;
    int ready_rows;    /* variable rows per block */
    int ready_dim;     /* 0: the row blocks partition K of this GEMM (B = W^T copy); 1: they partition N (B = W copy) */
    int ready_nblk;
    int ready_first;
    const uint32_t* bias_ready;   /* flag word of the bias variable, waited for by the epilogue warps */
    uint32_t* ready_status;       /* receives PSB_PUSH_ST_TIMEOUT_PARAMS if a wait ends on ready_timeout_ns */
    uint64_t ready_timeout_ns;
    int pull_mode;     /* psb_gemm_bf16_pull only: 0 = TMA on the peer-mapped address, 1 = 128-bit ld.global */
} psb_epilogue_t;

int psb_gemm_bf16(const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                  const psb_epilogue_t* epi, int block_n /*0 = auto*/, void* stream);

/* Parameter pull fused into the consuming GEMM.  Same contract as psb_gemm_bf16, but B (the
 * weight copy, [N, ldb] bf16, pad columns zero) may live in ANOTHER GPU's memory (a CUDA-IPC peer
 * mapping): its tiles are fetched over NVLink by TMA -- a tensor map over the peer address --
 * straight into the swizzled shared-memory stages while the tensor cores work on the previous
 * stages, optionally gated tile by tile on the owners' readiness flags (epi->ready ...).
 * epi->pull_mode = 1 selects the older producer-warp variant (128-bit ld.global into the swizzled
 * stage).  A arrives by TMA from local memory; `epi->bias` may be a peer pointer too.
 * ref: the per-step ps->worker variable reads TF inserts for encoder:69-83, gan:77-90, lstm:53. */
int psb_gemm_bf16_pull(const void* A, int lda, const void* B_peer, int ldb, int M, int N, int K,
                       const psb_epilogue_t* epi, int block_n /*0 = auto*/, void* stream);


/* ------------------------------------------------------------------ persistent LSTM (H = 128)
 * Forward: BasicLSTMCell unrolled over T steps from a zero state, one CTA per 16 batch rows,
 * cell kernel resident in shared memory, gate GEMM on tcgen05 with the batch slice as the MMA
 * N dimension, sigmoid/tanh pointwise on the TMEM accumulator, h fed back through shared
 * memory.  `wt` is the cell kernel TRANSPOSED, bf16 [4H, ld_wt] (gate order i,j,f,o); it may
 * point into a peer GPU's memory (the parameter pull is then part of this kernel).
 * Optional outputs: h_last bf16 [B, ld_h], h_last_t bf16 [H, ld_ht]; for training also the
 * two saved-activation blobs `gates` / `c_all` (opaque, 16-byte aligned, sized by
 * psb_lstm_saved_bytes, consumed by psb_lstm_bwd) and xh_t bf16 [I+H, ld_xht], the
 * weight-gradient GEMM's A operand: column (cta*T + t)*16 + b16 holds [x_t, h_{t-1}] of batch
 * row cta*16 + b16, so it needs ceil(B/16)*T*16 columns (psb_lstm_bwd writes dg_t in the same
 * column order; four consecutive timesteps of a row are one 128-byte line).
 * ref: rnn.BasicLSTMCell(128, forget_bias=1.0) + rnn.static_rnn, lstm:50-52. */
/* Bytes of the saved-activation blobs: which = 0 activated gates (bf16), which = 1 cell state
 * (fp32); private layouts documented in csrc/lstm.cu, one block per 16 batch rows. */
size_t psb_lstm_saved_bytes(int B, int T, int H, int which);
int psb_lstm_fwd(const float* x, const void* wt, int ld_wt, const float* bias, float forget_bias,
                 int B, int T, int I, int H, void* h_last, int ld_h, void* h_last_t, int ld_ht,
                 void* gates, float* c_all, void* xh_t, int ld_xht, void* stream);

/* BPTT for psb_lstm_fwd when only the last hidden state feeds the loss.  `w` is the cell kernel
 * in its natural layout, bf16 [I+H, ld_w]; dh_last fp32 [B, ld_dh].  Writes the gate gradients
 * transposed, dg_t bf16 [4H, ld_dgt] (column t*B+b), and adds the bias gradient into dbias fp32
 * [4H] (per-CTA partials in dbias_scratch, ceil(B/16)*4H + 1 floats, zero before first use,
 * summed in CTA order by the last CTA: bit-reproducible).  The kernel gradient is then one
 * GEMM: dK = xh_t * dg_t^T (psb_gemm_bf16).
 * ref: the gradient graph optimizer.minimize builds for static_rnn, lstm:97. */
int psb_lstm_bwd(const void* w, int ld_w, const float* dh_last, int ld_dh, const void* gates,
                 const float* c_all, int B, int T, int I, int H, void* dg_t, int ld_dgt,
                 float* dbias, float* dbias_scratch, void* stream);

/* psb_lstm_bwd with the classifier head of the LSTM model fused into its prologue: instead of
 * reading dh_last, each CTA computes for its 16 batch rows logits = h_last W_out + b_out, the
 * softmax cross-entropy against `labels`, dlogits, and dh_last = dlogits W_out^T on the spot, and
 * leaves partial dW_out / db_out / loss / correct sums in `scratch` that the last CTA adds in CTA
 * order (bit-reproducible).  d_w_out and d_b_out are accumulated (+=), loss (mean over B) and
 * correct (count of argmax matches) are overwritten; logits is optional.  C <= 16.
 * `scratch` holds psb_lstm_head_scratch_floats(B) floats; dbias_scratch is required (it carries
 * the completion counter).
 * ref: outputs[-1] W + b, softmax_cross_entropy_with_logits, accuracy and their gradients,
 * lstm:53,92-100. */
typedef struct {
    const void* h_last; int ld_h;        /* bf16 [B, ld_h], written by psb_lstm_fwd */
    const void* w_out; int ld_wo;        /* bf16 [H, ld_wo] */
    const float* b_out;                  /* [C] */
    const float* labels; int ld_l;       /* fp32 [B, ld_l] one-hot or soft */
    int C;
    float* logits; int ld_logits;        /* optional fp32 [B, ld_logits] */
    float* loss; float* correct;         /* optional scalars */
    float* d_w_out; int ld_dwo;          /* fp32 [H, ld_dwo] += */
    float* d_b_out;                      /* fp32 [C] += */
    float* scratch;
    int accumulate;                      /* 0: d_w_out, d_b_out and dbias are overwritten instead of += */
    void* zero_ptr; size_t zero_bytes;   /* optional buffer cleared in the prologue (16-byte multiples),
                                            e.g. the gradient a following split-K GEMM accumulates into */
} psb_lstm_head_t;
size_t psb_lstm_head_scratch_floats(int B);
int psb_lstm_bwd_head(const void* w, int ld_w, const psb_lstm_head_t* head, const void* gates,
                      const float* c_all, int B, int T, int I, int H, void* dg_t, int ld_dgt,
                      float* dbias, float* dbias_scratch, void* stream);

/* ------------------------------------------------------------------ large-H LSTM path
 * Pointwise halves of a BasicLSTMCell step for hidden sizes whose kernel cannot be resident on
 * one SM (the scaled 2-layer H=4096 configuration): the gate GEMMs run through psb_gemm_bf16.
 * Forward: pre = gx + gh + bias (fp32 [B,4H], gate order i,j,f,o), c fp32 [B,H] updated in place,
 * h_out bf16 [B, ld_h]; optional saves for backward (activated gates bf16 [B,4H], c fp32 [B,H]).
 * Backward: dh = sum of the n_parts partial dh_rec buffers (part_stride elements apart: the
 * split-K outputs of the dh GEMM) + dy (either may be NULL), dc fp32 [B,H] updated in place, dg bf16
 * [B, ld_dg] = gate gradients, dbias fp32 [4H] += column sums (atomic).
 * ref: rnn.BasicLSTMCell / static_rnn, lstm:50-52, scaled per BASELINE.json config #4. */
/* One forward timestep of the large-H LSTM in ONE kernel: gates = h_prev Wh (tcgen05 GEMM, every
 * CTA gathers the i/j/f/o rows of its 32 hidden units with four 32-row TMA boxes) + gx + bias,
 * then the cell pointwise in the epilogue.  h_prev bf16 [B, ld_hp], wh_t bf16 [4H, ld_w] (the
 * recurrent half of kernel^T), gx fp32 [B, ld_gx], c fp32 [B,H] in/out, h_out bf16 [B, ld_h];
 * optional saves as in psb_lstm_cell_fwd.  H must be a multiple of 32. */
int psb_lstm_step_fwd(const void* h_prev, int ld_hp, const void* wh_t, int ld_w, int B, int H,
                      const float* gx, int ld_gx, const float* bias, float forget_bias, float* c,
                      void* h_out, int ld_h, void* gates_save, float* c_save, void* stream);
int psb_lstm_cell_fwd(const float* gx, int ld_gx, const float* gh, int ld_gh, const float* bias,
                      float forget_bias, float* c, void* h_out, int ld_h, void* gates_save,
                      float* c_save, int B, int H, void* stream);
int psb_lstm_cell_bwd(const float* dh_rec, int ld_dh, int n_parts, long long part_stride, const float* dy, int ld_dy, const void* gates,
                      const float* c_new, const float* c_prev, float* dc, void* dg, int ld_dg,
                      float* dbias, int B, int H, void* stream);
/* out[c][r] = in[r][c], bf16, leading dimensions in elements. */
int psb_transpose_bf16(const void* in, int ld_in, int R, int C, void* out, int ld_out, void* stream);

/* ------------------------------------------------------------------ input cast, losses
 * fp32 [M,N] -> bf16 [M, ld_out] (zero pad columns) and/or its transpose bf16 [N, ld_t].
 * ref: feed_dict of the float32 placeholders, encoder:165, gan:192-194, lstm:128. */
int psb_cast_transpose(const float* in, int ld_in, int M, int N, void* out, int ld_out,
                       void* out_t, int ld_t, void* stream);
/* The same, and the kernel also zeroes clear[0..n_clear) (n_clear <= 256): as the first kernel of
 * a step it resets the loss accumulators the later kernels add into, so a step needs no fill. */
int psb_cast_transpose_clear(const float* in, int ld_in, int M, int N, void* out, int ld_out,
                             void* out_t, int ld_t, float* clear, int n_clear, void* stream);

/* MSE between the sigmoid output y (fp32 [M, ld_y]) and the target x (fp32 [M, ld_x]):
 * *loss += mean((x-y)^2) (caller zeroes it), dz = dLoss/d(pre-activation) as bf16 [M, ld_dz],
 * bf16 transposed [N, ld_dzt] and fp32 column sums (+=).  ref: encoder:131-135. */
int psb_mse_sigmoid_grad(const float* y, int ld_y, const float* x, int ld_x, int M, int N,
                         float* loss, void* dz, int ld_dz, void* dz_t, int ld_dzt,
                         float* colsum, void* stream);

/* GAN log-losses from discriminator logits (fp32, 2B rows with stride ld: real then fake).
 * losses[0] = -mean(log D(G(z))), losses[1] = -mean(log D(x) + log(1 - D(G(z)))) (stable
 * softplus form); d_disc bf16 [2B, ld_d] and d_disc_t bf16 [1, 2B] = d disc_loss / d logit;
 * d_gen bf16 [B, ld_g] = d gen_loss / d logit_fake; *db_disc_out = sum(d_disc).
 * ref: gan:142-143. */
int psb_gan_loss_grad(const float* logit, int ld, int B, float* losses, void* d_disc, int ld_d,
                      void* d_disc_t, void* d_gen, int ld_g, float* db_disc_out, void* stream);

/* Mean softmax cross-entropy of logits fp32 [B,C] against labels fp32 [B,C] (C <= 32):
 * *loss (overwritten), dz bf16 [B, ld_dz], dz_t bf16 [C, ld_dzt], colsum fp32 [C] (+=), and
 * optionally *correct = number of rows whose argmax matches.  ref: lstm:92-94,99-100. */
int psb_softmax_xent_grad(const float* logits, int ld, const float* labels, int ld_l, int B, int C,
                          float* loss, void* dz, int ld_dz, void* dz_t, int ld_dzt, float* colsum,
                          float* correct, void* stream);

/* ------------------------------------------------------------------ fused gradient push
 * One variable (or one contiguous run of it) owned by the calling rank. */
typedef struct psb_push_seg {
    uint64_t grad_offset;   /* element offset inside every rank's flat fp32 gradient buffer */
    uint32_t numel;
    uint32_t cols;          /* logical [numel/cols, cols] view used for the bf16 copies */
    int32_t opt_kind;       /* 0 SGD, 1 RMSProp, 2 Adam, -1 none (only refresh the copies) */
    float lr, h1, h2, eps;  /* h1 = decay | beta1, h2 = momentum | beta2 */
    float* master;          /* fp32 [numel] master weights (owner memory; peer pointer in async mode) */
    float* slot0;           /* rms | m */
    float* slot1;           /* momentum | v */
    const float* beta_pow;  /* Adam: {beta1^t, beta2^t} */
    int64_t w_off;          /* element offset of the natural bf16 copy [rows, ld_w] in the arena, -1 = none */
    int32_t ld_w;
    int64_t wt_off;         /* element offset of the transposed bf16 copy [cols, ld_wt], -1 = none */
    int32_t ld_wt;
    int64_t f_off;          /* element offset of an fp32 copy [numel] in the fp32 arena, -1 = none */
    int32_t home;           /* index into arena_ptrs that receives the copies when !bcast */
    int32_t bcast;          /* 1: write the copies into every rank's arena (push-mode all-gather) */
    uint32_t flags;         /* PSB_SEG_* */
    uint16_t nblocks;       /* blocks that share this segment's work (block k of them takes units k, k+nblocks,
                               ...); 0 = grid.x */
    uint16_t block_x;       /* this entry is for block block_x + blockIdx.x of those.  A caller may list a
                               segment once (block_x = 0) and launch grid.x = blocks_per_seg blocks per entry, or
                               list it nblocks times with block_x = 0..nblocks-1 and launch ONE block per entry:
                               segments of very different sizes then share one wave without idle blocks */
    uint32_t ready_slot;    /* 1 + index of this segment's readiness word, 0 = none.  When the LAST block of the
                               segment has finished, the step token is published in that word of every rank's
                               readiness array (desc->ready_ptrs): consumers that pull the copies (psb_gemm_bf16 /
                               _pull with epi->ready) start on this row block without waiting for the rest of
                               the push */
    uint32_t reserved_;
} psb_push_seg_t;
/* After reading them, overwrite this segment's gradients with zero in EVERY rank's gradient
 * buffer, so producers that accumulate (bias column sums, split-K partial sums) need no separate
 * clear before the next step. */
#define PSB_SEG_ZERO_GRAD 1u
/* Matrix segment processed in 128x128 tiles whose bf16 copies leave in 16-byte stores forming
 * 256-byte runs (what large copies written into peers' arenas over NVLink need); for matrices of
 * at least 128 x 128.  Without it: 32x32 tiles, one wave of blocks for small variables. */
#define PSB_SEG_BIG_TILES 2u

typedef struct psb_beta_update {
    float* beta_pow;        /* {beta1_power, beta2_power} */
    float beta1, beta2;
} psb_beta_update_t;

typedef struct psb_push_desc {
    const psb_push_seg_t* segs;      /* DEVICE array */
    int nseg;
    const float* const* grad_ptrs;   /* DEVICE array [world]: every rank's gradient buffer */
    void* const* arena_ptrs;         /* DEVICE array [world]: every rank's bf16 arena */
    float* const* farena_ptrs;       /* DEVICE array [world]: every rank's fp32 arena */
    uint32_t* const* pad_ptrs;       /* DEVICE array [world]: every rank's signal pad (>= 32 words) */
    int world, rank;
    int n_arenas;                    /* arenas written when a segment has bcast=1 (0 = world) */
    float inv_replicas;              /* 1 / replicas_to_aggregate */
    uint32_t step_id;                /* monotone token, first call = 1; 0 = use local_ctrl[2] + 1
                                        (lets a captured CUDA graph replay the launch) */
    uint32_t* local_ctrl;            /* DEVICE, PSB_PUSH_CTRL_WORDS zero-initialised words private to this rank:
                                        [2] last step token, [3] status (PSB_PUSH_ST_*, sticky), [4] abort request
                                        (a non-zero value written by the host ends every wait), [7] updates applied */
    psb_beta_update_t beta[2];
    int nbeta;
    int32_t* global_step;            /* incremented by step_delta by the last block (NULL = not here) */
    int step_delta;
    int async_sgd_atomic;
    int signal_params;               /* publish (and wait for) params-ready on the pads (when phases == 0) */
    int blocks_per_seg;              /* grid.x; a segment uses its first `nblocks` of them.  0 = default */
    uint64_t* prof;                  /* optional DEVICE [8] u64, zeroed by the caller: the kernel adds the
                                        nanoseconds it spent waiting for gradients [4], reducing+applying [5],
                                        waiting for parameters [6], and the call count [7] (globaltimer) */
    int phases;                      /* PSB_PUSH_* bits; 0 = the fused step: signal + wait for all gradients, apply,
                                        and (signal_params) signal + wait for all parameters */
    int replicas;                    /* R of N: aggregate the first `replicas` gradients to arrive, mean over those
                                        actually taken; gradients that arrive later are never read.  <= 0 or >= world: all */
    uint64_t timeout_ns;             /* bound on every cross-GPU wait (globaltimer); 0 = none.  On expiry the kernel
                                        applies nothing, sets local_ctrl[3] and returns */
    uint32_t* const* ready_ptrs;     /* DEVICE array [world]: every rank's readiness array (NULL = no per-segment flags) */
    uint32_t* seg_done;              /* DEVICE, zeroed: one finished-block counter per readiness word (local) */
    int big_tiles;                   /* 1: some segment has PSB_SEG_BIG_TILES (the launch then carries 34 KiB of dynamic
                                        shared memory; without it those segments take the small-tile path) */
} psb_push_desc_t;

#define PSB_PUSH_CTRL_WORDS 64
/* phases */
#define PSB_PUSH_SIGNAL_GRADS 1   /* publish "my gradients for this step are complete" to every pad */
#define PSB_PUSH_WAIT_GRADS 2     /* block until `replicas` ranks have published theirs */
#define PSB_PUSH_TRY_GRADS 4      /* do not block: if fewer have, apply nothing and return (ps service polling) */
#define PSB_PUSH_APPLY 8
#define PSB_PUSH_SIGNAL_PARAMS 16 /* publish "parameters of this step are in place" */
#define PSB_PUSH_WAIT_PARAMS 32   /* block until every owner has */
#define PSB_PUSH_EACH 64          /* async owner-side apply: one optimizer step per newly arrived gradient, in rank
                                     order, then acknowledge each sender (pad word 32 + owner) */
/* local_ctrl[3] */
#define PSB_PUSH_ST_OK 0
#define PSB_PUSH_ST_TIMEOUT_GRADS 1
#define PSB_PUSH_ST_TIMEOUT_PARAMS 2
#define PSB_PUSH_ST_ABORTED 3

/* Reduce-scatter + optimizer apply (+ all-gather of the bf16 copies) in one kernel over peer
 * memory; see csrc/push.cu.  Allocates nothing.
 * ref: minimize() -> Apply* on the ps + AssignAdd(global_step), encoder:137 gan:158-159 lstm:97. */
int psb_push_apply(const psb_push_desc_t* desc, void* stream);
/* psb_push_apply for callers that cannot allocate device memory (an optimizer hook running
 * inside another framework's memory pool, a plug-in whose host forbids allocation after
 * start-up, a latency path where cudaMalloc's implicit synchronisation is unacceptable).
 * This entry point has no counterpart in the reference; it exists only for those callers.
 * The kernel's control words, the segment table and the pointer tables live in one static
 * device array inside the library image (9472 bytes = 9.25 KiB per GPU, held until process
 * exit), so the tables are given as HOST arrays: host_segs[nseg], and host_grad_ptrs[world],
 * host_arena_ptrs[n_arenas], host_farena_ptrs[n_arenas], host_pad_ptrs[world] (may be NULL
 * when world == 1); they are copied with cudaMemcpyAsync on `stream`, so keep them valid and
 * unchanged until the stream has passed those copies.  Of `desc`, the fields segs, grad_ptrs,
 * arena_ptrs, farena_ptrs, pad_ptrs and local_ctrl are ignored; all others mean what they mean
 * for psb_push_apply.
 * Limits: nseg <= PSB_NOALLOC_MAX_SEGS, world and n_arenas <= PSB_NOALLOC_MAX_WORLD; beyond
 * them the call fails with an error text and launches nothing.  blocks_per_seg is capped at 64.
 * The first call on a device resolves the array's device address and caches it: make that
 * call before any stream capture.  Calls on one device share the array (the finished-block
 * counter is cleared per call; the step token, launch number and status word are kept), so
 * calls that overlap on the GPU race: serialise them.  Because the step token is kept for the
 * life of the process, explicit step ids (desc->step_id != 0) must keep increasing per device
 * across all users of this entry point, and one set of signal pads must be driven through
 * EITHER this entry point OR psb_push_apply, never alternately: their automatic step tokens are
 * separate (static array here, the caller's local_ctrl there) and would diverge.
 * Calls no cudaMalloc/cudaFree and no allocating library routine. */
#define PSB_NOALLOC_MAX_SEGS 64
#define PSB_NOALLOC_MAX_WORLD 16
int psb_push_apply_noalloc(const psb_push_seg_t* host_segs, int nseg,
                           const float* const* host_grad_ptrs, void* const* host_arena_ptrs,
                           float* const* host_farena_ptrs, uint32_t* const* host_pad_ptrs,
                           const psb_push_desc_t* desc, void* stream);
/* Worker half of the decoupled protocol (the owner half is psb_push_apply with PSB_PUSH_TRY_GRADS,
 * launched by the ps service).  worker_ctrl: PSB_PUSH_CTRL_WORDS zeroed device words.
 *   async_mode 0: publish gradients for token worker_ctrl[2] (first = 1), wait until every owner has
 *                 applied an aggregate >= token, adopt min(applied)+1 as the next token -- a worker
 *                 that fell behind skips ahead and its late gradient is dropped as stale;
 *   async_mode 1: wait == 0 publishes gradient number worker_ctrl[2]+1; wait == 1 blocks until every
 *                 owner has acknowledged gradient number worker_ctrl[2] (buffer free to overwrite).
 * ref: token_queue dequeue / accumulator apply_grad of SyncReplicasOptimizer; plain minimize(). */
int psb_push_commit(uint32_t* const* pad_ptrs, int world, int rank, uint32_t* worker_ctrl,
                    int async_mode, int wait, uint64_t timeout_ns, void* stream);
/* The ps role of this rank as a native service thread (csrc/ps_service.cu): a std::thread with its
 * own high-priority stream launches psb_push_apply(desc) over and over; `desc->phases` must be
 * the non-blocking form (PSB_PUSH_TRY_GRADS | PSB_PUSH_APPLY | PSB_PUSH_SIGNAL_PARAMS, plus
 * PSB_PUSH_EACH for async), so no kernel of the service ever waits for anybody.  The shard keeps
 * serving its peers whatever this rank's worker thread does.  `desc` is copied; the device memory
 * it points to must outlive the service.  poll_us > 0 sleeps between launches.
 * ref: the worker service of a ps task, encoder:96-109; Apply* running on the ps, encoder:137. */
int psb_ps_service_start(const psb_push_desc_t* desc, int poll_us, void** handle);
int psb_ps_service_stop(void* handle, uint64_t* launches /* optional out */);
/* Block the stream until every rank has published params-ready >= step_id on `pad`. */
int psb_wait_params(uint32_t* pad, int world, uint32_t step_id, void* stream);
/* The same against the step token in local_ctrl[2] (what this rank's last push recorded), bounded by
 * timeout_ns with the outcome in local_ctrl[3]: for consumers that do not watch readiness flags when
 * the push itself no longer waits for the other owners (pull mode). */
int psb_wait_params_token(uint32_t* pad, int world, uint32_t* local_ctrl, uint64_t timeout_ns, void* stream);

/* ------------------------------------------------------------------ peer memory (CUDA IPC)
 * psb_heap_alloc cudaMalloc's and zeroes one "symmetric heap" (THE allocating call of this
 * library; every buffer the store uses is carved from it).  psb_ipc_export/psb_ipc_open move
 * a 64-byte cudaIpcMemHandle_t between the ranks of one box so kernels can dereference peer
 * pointers over NVLink.  ref: the ps<->worker gRPC edges set up at encoder:96-98,157. */
int psb_heap_alloc(size_t bytes, void** out);
int psb_heap_free(void* p);
int psb_ipc_export(void* p, unsigned char* handle64);
int psb_ipc_open(const unsigned char* handle64, void** out);
int psb_ipc_close(void* p);

/* Zero `rows` runs of `row_bytes` bytes that lie `pitch` bytes apart (cudaMemset[2D]Async: a memset
 * node, not a kernel) -- recurrent state that must start at zero (c_0, dc_T, h_{-1}). */
int psb_memset_zero(void* p, size_t pitch, size_t row_bytes, size_t rows, void* stream);

/* U(lo,hi) noise [rows, cols] generated on the device from (seed, *step_counter, index); the
 * kernel advances *step_counter by one, so a captured graph draws fresh noise on every replay.
 * Any of the fp32 / bf16 / bf16-transposed outputs may be NULL.  ref: np.random.uniform at gan:189. */
int psb_uniform_noise(float* out, int rows, int cols, int ld, float lo, float hi, uint32_t seed,
                      uint32_t* step_counter, void* out_bf16, int ld_bf16, void* out_t,
                      int ld_t, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PSB200_H_ */
