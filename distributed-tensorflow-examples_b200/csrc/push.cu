// Gradient push fused with the owner-side optimizer update, over NVLink peer memory.
//
// Every rank keeps its gradients of ALL variables in one flat fp32 buffer that its peers
// can address (CUDA IPC).  The rank that owns a variable -- or a row block of one: big
// variables are split over all ranks -- ("its ps shard") runs this kernel over the segment:
//
//   wait until every rank (or the first R of N) has published "gradients of step s are
//   complete"                                                            (signal pads)
//   g      = (sum over contributing ranks r of grad_r[i]) / contributors -- P2P 128-bit loads,
//            all ranks' loads in flight together, summed in rank order (deterministic)
//   w,slots = optimizer(w, slots, g)                             -- SGD / RMSProp / Adam, fp32
//   bf16 compute copies of w (natural and transposed layout) written to the owner's arena,
//   or to every rank's arena (push-mode all-gather) when `bcast` is set
//   the last block bumps Adam's beta powers and global_step and publishes
//   "parameters of step s+1 are ready" to every rank.
//
// i.e. reduce-scatter + optimizer apply (+ all-gather) in ONE kernel, no NCCL and no
// intermediate reduced-gradient buffer.  With world == 1 the same kernel is the plain
// local optimizer step.
//
// The same kernel also serves the decoupled ps role (csrc/ps_service.cu): launched with
// PSB_PUSH_TRY_GRADS it looks at the pads once, applies only if enough gradients have
// arrived (sync, R of N: late gradients are never read -- they are stale) or applies every
// newly arrived gradient on its own and acknowledges its sender (async, PSB_PUSH_EACH),
// and returns without ever spinning.  Workers then only signal and wait (psb_push_commit).
//
// In the legacy async mode a worker launches it with world == 1 and `master`/`slot`
// pointers that live in the OWNER's HBM: a lock-free remote read-modify-write; SGD
// degenerates to a peer atomic add.
//
// Every cross-GPU wait is bounded: a %globaltimer deadline (timeout_ns) and a host-settable
// abort word end the spin, the kernel records why in local_ctrl[3] and finishes without
// applying; the trainer turns that into an exception instead of a hung GPU.
//
// ref: Optimizer.minimize -> Send(grad) -> Apply{GradientDescent,RMSProp,Adam} on the ps +
//      AssignAdd(global_step): encoder:137, gan:158-159, lstm:97; sync aggregation per
//      SyncReplicasOptimizer.
#include "../../include/psb200.h"
#include "host_common.h"
#include "ptx.cuh"

#include <mutex>

namespace psb {

constexpr int kPushThreads = 256;
constexpr int kMaxWorld = 16;
constexpr int kPadGrads = 0;      // pad[kPadGrads + r]  : rank r's gradients for step >= value are complete
constexpr int kPadParams = 16;    // pad[kPadParams + r] : rank r has applied step >= value
constexpr int kPadAck = 32;       // pad[kPadAck + r]    : owner r has consumed this rank's gradient number `value` (async)

// words of local_ctrl (PSB_PUSH_CTRL_WORDS of them, zero-initialised by the caller)
enum : int {
    kCtrlGo = 0,        // launch sequence number block (0,0) publishes once it has decided the contributor mask
    kCtrlDone = 1,      // finished-block counter
    kCtrlStep = 2,      // last applied step token
    kCtrlStatus = 3,    // 0 ok | PSB_PUSH_ST_*: sticky until the host clears it
    kCtrlAbort = 4,     // host writes non-zero: every spin ends, status = PSB_PUSH_ST_ABORTED
    kCtrlMask = 5,      // contributor mask of this launch (0 = nothing applied)
    kCtrlSeq = 6,       // launches completed
    kCtrlApplied = 7,   // aggregates / gradients applied so far (tests, service statistics)
    kCtrlConsumed = 8,  // [8 .. 8+16): async, last gradient number consumed per rank
};

struct PushArgs {
    const psb_push_seg_t* segs;
    int nseg;
    const float* const* grad_ptrs;     // [world]
    void* const* arena_ptrs;           // [world] bf16 arenas
    float* const* farena_ptrs;         // [world] fp32 arenas (bias copies)
    uint32_t* const* pad_ptrs;         // [world]
    int world, rank;
    int n_arenas;                      // entries of arena_ptrs/farena_ptrs written when bcast
    float inv_replicas;
    uint32_t step_id;                  // 0 = take local_ctrl[2] + 1
    uint32_t* local_ctrl;
    psb_beta_update_t beta[2];
    int nbeta;
    int32_t* global_step;
    int step_delta;
    int async_sgd_atomic;              // 1: SGD segments use atomicAdd on master (peer atomics)
    int phases;                        // PSB_PUSH_* bits
    int replicas;                      // gradients to aggregate (R of N); <= 0 or >= world: all
    int big_tiles;                     // the launch carries the dynamic shared memory of the 128x128-tile path
    unsigned long long* prof;          // optional [8]: t0, t1, -, -, sum(wait grads), sum(apply), sum(wait params), calls
    unsigned long long timeout_ns;     // 0 = wait forever
    uint32_t* const* ready_ptrs;       // [world] readiness arrays (per-segment "this row block is refreshed" flags)
    uint32_t* seg_done;                // finished-block counters, one per readiness word
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// L2-coherent (L1-bypassing) loads; no "memory" clobber so independent loads can be
// batched -- ordering against the flag wait comes from the __syncthreads() after it.
__device__ __forceinline__ float4 ld_cg_v4(const float* p) {
    float4 r;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ float ld_cg_f32(const float* p) {
    float r;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t r;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Release fence for what this block wrote (possibly into peers' memory) before it is counted as done.
// With peers it must be system scope; alone on the box, device scope is enough and much cheaper
// (a system-scope fence costs about 3 us here).
__device__ __forceinline__ void release_fence(bool peers) {
    if (peers) asm volatile("fence.acq_rel.sys;" ::: "memory");
    else asm volatile("fence.acq_rel.gpu;" ::: "memory");
}

// Four consecutive floats: one 128-bit access when the group is whole and aligned, otherwise
// `nvalid` scalar ones (rows whose length is not a multiple of four, the tail of a vector).
__device__ __forceinline__ float4 ld4_cg(const float* p, bool v4, int nvalid) {
    if (v4) return ld_cg_v4(p);
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    if (nvalid > 0) r.x = ld_cg_f32(p);
    if (nvalid > 1) r.y = ld_cg_f32(p + 1);
    if (nvalid > 2) r.z = ld_cg_f32(p + 2);
    if (nvalid > 3) r.w = ld_cg_f32(p + 3);
    return r;
}
__device__ __forceinline__ float4 ld4(const float* p, bool v4, int nvalid) {
    if (v4) return *reinterpret_cast<const float4*>(p);
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    if (nvalid > 0) r.x = p[0];
    if (nvalid > 1) r.y = p[1];
    if (nvalid > 2) r.z = p[2];
    if (nvalid > 3) r.w = p[3];
    return r;
}
__device__ __forceinline__ void st4(float* p, const float4& v, bool v4, int nvalid) {
    if (v4) { *reinterpret_cast<float4*>(p) = v; return; }
    if (nvalid > 0) p[0] = v.x;
    if (nvalid > 1) p[1] = v.y;
    if (nvalid > 2) p[2] = v.z;
    if (nvalid > 3) p[3] = v.w;
}

struct Spin {
    const uint32_t* abort_word;
    unsigned long long timeout_ns;
};

// Wait until *flag >= want (wrap-safe).  0 = seen, otherwise the PSB_PUSH_ST_* reason it gave up.
__device__ __forceinline__ int spin_flag(const uint32_t* flag, uint32_t want, const Spin& sp, int timeout_code) {
    unsigned long long t0 = 0;
    uint32_t it = 0;
    while (static_cast<int32_t>(ld_acquire_sys(flag) - want) < 0) {
        if ((++it & 1023u) == 0) {
            if (ld_volatile_u32(sp.abort_word) != 0) return PSB_PUSH_ST_ABORTED;
            if (sp.timeout_ns != 0) {
                const unsigned long long now = global_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > sp.timeout_ns) return timeout_code;
            }
        }
    }
    return 0;
}

struct OptCtx {
    int kind;
    float lr, h1, h2, eps, lr_t;
};

// One element of the TF dense Apply* kernels on register values; slots are updated in place.
__device__ __forceinline__ float apply_val(const OptCtx& o, float w, float g, float& s0, float& s1) {
    if (o.kind < 0) return w;                 // refresh the compute copies only
    if (o.kind == 0) return w - o.lr * g;
    if (o.kind == 1) {
        s0 = o.h1 * s0 + (1.0f - o.h1) * g * g;
        s1 = o.h2 * s1 + o.lr * g * rsqrtf(s0 + o.eps);
        return w - s1;
    }
    s0 += (1.0f - o.h1) * (g - s0);
    s1 += (1.0f - o.h2) * (g * g - s1);
    return w - o.lr_t * s0 / (sqrtf(s1) + o.eps);
}

// The gradient of one group of four elements from every contributing rank (t[r], loaded by the
// caller with all loads in flight) applied to w/v0/v1: the mean in one optimizer step (sync), or
// one optimizer step per rank (async, `each`).
template <int MAXW>
__device__ __forceinline__ void apply4(const OptCtx& o, const float4 (&t)[MAXW], int world, uint32_t mask, bool each, float inv,
                                       float4& w, float4& v0, float4& v1) {
    if (!each) {
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < MAXW; ++r)
            if (r < world && ((mask >> r) & 1u)) { g.x += t[r].x; g.y += t[r].y; g.z += t[r].z; g.w += t[r].w; }
        w.x = apply_val(o, w.x, g.x * inv, v0.x, v1.x);
        w.y = apply_val(o, w.y, g.y * inv, v0.y, v1.y);
        w.z = apply_val(o, w.z, g.z * inv, v0.z, v1.z);
        w.w = apply_val(o, w.w, g.w * inv, v0.w, v1.w);
    } else {
#pragma unroll
        for (int r = 0; r < MAXW; ++r)
            if (r < world && ((mask >> r) & 1u)) {
                w.x = apply_val(o, w.x, t[r].x, v0.x, v1.x);
                w.y = apply_val(o, w.y, t[r].y, v0.y, v1.y);
                w.z = apply_val(o, w.z, t[r].z, v0.z, v1.z);
                w.w = apply_val(o, w.w, t[r].w, v0.w, v1.w);
            }
    }
}

// Templated on the largest world the register-resident batch of peer loads is sized for.
template <int MAXW, bool kBig>
__global__ void __launch_bounds__(kPushThreads, 2) push_apply_kernel(PushArgs a) {
    __shared__ const float* s_grad[kMaxWorld];
    __shared__ __nv_bfloat16* s_arena[kMaxWorld];
    __shared__ float* s_farena[kMaxWorld];
    __shared__ int s_bad;
    __shared__ uint32_t s_mask;
    __shared__ bool is_last;
    __shared__ bool seg_last;
    __shared__ __nv_bfloat16 tile[2][32][34];
    // 128x128 tile of new weights (+8 pad: 272-byte rows): dynamic, only launches that have
    // PSB_SEG_BIG_TILES segments ask for it (34816 bytes)
    extern __shared__ __align__(16) uint8_t big_raw[];
    __nv_bfloat16 (*big)[136] = reinterpret_cast<__nv_bfloat16 (*)[136]>(big_raw);      // kBig instantiation only

    const int tid = threadIdx.x;
    const bool first_block = blockIdx.x == 0 && blockIdx.y == 0;
    const uint32_t total_blocks = gridDim.x * gridDim.y;
    uint32_t* const ctrl = a.local_ctrl;
    const Spin sp{ctrl + kCtrlAbort, a.timeout_ns};

    // ---------------- prologue on static data only: overlaps the tail of the kernel launched
    // before this one (programmatic dependent launch)
    const bool has_seg = static_cast<int>(blockIdx.y) < a.nseg;
    psb_push_seg_t s;
    if (has_seg) s = a.segs[blockIdx.y];
    uint32_t nb = 0, bx = 0;
    if (has_seg) {
        nb = s.nblocks != 0 ? s.nblocks : gridDim.x;
        bx = s.block_x + blockIdx.x;
    }
    const bool has_work = has_seg && bx < nb && (a.phases & PSB_PUSH_APPLY);
    if (tid < kMaxWorld) {
        if (tid < a.world) s_grad[tid] = a.grad_ptrs[tid];
        if (tid < a.n_arenas) {
            s_arena[tid] = static_cast<__nv_bfloat16*>(a.arena_ptrs[tid]);
            s_farena[tid] = a.farena_ptrs[tid];
        }
    }
    if (tid == 0) { s_bad = 0; s_mask = 0; }
    uint32_t* my_pad = nullptr;
    if (a.world > 1) my_pad = a.pad_ptrs[a.rank];
    griddep_wait();

    const bool auto_step = a.step_id == 0;
    if (auto_step) a.step_id = ld_volatile_u32(ctrl + kCtrlStep) + 1;
    const uint32_t seq = ld_volatile_u32(ctrl + kCtrlSeq) + 1;       // this launch
    const bool each = (a.phases & PSB_PUSH_EACH) != 0;
    const bool all_ranks = !each && (a.replicas <= 0 || a.replicas >= a.world);
    uint32_t mask = a.world >= 32 ? 0xffffffffu : ((1u << a.world) - 1u);
    float inv = a.inv_replicas;
    __syncthreads();

    if (a.prof != nullptr && first_block && tid == 0) a.prof[0] = global_ns();
    // ---------------- cross-GPU: the gradients this launch consumes are in place
    if (a.world > 1) {
        if (first_block && (a.phases & PSB_PUSH_SIGNAL_GRADS) && tid < a.world) {
            // the gradients were written by earlier kernels of this stream: complete and in L2 (where
            // peers read them) since those kernels ended; the release store needs no extra fence
            st_release_sys(a.pad_ptrs[tid] + kPadGrads + a.rank, a.step_id);
        }
        if (all_ranks && (a.phases & PSB_PUSH_WAIT_GRADS)) {
            // every block watches the pad itself: no hop through a block-0 "go" flag
            if ((has_work || first_block) && tid < a.world) {
                const int bad = spin_flag(my_pad + kPadGrads + tid, a.step_id, sp, PSB_PUSH_ST_TIMEOUT_GRADS);
                if (bad) atomicMax(&s_bad, bad);
            }
        } else if (a.phases & (PSB_PUSH_WAIT_GRADS | PSB_PUSH_TRY_GRADS)) {
            // R of N / async: block (0,0) decides which gradients this launch takes, everybody follows
            if (first_block) {
                if (tid == 0) {
                    uint32_t m = 0;
                    int bad = 0;
                    unsigned long long t0 = 0;
                    for (uint32_t it = 0;; ++it) {
                        m = 0;
                        for (int r = 0; r < a.world; ++r) {
                            const uint32_t f = ld_acquire_sys(my_pad + kPadGrads + r);
                            const bool in = each ? (static_cast<int32_t>(f - ld_volatile_u32(ctrl + kCtrlConsumed + r)) > 0)
                                                 : (static_cast<int32_t>(f - a.step_id) >= 0);
                            if (in) m |= 1u << r;
                        }
                        if (__popc(m) >= (each ? 1 : a.replicas)) break;
                        m = 0;
                        if (a.phases & PSB_PUSH_TRY_GRADS) break;
                        if ((it & 255u) == 255u) {
                            if (ld_volatile_u32(sp.abort_word) != 0) { bad = PSB_PUSH_ST_ABORTED; break; }
                            if (sp.timeout_ns != 0) {
                                const unsigned long long now = global_ns();
                                if (t0 == 0) t0 = now;
                                else if (now - t0 > sp.timeout_ns) { bad = PSB_PUSH_ST_TIMEOUT_GRADS; break; }
                            }
                        }
                    }
                    if (each && m != 0) {
                        // remember WHICH gradient of each rank this is: acknowledged at the end
                        for (int r = 0; r < a.world; ++r)
                            if ((m >> r) & 1u) ctrl[kCtrlConsumed + 16 + r] = ld_acquire_sys(my_pad + kPadGrads + r);
                    }
                    if (bad) atomicMax(&s_bad, bad);
                    s_mask = m;
                    *reinterpret_cast<volatile uint32_t*>(ctrl + kCtrlMask) = m;
                    __threadfence();
                    atomicExch(ctrl + kCtrlGo, seq);
                }
            } else if (tid == 0) {
                while (static_cast<int32_t>(atomicAdd(ctrl + kCtrlGo, 0u) - seq) < 0) {
                }
                s_mask = ld_volatile_u32(ctrl + kCtrlMask);
                __threadfence_system();
            }
        }
        __syncthreads();
        if (!all_ranks && (a.phases & (PSB_PUSH_WAIT_GRADS | PSB_PUSH_TRY_GRADS))) {
            mask = s_mask;
            if (!each && mask != 0) inv = 1.0f / static_cast<float>(__popc(mask));
        }
        if (a.prof != nullptr && first_block && tid == 0) a.prof[1] = global_ns();
    }
    const int bad_grads = s_bad;
    const bool go = bad_grads == 0 && mask != 0;

    // ---------------- reduce + apply + cast for this block's share of its segment
    if (has_work && go) {
        OptCtx o;
        o.kind = s.opt_kind; o.lr = s.lr; o.h1 = s.h1; o.h2 = s.h2; o.eps = s.eps; o.lr_t = s.lr;
        if (s.opt_kind == 2) {
            const float b1p = s.beta_pow[0], b2p = s.beta_pow[1];
            o.lr_t = s.lr * sqrtf(1.0f - b2p) / (1.0f - b1p);
        }
        float* const master = s.master;
        float* const s0 = s.slot0;
        float* const s1 = s.slot1;
        const int cols = static_cast<int>(s.cols);
        const bool aligned = (s.grad_offset % 4 == 0) && ((reinterpret_cast<uintptr_t>(master) & 15) == 0);
        const bool atomic_sgd = a.async_sgd_atomic && o.kind == 0;
        const bool need_g = o.kind >= 0;
        const bool zero_g = (s.flags & PSB_SEG_ZERO_GRAD) != 0 && need_g;
        const int n_dst = s.bcast ? a.n_arenas : 1;
        if (kBig && (s.w_off >= 0 || s.wt_off >= 0) && (s.flags & PSB_SEG_BIG_TILES) && !atomic_sgd) {
            // ---- big matrices: 128x128 tiles.  The reduction and the optimizer step run over the tile in
            // sixteen 32x32 pieces (MAXW peer loads in flight per thread each), the new weights are
            // collected as bf16 in shared memory, and BOTH copies then leave in 16-byte stores that form
            // 256-byte runs per row (natural copy) and per column (transposed copy).  Over NVLink that is
            // what matters: the 64-byte runs of the small-tile path reach about a tenth of the link rate.
            const int rows = static_cast<int>(s.numel / s.cols);
            const int tc = (cols + 127) >> 7, trw = (rows + 127) >> 7;
            const int ntiles = tc * trw;
            const int ty = tid >> 3, tx4 = (tid & 7) * 4;
            const bool v4 = aligned && (cols % 4 == 0);
            for (int T = bx; T < ntiles; T += nb) {
                const int R0 = (T / tc) * 128, C0 = (T % tc) * 128;
#pragma unroll 1
                for (int sub = 0; sub < 16; ++sub) {
                    const int lr = (sub >> 2) * 32 + ty, lc = (sub & 3) * 32 + tx4;
                    const int r = R0 + lr, c = C0 + lc;
                    const bool live = r < rows && c < cols;
                    const int nv = live ? min(4, cols - c) : 0;
                    float4 t[MAXW];
                    float4 w = make_float4(0.f, 0.f, 0.f, 0.f), v0 = w, v1 = w;
                    if (live) {
                        const size_t i = static_cast<size_t>(r) * cols + c;
                        if (need_g) {
#pragma unroll
                            for (int rk = 0; rk < MAXW; ++rk)
                                if (rk < a.world && ((mask >> rk) & 1u)) t[rk] = ld4_cg(s_grad[rk] + s.grad_offset + i, v4, nv);
                        }
                        w = ld4(master + i, v4, nv);
                        if (o.kind > 0) { v0 = ld4(s0 + i, v4, nv); v1 = ld4(s1 + i, v4, nv); }
                        if (need_g) apply4<MAXW>(o, t, a.world, mask, each, inv, w, v0, v1);
                        st4(master + i, w, v4, nv);
                        if (o.kind > 0) { st4(s0 + i, v0, v4, nv); st4(s1 + i, v1, v4, nv); }
                        if (zero_g) {
                            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                            for (int rk = 0; rk < a.world; ++rk)
                                st4(const_cast<float*>(s_grad[rk]) + s.grad_offset + i, z, v4, nv);
                        }
                    }
                    const float wv[4] = {w.x, w.y, w.z, w.w};
                    uint2 pk;
                    pk.x = pack_bf16(nv > 0 ? wv[0] : 0.f, nv > 1 ? wv[1] : 0.f);
                    pk.y = pack_bf16(nv > 2 ? wv[2] : 0.f, nv > 3 ? wv[3] : 0.f);
                    *reinterpret_cast<uint2*>(&big[lr][lc]) = pk;
                }
                __syncthreads();
                // natural copy: 128 rows x 16 chunks of 8 columns (16 bytes)
                if (s.w_off >= 0) {
                    for (int idx = tid; idx < 128 * 16; idx += kPushThreads) {
                        const int lr = idx >> 4, lc = (idx & 15) * 8;
                        const int r = R0 + lr, c = C0 + lc;
                        if (r >= rows || c >= cols) continue;
                        const uint4 v = *reinterpret_cast<const uint4*>(&big[lr][lc]);
                        const size_t off = static_cast<size_t>(s.w_off) + static_cast<size_t>(r) * s.ld_w + c;
                        const bool full = (c + 8 <= cols) && ((off & 7) == 0);
                        for (int d = 0; d < n_dst; ++d) {
                            __nv_bfloat16* pd = s_arena[s.bcast ? d : s.home] + off;
                            if (full) {
                                *reinterpret_cast<uint4*>(pd) = v;
                            } else {
                                for (int e = 0; e < 8 && c + e < cols; ++e) pd[e] = big[lr][lc + e];
                            }
                        }
                    }
                }
                // transposed copy: 128 columns x 16 chunks of 8 rows (16 bytes of one wt row)
                if (s.wt_off >= 0) {
                    for (int idx = tid; idx < 128 * 16; idx += kPushThreads) {
                        const int lc = idx >> 4, lr0 = (idx & 15) * 8;
                        const int cg = C0 + lc, rg = R0 + lr0;
                        if (cg >= cols || rg >= rows) continue;
                        __nv_bfloat16 col[8];
#pragma unroll
                        for (int k = 0; k < 8; ++k) col[k] = big[lr0 + k][lc];
                        const size_t off = static_cast<size_t>(s.wt_off) + static_cast<size_t>(cg) * s.ld_wt + rg;
                        const bool full = (rg + 8 <= rows) && ((off & 7) == 0);
                        uint4 v;
                        v.x = static_cast<uint32_t>(__bfloat16_as_ushort(col[0])) | (static_cast<uint32_t>(__bfloat16_as_ushort(col[1])) << 16);
                        v.y = static_cast<uint32_t>(__bfloat16_as_ushort(col[2])) | (static_cast<uint32_t>(__bfloat16_as_ushort(col[3])) << 16);
                        v.z = static_cast<uint32_t>(__bfloat16_as_ushort(col[4])) | (static_cast<uint32_t>(__bfloat16_as_ushort(col[5])) << 16);
                        v.w = static_cast<uint32_t>(__bfloat16_as_ushort(col[6])) | (static_cast<uint32_t>(__bfloat16_as_ushort(col[7])) << 16);
                        for (int d = 0; d < n_dst; ++d) {
                            __nv_bfloat16* pd = s_arena[s.bcast ? d : s.home] + off;
                            if (full) {
                                *reinterpret_cast<uint4*>(pd) = v;
                            } else {
                                for (int k = 0; k < 8 && rg + k < rows; ++k) pd[k] = col[k];
                            }
                        }
                    }
                }
                __syncthreads();
            }
        } else if (s.w_off >= 0 || s.wt_off >= 0) {
            // ---- matrices: 32x32 tiles so that BOTH bf16 copies are written in full 32-byte
            // sectors (the transposed copy goes through a shared-memory transpose).  Two tiles
            // per iteration: every peer load of both tiles first, then the math, then the stores.
            const int rows = static_cast<int>(s.numel / s.cols);
            const int tcols = (cols + 31) >> 5, trows = (rows + 31) >> 5;
            const int ntiles = tcols * trows;
            const int ty = tid >> 3, tx4 = (tid & 7) * 4;
            const bool v4 = aligned && (cols % 4 == 0);
            constexpr int TU = MAXW >= 8 ? 1 : 2;      // tiles per iteration (register budget: TU * MAXW peer loads)
            for (int t0 = bx * TU; t0 < ntiles; t0 += nb * TU) {
                float4 t[TU][MAXW];
                float4 w[TU], v0[TU], v1[TU];
                bool live[TU];
                int nv[TU];
#pragma unroll
                for (int u = 0; u < TU; ++u) {
                    const int tI = t0 + u;
                    const int r = (tI / tcols) * 32 + ty, c = (tI % tcols) * 32 + tx4;
                    live[u] = tI < ntiles && r < rows && c < cols;
                    nv[u] = live[u] ? min(4, cols - c) : 0;
                    w[u] = v0[u] = v1[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (live[u]) {
                        const size_t i = static_cast<size_t>(r) * cols + c;
                        if (need_g) {
#pragma unroll
                            for (int rk = 0; rk < MAXW; ++rk)
                                if (rk < a.world && ((mask >> rk) & 1u)) t[u][rk] = ld4_cg(s_grad[rk] + s.grad_offset + i, v4, nv[u]);
                        }
                        if (!atomic_sgd) w[u] = ld4(master + i, v4, nv[u]);
                        if (o.kind > 0) { v0[u] = ld4(s0 + i, v4, nv[u]); v1[u] = ld4(s1 + i, v4, nv[u]); }
                    }
                }
#pragma unroll
                for (int u = 0; u < TU; ++u) {
                    const int tI = t0 + u;
                    const int r = (tI / tcols) * 32 + ty, c = (tI % tcols) * 32 + tx4;
                    float wv[4] = {0.f, 0.f, 0.f, 0.f};
                    if (live[u]) {
                        const size_t i = static_cast<size_t>(r) * cols + c;
                        if (atomic_sgd) {
                            float gs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                            for (int rk = 0; rk < MAXW; ++rk)
                                if (rk < a.world && ((mask >> rk) & 1u)) { gs[0] += t[u][rk].x; gs[1] += t[u][rk].y; gs[2] += t[u][rk].z; gs[3] += t[u][rk].w; }
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                if (e < nv[u]) {
                                    const float d = -o.lr * gs[e] * inv;
                                    wv[e] = atomicAdd(master + i + e, d) + d;
                                }
                            }
                        } else {
                            if (need_g) apply4(o, t[u], a.world, mask, each, inv, w[u], v0[u], v1[u]);
                            st4(master + i, w[u], v4, nv[u]);
                            if (o.kind > 0) { st4(s0 + i, v0[u], v4, nv[u]); st4(s1 + i, v1[u], v4, nv[u]); }
                            wv[0] = w[u].x; wv[1] = w[u].y; wv[2] = w[u].z; wv[3] = w[u].w;
                        }
                        if (zero_g) {
                            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                            for (int rk = 0; rk < a.world; ++rk)
                                st4(const_cast<float*>(s_grad[rk]) + s.grad_offset + i, z, v4, nv[u]);
                        }
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e) tile[u][ty][tx4 + e] = __float2bfloat16_rn(e < nv[u] ? wv[e] : 0.0f);
                    if (live[u] && s.w_off >= 0) {
                        const uint32_t lo = pack_bf16(wv[0], wv[1]), hi = pack_bf16(wv[2], wv[3]);
                        for (int d = 0; d < n_dst; ++d) {
                            __nv_bfloat16* p = s_arena[s.bcast ? d : s.home] + s.w_off + static_cast<size_t>(r) * s.ld_w + c;
                            if (nv[u] == 4) {
                                *reinterpret_cast<uint2*>(p) = make_uint2(lo, hi);
                            } else {
#pragma unroll
                                for (int e = 0; e < 4; ++e)
                                    if (e < nv[u]) p[e] = __float2bfloat16_rn(wv[e]);
                            }
                        }
                    }
                }
                __syncthreads();
                if (s.wt_off >= 0) {
#pragma unroll
                    for (int u = 0; u < TU; ++u) {
                        const int tI = t0 + u;
                        // thread (cc = ty, rr4 = tx4): column cc of the tile, rows rr4..rr4+3 -> 8 contiguous bytes of wt
                        const int cg = (tI % tcols) * 32 + ty, rg = (tI / tcols) * 32 + tx4;
                        if (tI < ntiles && cg < cols && rg < rows) {
                            const __nv_bfloat16 b0 = tile[u][tx4][ty], b1 = tile[u][tx4 + 1][ty], b2 = tile[u][tx4 + 2][ty],
                                                b3 = tile[u][tx4 + 3][ty];
                            for (int d = 0; d < n_dst; ++d) {
                                __nv_bfloat16* p = s_arena[s.bcast ? d : s.home] + s.wt_off + static_cast<size_t>(cg) * s.ld_wt + rg;
                                if (rg + 3 < rows) {
                                    uint2 pk;
                                    pk.x = static_cast<uint32_t>(__bfloat16_as_ushort(b0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(b1)) << 16);
                                    pk.y = static_cast<uint32_t>(__bfloat16_as_ushort(b2)) | (static_cast<uint32_t>(__bfloat16_as_ushort(b3)) << 16);
                                    *reinterpret_cast<uint2*>(p) = pk;
                                } else {
                                    p[0] = b0;
                                    if (rg + 1 < rows) p[1] = b1;
                                    if (rg + 2 < rows) p[2] = b2;
                                }
                            }
                        }
                    }
                }
                __syncthreads();
            }
        } else {
            // ---- vectors (biases, flat shards): groups of four, U groups per thread per iteration
            constexpr int U = MAXW >= 8 ? 1 : (MAXW >= 4 ? 2 : 4);   // >= 4 (8 from MAXW = 2) 128-bit loads in flight per thread
            const uint32_t n4 = (s.numel + 3) / 4;
            for (uint32_t base = bx * (kPushThreads * U); base < n4; base += nb * (kPushThreads * U)) {
                float4 t[U][MAXW];
                float4 w[U], v0[U], v1[U];
                int nv[U];
#pragma unroll
                for (int j = 0; j < U; ++j) {
                    const uint32_t i4 = base + j * kPushThreads + tid;
                    nv[j] = i4 < n4 ? min(4u, s.numel - i4 * 4) : 0;
                    w[j] = v0[j] = v1[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (nv[j] > 0) {
                        const size_t i = static_cast<size_t>(i4) * 4;
                        const bool v4 = aligned && nv[j] == 4;
                        if (need_g) {
#pragma unroll
                            for (int rk = 0; rk < MAXW; ++rk)
                                if (rk < a.world && ((mask >> rk) & 1u)) t[j][rk] = ld4_cg(s_grad[rk] + s.grad_offset + i, v4, nv[j]);
                        }
                        if (!atomic_sgd) w[j] = ld4(master + i, v4, nv[j]);
                        if (o.kind > 0) { v0[j] = ld4(s0 + i, v4, nv[j]); v1[j] = ld4(s1 + i, v4, nv[j]); }
                    }
                }
#pragma unroll
                for (int j = 0; j < U; ++j) {
                    if (nv[j] == 0) continue;
                    const uint32_t i4 = base + j * kPushThreads + tid;
                    const size_t i = static_cast<size_t>(i4) * 4;
                    const bool v4 = aligned && nv[j] == 4;
                    if (atomic_sgd) {
                        float gs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (int rk = 0; rk < MAXW; ++rk)
                            if (rk < a.world && ((mask >> rk) & 1u)) { gs[0] += t[j][rk].x; gs[1] += t[j][rk].y; gs[2] += t[j][rk].z; gs[3] += t[j][rk].w; }
                        float wv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            if (e < nv[j]) {
                                const float d = -o.lr * gs[e] * inv;
                                wv[e] = atomicAdd(master + i + e, d) + d;
                            }
                        }
                        w[j] = make_float4(wv[0], wv[1], wv[2], wv[3]);
                    } else {
                        if (need_g) apply4(o, t[j], a.world, mask, each, inv, w[j], v0[j], v1[j]);
                        st4(master + i, w[j], v4, nv[j]);
                        if (o.kind > 0) { st4(s0 + i, v0[j], v4, nv[j]); st4(s1 + i, v1[j], v4, nv[j]); }
                    }
                    if (zero_g) {
                        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                        for (int rk = 0; rk < a.world; ++rk)
                            st4(const_cast<float*>(s_grad[rk]) + s.grad_offset + i, z, v4, nv[j]);
                    }
                    if (s.f_off >= 0) {
                        // the fp32 arena is padded to whole groups of four: full 128-bit stores
                        const bool fv4 = ((s.f_off + i) % 4 == 0);
                        for (int d = 0; d < n_dst; ++d)
                            st4(s_farena[s.bcast ? d : s.home] + s.f_off + i, w[j], fv4 && nv[j] == 4, nv[j]);
                    }
                }
            }
        }
    }

    // ---------------- completion: last block finishes the step
    __syncthreads();
    // remote writes happen when there are peers' arenas / gradient buffers to write, or (legacy async)
    // when master and slots are peer pointers
    const bool peers = a.world > 1 || a.n_arenas > 1 || a.async_sgd_atomic;
    if (tid == 0) {
        if (bad_grads) atomicMax(ctrl + kCtrlStatus, static_cast<uint32_t>(bad_grads));
        release_fence(peers);
        seg_last = false;
        if (has_work && go && s.ready_slot != 0 && a.ready_ptrs != nullptr) {
            // last block of this segment: its row block is complete in every arena it goes to
            uint32_t* cnt = a.seg_done + (s.ready_slot - 1);
            if (atomicAdd(cnt, 1u) == nb - 1) {
                *cnt = 0;
                __threadfence();
                seg_last = true;
            }
        }
        const uint32_t prev = atomicAdd(ctrl + kCtrlDone, 1u);
        is_last = (prev == total_blocks - 1);
    }
    __syncthreads();
    if (seg_last && tid < a.world) st_release_sys(a.ready_ptrs[tid] + (s.ready_slot - 1), a.step_id);
    unsigned long long t_applied = 0;
    if (is_last) {
        const bool failed = ld_volatile_u32(ctrl + kCtrlStatus) != 0;
        const bool applied = go && !failed && (a.phases & PSB_PUSH_APPLY);
        if (tid == 0) {
            if (a.prof != nullptr) t_applied = global_ns();
            ctrl[kCtrlDone] = 0;
            ctrl[kCtrlSeq] = seq;
            if (applied) {
                const int napplied = each ? __popc(mask) : 1;
                if (auto_step) ctrl[kCtrlStep] = a.step_id;
                ctrl[kCtrlApplied] += napplied;
                for (int k = 0; k < a.nbeta; ++k) {
                    for (int n = 0; n < napplied; ++n) {
                        a.beta[k].beta_pow[0] *= a.beta[k].beta1;
                        a.beta[k].beta_pow[1] *= a.beta[k].beta2;
                    }
                }
                if (a.global_step != nullptr && a.step_delta != 0) atomicAdd(a.global_step, a.step_delta * napplied);
                if (each)
                    for (int r = 0; r < a.world; ++r)
                        if ((mask >> r) & 1u) ctrl[kCtrlConsumed + r] = ctrl[kCtrlConsumed + 16 + r];
            }
            // no fence here: the release stores below order everything this block has seen
            // (bar.sync makes thread 0's writes part of the signalling threads' causality order)
            release_fence(false);
        }
        __syncthreads();
        if (a.world > 1 && applied && tid < a.world) {
            if (each) {
                // tell every sender whose gradient was consumed that its buffer is free again
                if ((mask >> tid) & 1u) st_release_sys(a.pad_ptrs[tid] + kPadAck + a.rank, ctrl[kCtrlConsumed + tid]);
            } else if (a.phases & PSB_PUSH_SIGNAL_PARAMS) {
                st_release_sys(a.pad_ptrs[tid] + kPadParams + a.rank, a.step_id);
            }
        }
        if (a.world > 1 && !failed && (a.phases & PSB_PUSH_WAIT_PARAMS) && tid < a.world) {
            // ...and do not let this stream run ahead until every owner has finished writing
            // this step's parameters (they may be writing into OUR arena).
            const int bad = spin_flag(my_pad + kPadParams + tid, a.step_id, sp, PSB_PUSH_ST_TIMEOUT_PARAMS);
            if (bad) atomicMax(ctrl + kCtrlStatus, static_cast<uint32_t>(bad));
        }
        if (a.prof != nullptr) {
            __syncthreads();
            if (tid == 0) {
                const unsigned long long t_end = global_ns();
                const unsigned long long t0 = *reinterpret_cast<volatile unsigned long long*>(a.prof);
                const unsigned long long t1 = a.world > 1 ? *reinterpret_cast<volatile unsigned long long*>(a.prof + 1) : t0;
                a.prof[4] += t1 - t0;
                a.prof[5] += t_applied - t1;
                a.prof[6] += t_end - t_applied;
                a.prof[7] += 1;
            }
        }
    }
}

// Worker half of the decoupled protocol (the ps half is push_apply_kernel launched by the ps
// service).  One block.
//   sync  : publish "my gradients for token s are complete", wait until every owner has applied
//           aggregate >= s, adopt min(applied) + 1 as the next token (a late worker skips ahead:
//           its gradient was stale and has been dropped)
//   async : publish "my gradient number n is complete" (wait = 0), or wait until every owner has
//           acknowledged gradient number ctrl[2] (wait = 1; issued before the next backward pass
//           overwrites the gradient buffer)
struct CommitArgs {
    uint32_t* const* pad_ptrs;
    int world, rank;
    uint32_t* ctrl;          // worker control words: [2] token / gradient number, [3] status, [4] abort
    int async_mode;
    int wait;
    unsigned long long timeout_ns;
};

__global__ void push_commit_kernel(CommitArgs a) {
    const int tid = threadIdx.x;
    __shared__ uint32_t s_min[kMaxWorld];
    const Spin sp{a.ctrl + kCtrlAbort, a.timeout_ns};
    uint32_t* my_pad = a.pad_ptrs[a.rank];
    if (a.async_mode) {
        if (a.wait) {
            const uint32_t n = ld_volatile_u32(a.ctrl + kCtrlStep);
            if (tid < a.world && n != 0) {
                const int bad = spin_flag(my_pad + kPadAck + tid, n, sp, PSB_PUSH_ST_TIMEOUT_PARAMS);
                if (bad) atomicMax(a.ctrl + kCtrlStatus, static_cast<uint32_t>(bad));
            }
        } else {
            const uint32_t n = ld_volatile_u32(a.ctrl + kCtrlStep) + 1;
            if (tid < a.world) {
                __threadfence_system();
                st_release_sys(a.pad_ptrs[tid] + kPadGrads + a.rank, n);
            }
            __syncthreads();
            if (tid == 0) a.ctrl[kCtrlStep] = n;
        }
        return;
    }
    uint32_t token = ld_volatile_u32(a.ctrl + kCtrlStep);
    if (token == 0) token = 1;
    if (tid < a.world) {
        __threadfence_system();
        st_release_sys(a.pad_ptrs[tid] + kPadGrads + a.rank, token);
        const int bad = spin_flag(my_pad + kPadParams + tid, token, sp, PSB_PUSH_ST_TIMEOUT_PARAMS);
        if (bad) atomicMax(a.ctrl + kCtrlStatus, static_cast<uint32_t>(bad));
        s_min[tid] = ld_acquire_sys(my_pad + kPadParams + tid);
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t m = s_min[0];
        for (int r = 1; r < a.world; ++r)
            if (static_cast<int32_t>(s_min[r] - m) < 0) m = s_min[r];
        a.ctrl[kCtrlStep] = m + 1;
    }
}

// Spin until every rank has published params-ready >= step_id (used by consumers that read
// peers' arenas directly and by tests).
__global__ void wait_params_kernel(uint32_t* pad, int world, uint32_t step_id) {
    if (threadIdx.x < world) {
        while (static_cast<int32_t>(ld_acquire_sys(pad + kPadParams + threadIdx.x) - step_id) < 0) {
        }
    }
}

// The same against the token this rank's last push recorded, bounded.
__global__ void wait_params_token_kernel(uint32_t* pad, int world, uint32_t* ctrl, unsigned long long timeout_ns) {
    const Spin sp{ctrl + kCtrlAbort, timeout_ns};
    const uint32_t step = ld_volatile_u32(ctrl + kCtrlStep);
    if (static_cast<int>(threadIdx.x) < world) {
        const int bad = spin_flag(pad + kPadParams + threadIdx.x, step, sp, PSB_PUSH_ST_TIMEOUT_PARAMS);
        if (bad) atomicMax(ctrl + kCtrlStatus, static_cast<uint32_t>(bad));
    }
}

static int check_desc(const psb_push_desc_t* d, const char* who, bool device_tables) {
    if (!d) return fail(PSB_ERR_ARG, "%s: null descriptor", who);
    if (d->world < 1 || d->world > kMaxWorld || d->rank < 0 || d->rank >= d->world)
        return fail(PSB_ERR_ARG, "%s: world must be 1..16 and 0 <= rank < world", who);
    if (d->nbeta < 0 || d->nbeta > 2) return fail(PSB_ERR_ARG, "%s: at most two Adam optimizers", who);
    if (device_tables) {
        if (d->nseg < 0 || (d->nseg > 0 && !d->segs)) return fail(PSB_ERR_ARG, "%s: bad segment table", who);
        if (!d->grad_ptrs || !d->arena_ptrs || !d->farena_ptrs || !d->local_ctrl) return fail(PSB_ERR_ARG, "%s: null table", who);
        if (d->world > 1 && !d->pad_ptrs) return fail(PSB_ERR_ARG, "%s: world > 1 needs signal pads", who);
    }
    if ((d->phases & PSB_PUSH_EACH) && d->async_sgd_atomic)
        return fail(PSB_ERR_ARG, "%s: PSB_PUSH_EACH applies on the owner; it excludes async_sgd_atomic", who);
    return PSB_OK;
}

static void fill_args(PushArgs& a, const psb_push_desc_t* d) {
    a.world = d->world; a.rank = d->rank; a.n_arenas = d->n_arenas > 0 ? d->n_arenas : d->world;
    a.inv_replicas = d->inv_replicas;
    a.step_id = d->step_id;
    a.beta[0] = d->beta[0]; a.beta[1] = d->beta[1]; a.nbeta = d->nbeta;
    a.global_step = d->global_step; a.step_delta = d->step_delta;
    a.async_sgd_atomic = d->async_sgd_atomic;
    a.phases = d->phases;
    if (a.phases == 0) {
        a.phases = PSB_PUSH_APPLY;
        if (d->world > 1) {
            a.phases |= PSB_PUSH_SIGNAL_GRADS | PSB_PUSH_WAIT_GRADS;
            if (d->signal_params) a.phases |= PSB_PUSH_SIGNAL_PARAMS | PSB_PUSH_WAIT_PARAMS;
        }
    }
    a.replicas = d->replicas;
    a.big_tiles = d->big_tiles;
    a.prof = reinterpret_cast<unsigned long long*>(d->prof);
    a.timeout_ns = d->timeout_ns;
    a.ready_ptrs = d->seg_done != nullptr ? d->ready_ptrs : nullptr;
    a.seg_done = d->seg_done;
}

static int launch_push(const PushArgs& a, dim3 grid, cudaStream_t st, bool pdl, bool big_tiles = false) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kPushThreads);
    cfg.dynamicSmemBytes = big_tiles ? 128 * 136 * 2 : 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t e;
    // the 128x128-tile path is a separate instantiation: launches without big segments run exactly
    // the small-tile code (registers, shared memory) they always ran
    if (big_tiles) {
        if (a.world <= 1) e = cudaLaunchKernelEx(&cfg, push_apply_kernel<1, true>, a);
        else if (a.world <= 2) e = cudaLaunchKernelEx(&cfg, push_apply_kernel<2, true>, a);
        else if (a.world <= 4) e = cudaLaunchKernelEx(&cfg, push_apply_kernel<4, true>, a);
        else if (a.world <= 8) e = cudaLaunchKernelEx(&cfg, push_apply_kernel<8, true>, a);
        else e = cudaLaunchKernelEx(&cfg, push_apply_kernel<16, true>, a);
    } else if (a.world <= 1) e = cudaLaunchKernelEx(&cfg, push_apply_kernel<1, false>, a);
    else if (a.world <= 2) e = cudaLaunchKernelEx(&cfg, push_apply_kernel<2, false>, a);
    else if (a.world <= 4) e = cudaLaunchKernelEx(&cfg, push_apply_kernel<4, false>, a);
    else if (a.world <= 8) e = cudaLaunchKernelEx(&cfg, push_apply_kernel<8, false>, a);
    else e = cudaLaunchKernelEx(&cfg, push_apply_kernel<16, false>, a);
    if (e != cudaSuccess) return fail(PSB_ERR_CUDA, "push_apply launch failed: %s", cudaGetErrorString(e));
    return PSB_OK;
}

}  // namespace psb

extern "C" int psb_push_apply(const psb_push_desc_t* d, void* stream) {
    using namespace psb;
    int rc = check_desc(d, "psb_push_apply", true);
    if (rc != PSB_OK) return rc;
    PushArgs a;
    a.segs = d->segs; a.nseg = d->nseg; a.grad_ptrs = d->grad_ptrs; a.arena_ptrs = d->arena_ptrs;
    a.farena_ptrs = d->farena_ptrs; a.pad_ptrs = d->pad_ptrs; a.local_ctrl = d->local_ctrl;
    fill_args(a, d);
    int gx = d->blocks_per_seg > 0 ? d->blocks_per_seg : 16;
    if (gx > 1024) gx = 1024;
    if (d->nseg > 65535) return fail(PSB_ERR_LIMIT, "psb_push_apply: at most 65535 table entries");
    dim3 grid(gx, d->nseg > 0 ? d->nseg : 1);
    // No co-residency requirement: a block only ever waits for peers' block (0,0) -- which is in the
    // first wave of its grid and waits for nobody in its own grid -- and the last block to finish for
    // the owners' final flags.
    rc = launch_push(a, grid, static_cast<cudaStream_t>(stream), true, d->big_tiles != 0);
    if (rc != PSB_OK) return rc;
    return after_launch("push_apply");
}

extern "C" int psb_push_commit(uint32_t* const* pad_ptrs, int world, int rank, uint32_t* worker_ctrl,
                               int async_mode, int wait, uint64_t timeout_ns, void* stream) {
    using namespace psb;
    if (!pad_ptrs || !worker_ctrl || world < 1 || world > kMaxWorld || rank < 0 || rank >= world)
        return fail(PSB_ERR_ARG, "psb_push_commit: bad argument");
    CommitArgs a{pad_ptrs, world, rank, worker_ctrl, async_mode, wait, timeout_ns};
    push_commit_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a);
    return after_launch("push_commit");
}

extern "C" int psb_wait_params_token(uint32_t* pad, int world, uint32_t* local_ctrl, uint64_t timeout_ns, void* stream) {
    using namespace psb;
    if (!pad || !local_ctrl || world < 1 || world > 16) return fail(PSB_ERR_ARG, "psb_wait_params_token: bad argument");
    wait_params_token_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(pad, world, local_ctrl, timeout_ns);
    return after_launch("wait_params_token");
}

extern "C" int psb_wait_params(uint32_t* pad, int world, uint32_t step_id, void* stream) {
    using namespace psb;
    if (!pad || world < 1 || world > 16) return fail(PSB_ERR_ARG, "psb_wait_params: bad argument");
    wait_params_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(pad, world, step_id);
    return after_launch("wait_params");
}

// ======================================================================================
// psb_push_apply_noalloc -- the same kernel for callers that are not allowed to allocate.
//
// psb_push_apply needs three kinds of device memory besides the tensors it works on: the
// control words the kernel uses as scratch (finished-block counter, step token, status),
// the segment table and the per-rank pointer tables.  Somebody has to cudaMalloc those.
// Some callers cannot: an optimizer hook that runs inside another framework's memory pool
// and must not touch the CUDA allocator, a plug-in whose host forbids allocations after
// start-up, or code on a latency path where an allocation (an implicit device
// synchronisation) is not acceptable.  For them this entry point keeps all of that in ONE
// file-scope device array of fixed size, so the library's own image provides the memory:
// every GPU that loads this library holds sizeof(NoAllocScratch) = 9472 bytes (9.25 KiB)
// until the process exits, whether or not this entry point is ever called.
//
// What the caller must know
//   * Limits: at most PSB_NOALLOC_MAX_SEGS (64) segments and PSB_NOALLOC_MAX_WORLD (16)
//     ranks/arenas; larger requests are refused with PSB_ERR_LIMIT and a message.  The grid
//     is blocks_per_seg (default 16, at most 64) x nseg for every nseg up to the limit.
//   * The segment table and the pointer tables are passed as HOST arrays and copied into the
//     scratch with cudaMemcpyAsync on `stream` before the launch: keep them valid and
//     unchanged until the stream has passed those copies.
//   * The first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress, with that device current) and caches it.  Make that first call
//     outside any stream capture.
//   * All calls on a device share the one array and it keeps what the previous call left
//     there.  The finished-block counter is cleared on every call; the step token, the
//     launch sequence number and the status word are deliberately kept, because the
//     multi-GPU handshake compares against the token.  So, per device and for the life of
//     the process: explicit step ids (desc->step_id != 0) must keep increasing across ALL
//     callers of this entry point, and a caller must not alternate between this entry point
//     and psb_push_apply (whose token lives in the caller's own local_ctrl) on the same
//     signal pads -- the two tokens would drift apart and the ranks would wait for
//     different steps.
//   * Two calls that overlap on the GPU (different streams, or graphs replayed concurrently)
//     race on the array: serialise them.
//   * Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
namespace psb {

struct NoAllocScratch {
    uint32_t ctrl[PSB_PUSH_CTRL_WORDS];
    psb_push_seg_t segs[PSB_NOALLOC_MAX_SEGS];
    const float* grad_ptrs[PSB_NOALLOC_MAX_WORLD];
    void* arena_ptrs[PSB_NOALLOC_MAX_WORLD];
    float* farena_ptrs[PSB_NOALLOC_MAX_WORLD];
    uint32_t* pad_ptrs[PSB_NOALLOC_MAX_WORLD];
};
static_assert(sizeof(NoAllocScratch) == 9472, "update the size quoted in psb200.h, DESIGN.md and INTEGRATION.md");

// In host code this name is only a handle for cudaGetSymbolAddress; the device address it
// resolves to is what gets passed to the kernel.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

static int noalloc_scratch_for_current_device(NoAllocScratch** out) {
    constexpr int kMaxDevices = 64;
    static NoAllocScratch* cache[kMaxDevices] = {nullptr};
    static std::mutex mu;
    int dev = -1;
    PSB_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(PSB_ERR_LIMIT, "psb_push_apply_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (cache[dev] == nullptr) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr)
            return fail(PSB_ERR_CUDA, "psb_push_apply_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                        cudaGetErrorString(e));
        cache[dev] = static_cast<NoAllocScratch*>(p);
    }
    *out = cache[dev];
    return PSB_OK;
}

}  // namespace psb

extern "C" int psb_push_apply_noalloc(const psb_push_seg_t* host_segs, int nseg,
                                      const float* const* host_grad_ptrs, void* const* host_arena_ptrs,
                                      float* const* host_farena_ptrs, uint32_t* const* host_pad_ptrs,
                                      const psb_push_desc_t* d, void* stream) {
    using namespace psb;
    if (!d) return fail(PSB_ERR_ARG, "psb_push_apply_noalloc: null descriptor");
    if (nseg < 0 || (nseg > 0 && !host_segs)) return fail(PSB_ERR_ARG, "psb_push_apply_noalloc: bad segment table");
    if (nseg > PSB_NOALLOC_MAX_SEGS)
        return fail(PSB_ERR_LIMIT, "psb_push_apply_noalloc: %d segments exceed the static scratch (max %d)", nseg,
                    PSB_NOALLOC_MAX_SEGS);
    if (d->world < 1 || d->rank < 0 || d->rank >= d->world)
        return fail(PSB_ERR_ARG, "psb_push_apply_noalloc: need world >= 1 and 0 <= rank < world");
    const int n_arenas = d->n_arenas > 0 ? d->n_arenas : d->world;
    if (d->world > PSB_NOALLOC_MAX_WORLD || n_arenas > PSB_NOALLOC_MAX_WORLD)
        return fail(PSB_ERR_LIMIT, "psb_push_apply_noalloc: world %d / arenas %d exceed the static scratch (max %d)",
                    d->world, n_arenas, PSB_NOALLOC_MAX_WORLD);
    if (!host_grad_ptrs || !host_arena_ptrs || !host_farena_ptrs)
        return fail(PSB_ERR_ARG, "psb_push_apply_noalloc: null pointer table");
    if (d->world > 1 && !host_pad_ptrs) return fail(PSB_ERR_ARG, "psb_push_apply_noalloc: world > 1 needs signal pads");
    int rc = check_desc(d, "psb_push_apply_noalloc", false);
    if (rc != PSB_OK) return rc;
    int gx = d->blocks_per_seg > 0 ? d->blocks_per_seg : 16;
    if (gx > 64) gx = 64;
    dim3 grid(gx, nseg > 0 ? nseg : 1);

    NoAllocScratch* sc = nullptr;
    rc = noalloc_scratch_for_current_device(&sc);
    if (rc != PSB_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // stage the tables; clear only what must start at zero (the finished-block counter)
    if (nseg > 0) PSB_CUDA_OK(cudaMemcpyAsync(sc->segs, host_segs, sizeof(psb_push_seg_t) * nseg, cudaMemcpyHostToDevice, st));
    PSB_CUDA_OK(cudaMemcpyAsync(sc->grad_ptrs, host_grad_ptrs, sizeof(void*) * d->world, cudaMemcpyHostToDevice, st));
    PSB_CUDA_OK(cudaMemcpyAsync(sc->arena_ptrs, host_arena_ptrs, sizeof(void*) * n_arenas, cudaMemcpyHostToDevice, st));
    PSB_CUDA_OK(cudaMemcpyAsync(sc->farena_ptrs, host_farena_ptrs, sizeof(void*) * n_arenas, cudaMemcpyHostToDevice, st));
    if (host_pad_ptrs)
        PSB_CUDA_OK(cudaMemcpyAsync(sc->pad_ptrs, host_pad_ptrs, sizeof(void*) * d->world, cudaMemcpyHostToDevice, st));
    PSB_CUDA_OK(cudaMemsetAsync(&sc->ctrl[1], 0, sizeof(uint32_t), st));

    PushArgs a;
    a.segs = sc->segs; a.nseg = nseg; a.grad_ptrs = sc->grad_ptrs; a.arena_ptrs = sc->arena_ptrs;
    a.farena_ptrs = sc->farena_ptrs; a.pad_ptrs = sc->pad_ptrs; a.local_ctrl = sc->ctrl;
    fill_args(a, d);
    a.n_arenas = n_arenas;
    // the launch follows copies, not a kernel: no programmatic dependent launch here
    rc = launch_push(a, grid, st, false, d->big_tiles != 0);
    if (rc != PSB_OK) return rc;
    return after_launch("push_apply_noalloc");
}
