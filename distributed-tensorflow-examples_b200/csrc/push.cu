// Gradient push fused with the owner-side optimizer update, over NVLink peer memory.
//
// Every rank keeps its gradients of ALL variables in one flat fp32 buffer that its peers
// can address (CUDA IPC).  The rank that owns a variable ("its ps shard") runs this kernel
// over the variable's segment:
//
//   wait until every rank has published "gradients of step s are complete"   (signal pads)
//   g      = (sum over ranks r of grad_r[i]) * inv_replicas     -- P2P 128-bit loads, fixed
//                                                                   rank order (deterministic)
//   w,slots = optimizer(w, slots, g)                             -- SGD / RMSProp / Adam, fp32
//   bf16 compute copies of w (natural and transposed layout) written to the owner's arena,
//   or to every rank's arena (push-mode all-gather) when `bcast` is set
//   the last block bumps Adam's beta powers and global_step and publishes
//   "parameters of step s+1 are ready" to every rank.
//
// i.e. reduce-scatter + optimizer apply (+ all-gather) in ONE kernel, no NCCL and no
// intermediate reduced-gradient buffer.  With world == 1 the same kernel is the plain
// local optimizer step.  In async (Hogwild) mode a worker launches it with world == 1 and
// `master`/`slot` pointers that live in the OWNER's HBM: the update is then a lock-free
// remote read-modify-write, and SGD degenerates to a peer atomic add.
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
constexpr int kPadGrads = 0;      // pad[kPadGrads + r]  : rank r's gradients for step >= value are complete
constexpr int kPadParams = 16;    // pad[kPadParams + r] : rank r has applied step >= value

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
    uint32_t* local_ctrl;              // [0] go flag, [1] finished-block counter, [2] last step id
    psb_beta_update_t beta[2];
    int nbeta;
    int32_t* global_step;
    int step_delta;
    int async_sgd_atomic;              // 1: SGD segments use atomicAdd on master (peer atomics)
    int signal_params;                 // 1: publish params-ready to peers at the end
    unsigned long long* prof;          // optional [8]: t0, t1, -, -, sum(wait grads), sum(apply), sum(wait params), calls
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// L2-coherent (L1-bypassing) 128-bit load; no "memory" clobber so independent loads can be
// batched — ordering against the flag wait comes from the __syncthreads() after it.
__device__ __forceinline__ float4 ld_volatile_v4(const float* p) {
    float4 r;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ float ld_volatile_f32(const float* p) {
    float r;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
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

__global__ void __launch_bounds__(kPushThreads) push_apply_kernel(PushArgs a) {
    const int tid = threadIdx.x;
    const bool first_block = blockIdx.x == 0 && blockIdx.y == 0;
    const uint32_t total_blocks = gridDim.x * gridDim.y;
    const bool auto_step = a.step_id == 0;
    if (auto_step) a.step_id = *reinterpret_cast<volatile uint32_t*>(a.local_ctrl + 2) + 1;

    if (a.prof != nullptr && first_block && tid == 0) a.prof[0] = global_ns();
    // ---------------- cross-GPU: all gradients of this step are in place
    if (a.world > 1) {
        if (first_block) {
            if (tid < a.world) {
                __threadfence_system();
                st_release_sys(a.pad_ptrs[tid] + kPadGrads + a.rank, a.step_id);
                const uint32_t* mine = a.pad_ptrs[a.rank] + kPadGrads + tid;
                while (static_cast<int32_t>(ld_acquire_sys(mine) - a.step_id) < 0) {
                }
            }
            __syncthreads();
            if (tid == 0) {
                if (a.prof != nullptr) a.prof[1] = global_ns();
                __threadfence();
                atomicExch(a.local_ctrl, a.step_id);
            }
        } else {
            if (tid == 0) {
                while (static_cast<int32_t>(atomicAdd(a.local_ctrl, 0u) - a.step_id) < 0) {
                }
                __threadfence_system();
            }
            __syncthreads();
        }
    }

    // ---------------- reduce + apply + cast for this block's segment
    if (static_cast<int>(blockIdx.y) < a.nseg) {
        const psb_push_seg_t s = a.segs[blockIdx.y];
        OptCtx o;
        o.kind = s.opt_kind; o.lr = s.lr; o.h1 = s.h1; o.h2 = s.h2; o.eps = s.eps; o.lr_t = s.lr;
        if (s.opt_kind == 2) {
            const float b1p = s.beta_pow[0], b2p = s.beta_pow[1];
            o.lr_t = s.lr * sqrtf(1.0f - b2p) / (1.0f - b1p);
        }
        float* master = s.master;
        float* s0 = s.slot0;
        float* s1 = s.slot1;
        const int cols = static_cast<int>(s.cols);
        const bool vec = (s.numel % 4 == 0) && (s.grad_offset % 4 == 0) && (cols % 4 == 0) &&
                         ((reinterpret_cast<uintptr_t>(master) & 15) == 0) &&
                         !(a.async_sgd_atomic && s.opt_kind == 0);
        const int n_dst = s.bcast ? a.n_arenas : 1;
        if (s.w_off >= 0 || s.wt_off >= 0) {
            // ---- matrices: 32x32 tiles so that BOTH bf16 copies are written in full 32-byte
            // sectors (the transposed copy goes through a shared-memory transpose).  Two tiles
            // per iteration: all loads, then the math, then the stores.
            __shared__ __nv_bfloat16 tile[2][32][34];
            const int rows = static_cast<int>(s.numel / s.cols);
            const int tcols = (cols + 31) >> 5, trows = (rows + 31) >> 5;
            const int ntiles = tcols * trows;
            const int ty = tid >> 3, tx4 = (tid & 7) * 4;
            const bool atomic_sgd = a.async_sgd_atomic && o.kind == 0;
            const bool v4 = (cols % 4 == 0) && (s.grad_offset % 4 == 0) && ((reinterpret_cast<uintptr_t>(master) & 15) == 0);
            for (int t0 = blockIdx.x * 2; t0 < ntiles; t0 += gridDim.x * 2) {
                float g[2][4], w[2][4], v0[2][4], v1[2][4];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int tI = t0 + u;
                    const int r = (tI / tcols) * 32 + ty, c = (tI % tcols) * 32 + tx4;
#pragma unroll
                    for (int e = 0; e < 4; ++e) g[u][e] = w[u][e] = v0[u][e] = v1[u][e] = 0.0f;
                    if (tI < ntiles && r < rows && c < cols) {
                        const size_t i = static_cast<size_t>(r) * cols + c;
                        if (v4) {
                            // all ranks' slices first (independent NVLink loads in flight together),
                            // then the sum in rank order (deterministic)
                            float4 t[8];
#pragma unroll
                            for (int rk = 0; rk < 8; ++rk)
                                if (rk < a.world) t[rk] = ld_volatile_v4(a.grad_ptrs[rk] + s.grad_offset + i);
#pragma unroll
                            for (int rk = 0; rk < 8; ++rk)
                                if (rk < a.world) { g[u][0] += t[rk].x; g[u][1] += t[rk].y; g[u][2] += t[rk].z; g[u][3] += t[rk].w; }
                            for (int rk = 8; rk < a.world; ++rk) {
                                const float4 tt = ld_volatile_v4(a.grad_ptrs[rk] + s.grad_offset + i);
                                g[u][0] += tt.x; g[u][1] += tt.y; g[u][2] += tt.z; g[u][3] += tt.w;
                            }
                            if (!atomic_sgd) {
                                const float4 t = *reinterpret_cast<const float4*>(master + i);
                                w[u][0] = t.x; w[u][1] = t.y; w[u][2] = t.z; w[u][3] = t.w;
                            }
                            if (o.kind > 0) {
                                const float4 p0 = *reinterpret_cast<const float4*>(s0 + i);
                                const float4 p1 = *reinterpret_cast<const float4*>(s1 + i);
                                v0[u][0] = p0.x; v0[u][1] = p0.y; v0[u][2] = p0.z; v0[u][3] = p0.w;
                                v1[u][0] = p1.x; v1[u][1] = p1.y; v1[u][2] = p1.z; v1[u][3] = p1.w;
                            }
                        } else {
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                if (c + e < cols) {
                                    for (int rk = 0; rk < a.world; ++rk) g[u][e] += ld_volatile_f32(a.grad_ptrs[rk] + s.grad_offset + i + e);
                                    if (!atomic_sgd) w[u][e] = master[i + e];
                                    if (o.kind > 0) { v0[u][e] = s0[i + e]; v1[u][e] = s1[i + e]; }
                                }
                            }
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int tI = t0 + u;
                    const int r = (tI / tcols) * 32 + ty, c = (tI % tcols) * 32 + tx4;
                    const bool live = tI < ntiles && r < rows && c < cols;
                    if (live) {
                        const size_t i = static_cast<size_t>(r) * cols + c;
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float gg = g[u][e] * a.inv_replicas;
                            if (c + e < cols) {
                                if (atomic_sgd) w[u][e] = atomicAdd(master + i + e, -o.lr * gg) - o.lr * gg;
                                else w[u][e] = apply_val(o, w[u][e], gg, v0[u][e], v1[u][e]);
                            }
                        }
                        if (v4) {
                            if (!atomic_sgd) *reinterpret_cast<float4*>(master + i) = make_float4(w[u][0], w[u][1], w[u][2], w[u][3]);
                            if (o.kind > 0) {
                                *reinterpret_cast<float4*>(s0 + i) = make_float4(v0[u][0], v0[u][1], v0[u][2], v0[u][3]);
                                *reinterpret_cast<float4*>(s1 + i) = make_float4(v1[u][0], v1[u][1], v1[u][2], v1[u][3]);
                            }
                        } else {
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                if (c + e < cols) {
                                    if (!atomic_sgd) master[i + e] = w[u][e];
                                    if (o.kind > 0) { s0[i + e] = v0[u][e]; s1[i + e] = v1[u][e]; }
                                }
                            }
                        }
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e) tile[u][ty][tx4 + e] = __float2bfloat16_rn(live && c + e < cols ? w[u][e] : 0.0f);
                    if (live && s.w_off >= 0) {
                        const uint32_t lo = pack_bf16(w[u][0], w[u][1]), hi = pack_bf16(w[u][2], w[u][3]);
                        for (int d = 0; d < n_dst; ++d) {
                            __nv_bfloat16* p = static_cast<__nv_bfloat16*>(a.arena_ptrs[s.bcast ? d : s.home]) + s.w_off +
                                               static_cast<size_t>(r) * s.ld_w + c;
                            if (c + 3 < cols) {
                                *reinterpret_cast<uint2*>(p) = make_uint2(lo, hi);
                            } else {
                                for (int e = 0; e < 4 && c + e < cols; ++e) p[e] = __float2bfloat16_rn(w[u][e]);
                            }
                        }
                    }
                }
                __syncthreads();
                if (s.wt_off >= 0) {
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int tI = t0 + u;
                        // thread (cc = ty, rr4 = tx4): column cc of the tile, rows rr4..rr4+3 -> 8 contiguous bytes of wt
                        const int cg = (tI % tcols) * 32 + ty, rg = (tI / tcols) * 32 + tx4;
                        if (tI < ntiles && cg < cols && rg < rows) {
                            const __nv_bfloat16 b0 = tile[u][tx4][ty], b1 = tile[u][tx4 + 1][ty], b2 = tile[u][tx4 + 2][ty],
                                                b3 = tile[u][tx4 + 3][ty];
                            for (int d = 0; d < n_dst; ++d) {
                                __nv_bfloat16* p = static_cast<__nv_bfloat16*>(a.arena_ptrs[s.bcast ? d : s.home]) + s.wt_off +
                                                   static_cast<size_t>(cg) * s.ld_wt + rg;
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
        constexpr int U = 4;      // independent vectors per thread per iteration: all loads, then math, then stores
        if (vec) {
            const uint32_t n4 = s.numel / 4;
            for (uint32_t base = blockIdx.x * (kPushThreads * U); base < n4; base += gridDim.x * (kPushThreads * U)) {
                float4 g[U], w[U], v0[U], v1[U];
#pragma unroll
                for (int j = 0; j < U; ++j) {
                    const uint32_t i4 = base + j * kPushThreads + tid;
                    g[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                    w[j] = v0[j] = v1[j] = g[j];
                    if (i4 < n4) {
                        const size_t i = static_cast<size_t>(i4) * 4;
                        float4 t[8];
#pragma unroll
                        for (int r = 0; r < 8; ++r)
                            if (r < a.world) t[r] = ld_volatile_v4(a.grad_ptrs[r] + s.grad_offset + i);
#pragma unroll
                        for (int r = 0; r < 8; ++r)
                            if (r < a.world) { g[j].x += t[r].x; g[j].y += t[r].y; g[j].z += t[r].z; g[j].w += t[r].w; }
                        for (int r = 8; r < a.world; ++r) {
                            const float4 tt = ld_volatile_v4(a.grad_ptrs[r] + s.grad_offset + i);
                            g[j].x += tt.x; g[j].y += tt.y; g[j].z += tt.z; g[j].w += tt.w;
                        }
                        w[j] = *reinterpret_cast<const float4*>(master + i);
                        if (o.kind > 0) {
                            v0[j] = *reinterpret_cast<const float4*>(s0 + i);
                            v1[j] = *reinterpret_cast<const float4*>(s1 + i);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < U; ++j) {
                    const uint32_t i4 = base + j * kPushThreads + tid;
                    if (i4 >= n4) continue;
                    const size_t i = static_cast<size_t>(i4) * 4;
                    const float sc = a.inv_replicas;
                    w[j].x = apply_val(o, w[j].x, g[j].x * sc, v0[j].x, v1[j].x);
                    w[j].y = apply_val(o, w[j].y, g[j].y * sc, v0[j].y, v1[j].y);
                    w[j].z = apply_val(o, w[j].z, g[j].z * sc, v0[j].z, v1[j].z);
                    w[j].w = apply_val(o, w[j].w, g[j].w * sc, v0[j].w, v1[j].w);
                    *reinterpret_cast<float4*>(master + i) = w[j];
                    if (o.kind > 0) {
                        *reinterpret_cast<float4*>(s0 + i) = v0[j];
                        *reinterpret_cast<float4*>(s1 + i) = v1[j];
                    }
                    const int row = static_cast<int>(i / cols), col = static_cast<int>(i % cols);
                    const __nv_bfloat162 lo = __floats2bfloat162_rn(w[j].x, w[j].y), hi = __floats2bfloat162_rn(w[j].z, w[j].w);
                    for (int d = 0; d < n_dst; ++d) {
                        __nv_bfloat16* arena = static_cast<__nv_bfloat16*>(a.arena_ptrs[s.bcast ? d : s.home]);
                        if (s.w_off >= 0) {
                            __nv_bfloat16* p = arena + s.w_off + static_cast<size_t>(row) * s.ld_w + col;
                            uint2 pk;
                            pk.x = *reinterpret_cast<const uint32_t*>(&lo);
                            pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                            *reinterpret_cast<uint2*>(p) = pk;
                        }
                        if (s.f_off >= 0)
                            *reinterpret_cast<float4*>(a.farena_ptrs[s.bcast ? d : s.home] + s.f_off + i) = w[j];
                        if (s.wt_off >= 0) {
                            __nv_bfloat16* p = arena + s.wt_off + static_cast<size_t>(col) * s.ld_wt + row;
                            p[0] = lo.x;
                            p[s.ld_wt] = lo.y;
                            p[2 * static_cast<size_t>(s.ld_wt)] = hi.x;
                            p[3 * static_cast<size_t>(s.ld_wt)] = hi.y;
                        }
                    }
                }
            }
        } else {
            constexpr int US = 8;
            for (uint32_t base = blockIdx.x * (kPushThreads * US); base < s.numel; base += gridDim.x * (kPushThreads * US)) {
                float g[US], w[US], v0[US], v1[US];
                const bool atomic_sgd = a.async_sgd_atomic && o.kind == 0;
#pragma unroll
                for (int j = 0; j < US; ++j) {
                    const uint32_t i = base + j * kPushThreads + tid;
                    g[j] = w[j] = v0[j] = v1[j] = 0.0f;
                    if (i < s.numel) {
                        for (int r = 0; r < a.world; ++r) g[j] += ld_volatile_f32(a.grad_ptrs[r] + s.grad_offset + i);
                        if (!atomic_sgd) w[j] = master[i];
                        if (o.kind > 0) { v0[j] = s0[i]; v1[j] = s1[i]; }
                    }
                }
#pragma unroll
                for (int j = 0; j < US; ++j) {
                    const uint32_t i = base + j * kPushThreads + tid;
                    if (i >= s.numel) continue;
                    const float gg = g[j] * a.inv_replicas;
                    float wn;
                    if (atomic_sgd) {
                        wn = atomicAdd(master + i, -o.lr * gg) - o.lr * gg;
                    } else {
                        wn = apply_val(o, w[j], gg, v0[j], v1[j]);
                        master[i] = wn;
                        if (o.kind > 0) { s0[i] = v0[j]; s1[i] = v1[j]; }
                    }
                    const int row = static_cast<int>(i / cols), col = static_cast<int>(i % cols);
                    const __nv_bfloat16 wb = __float2bfloat16_rn(wn);
                    for (int d = 0; d < n_dst; ++d) {
                        __nv_bfloat16* arena = static_cast<__nv_bfloat16*>(a.arena_ptrs[s.bcast ? d : s.home]);
                        if (s.w_off >= 0) arena[s.w_off + static_cast<size_t>(row) * s.ld_w + col] = wb;
                        if (s.f_off >= 0) a.farena_ptrs[s.bcast ? d : s.home][s.f_off + i] = wn;
                        if (s.wt_off >= 0) arena[s.wt_off + static_cast<size_t>(col) * s.ld_wt + row] = wb;
                    }
                }
            }
        }
        }
    }

    // ---------------- completion: last block finishes the step
    __syncthreads();
    __shared__ bool is_last;
    if (tid == 0) {
        __threadfence_system();
        const uint32_t prev = atomicAdd(a.local_ctrl + 1, 1u);
        is_last = (prev == total_blocks - 1);
    }
    __syncthreads();
    unsigned long long t_applied = 0;
    if (is_last) {
        if (tid == 0) {
            if (a.prof != nullptr) t_applied = global_ns();
            a.local_ctrl[1] = 0;
            if (auto_step) a.local_ctrl[2] = a.step_id;
            for (int k = 0; k < a.nbeta; ++k) {
                a.beta[k].beta_pow[0] *= a.beta[k].beta1;
                a.beta[k].beta_pow[1] *= a.beta[k].beta2;
            }
            if (a.global_step != nullptr && a.step_delta != 0) atomicAdd(a.global_step, a.step_delta);
            __threadfence_system();
        }
        __syncthreads();
        if (a.world > 1 && a.signal_params && tid < a.world) {
            st_release_sys(a.pad_ptrs[tid] + kPadParams + a.rank, a.step_id);
            // ...and do not let this stream run ahead until every owner has finished writing
            // this step's parameters (they may be writing into OUR arena).
            const uint32_t* mine = a.pad_ptrs[a.rank] + kPadParams + tid;
            while (static_cast<int32_t>(ld_acquire_sys(mine) - a.step_id) < 0) {
            }
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

// Spin until every rank has published params-ready >= step_id (used by consumers that read
// peers' arenas directly and by tests).
__global__ void wait_params_kernel(uint32_t* pad, int world, uint32_t step_id) {
    if (threadIdx.x < world) {
        while (static_cast<int32_t>(ld_acquire_sys(pad + kPadParams + threadIdx.x) - step_id) < 0) {
        }
    }
}

}  // namespace psb

extern "C" int psb_push_apply(const psb_push_desc_t* d, void* stream) {
    using namespace psb;
    if (!d) return fail(PSB_ERR_ARG, "psb_push_apply: null descriptor");
    if (d->world < 1 || d->world > 16 || d->rank < 0 || d->rank >= d->world)
        return fail(PSB_ERR_ARG, "psb_push_apply: world must be 1..16 and 0 <= rank < world");
    if (d->nseg < 0 || (d->nseg > 0 && !d->segs)) return fail(PSB_ERR_ARG, "psb_push_apply: bad segment table");
    if (!d->grad_ptrs || !d->arena_ptrs || !d->farena_ptrs || !d->local_ctrl) return fail(PSB_ERR_ARG, "psb_push_apply: null table");
    if (d->world > 1 && !d->pad_ptrs) return fail(PSB_ERR_ARG, "psb_push_apply: world > 1 needs signal pads");
    if (d->nbeta < 0 || d->nbeta > 2) return fail(PSB_ERR_ARG, "psb_push_apply: at most two Adam optimizers");
    PushArgs a;
    a.segs = d->segs; a.nseg = d->nseg; a.grad_ptrs = d->grad_ptrs; a.arena_ptrs = d->arena_ptrs;
    a.farena_ptrs = d->farena_ptrs; a.pad_ptrs = d->pad_ptrs; a.world = d->world; a.rank = d->rank; a.n_arenas = d->n_arenas > 0 ? d->n_arenas : d->world; a.inv_replicas = d->inv_replicas;
    a.step_id = d->step_id; a.local_ctrl = d->local_ctrl;
    a.beta[0] = d->beta[0]; a.beta[1] = d->beta[1]; a.nbeta = d->nbeta;
    a.global_step = d->global_step; a.step_delta = d->step_delta;
    a.async_sgd_atomic = d->async_sgd_atomic; a.signal_params = d->signal_params;
    a.prof = reinterpret_cast<unsigned long long*>(d->prof);
    int gx = d->blocks_per_seg > 0 ? d->blocks_per_seg : 16;
    if (gx > 296) gx = 296;
    dim3 grid(gx, d->nseg > 0 ? d->nseg : 1);
    // the blocks spin on each other: all of them must be resident (2 x 256 threads per SM fit easily)
    if (d->world > 1 && grid.x * grid.y > 296)
        return fail(PSB_ERR_LIMIT, "psb_push_apply: grid %u x %u would not be co-resident", grid.x, grid.y);
    push_apply_kernel<<<grid, kPushThreads, 0, static_cast<cudaStream_t>(stream)>>>(a);
    return after_launch("push_apply");
}

extern "C" int psb_wait_params(uint32_t* pad, int world, uint32_t step_id, void* stream) {
    using namespace psb;
    if (!pad || world < 1 || world > 16) return fail(PSB_ERR_ARG, "psb_wait_params: bad argument");
    wait_params_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(pad, world, step_id);
    return after_launch("wait_params");
}

// ======================================================================================
// psb_push_apply_noalloc — the same kernel for callers that are not allowed to allocate.
//
// psb_push_apply needs three kinds of device memory besides the tensors it works on: the
// control words the kernel uses as scratch (go flag, finished-block counter, step token),
// the segment table and the per-rank pointer tables.  Somebody has to cudaMalloc those.
// Some callers cannot: an optimizer hook that runs inside another framework's memory pool
// and must not touch the CUDA allocator, a plug-in whose host forbids allocations after
// start-up, or code on a latency path where an allocation (an implicit device
// synchronisation) is not acceptable.  For them this entry point keeps all of that in ONE
// file-scope device array of fixed size, so the library's own image provides the memory:
// every GPU that loads this library holds sizeof(NoAllocScratch) bytes (about 7 KiB) until
// the process exits, whether or not this entry point is ever called.
//
// What the caller must know
//   * Limits: at most PSB_NOALLOC_MAX_SEGS (64) segments and PSB_NOALLOC_MAX_WORLD (16)
//     ranks/arenas; larger requests are refused with PSB_ERR_LIMIT and a message.
//   * The segment table and the pointer tables are passed as HOST arrays and copied into the
//     scratch with cudaMemcpyAsync on `stream` before the launch.  Pageable arrays may be
//     reused as soon as the call returns; pinned arrays must stay unchanged until the stream
//     has passed the copies.
//   * The first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress, with that device current) and caches it.  Make that first call
//     outside any stream capture.
//   * All calls on a device share the one array and it keeps what the previous call left
//     there.  The finished-block counter is cleared on every call; the go flag and the step
//     token are deliberately kept, because the multi-GPU handshake compares against them.
//     Two calls that overlap on the GPU (different streams, or graphs replayed concurrently)
//     race on the array: serialise them.
//   * Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
namespace psb {

struct NoAllocScratch {
    uint32_t ctrl[64];                                  // [0] go flag [1] finished blocks [2] step token
    psb_push_seg_t segs[PSB_NOALLOC_MAX_SEGS];
    const float* grad_ptrs[PSB_NOALLOC_MAX_WORLD];
    void* arena_ptrs[PSB_NOALLOC_MAX_WORLD];
    float* farena_ptrs[PSB_NOALLOC_MAX_WORLD];
    uint32_t* pad_ptrs[PSB_NOALLOC_MAX_WORLD];
};

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
    if (d->nbeta < 0 || d->nbeta > 2) return fail(PSB_ERR_ARG, "psb_push_apply_noalloc: at most two Adam optimizers");
    int gx = d->blocks_per_seg > 0 ? d->blocks_per_seg : 16;
    if (gx > 64) gx = 64;
    dim3 grid(gx, nseg > 0 ? nseg : 1);
    if (grid.x * grid.y > 148)
        return fail(PSB_ERR_LIMIT, "psb_push_apply_noalloc: grid %u x %u would not be co-resident", grid.x, grid.y);

    NoAllocScratch* sc = nullptr;
    int rc = noalloc_scratch_for_current_device(&sc);
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
    a.farena_ptrs = sc->farena_ptrs; a.pad_ptrs = sc->pad_ptrs; a.world = d->world; a.rank = d->rank;
    a.n_arenas = n_arenas; a.inv_replicas = d->inv_replicas;
    a.step_id = d->step_id; a.local_ctrl = sc->ctrl;
    a.beta[0] = d->beta[0]; a.beta[1] = d->beta[1]; a.nbeta = d->nbeta;
    a.global_step = d->global_step; a.step_delta = d->step_delta;
    a.async_sgd_atomic = d->async_sgd_atomic; a.signal_params = d->signal_params;
    a.prof = reinterpret_cast<unsigned long long*>(d->prof);
    push_apply_kernel<<<grid, kPushThreads, 0, st>>>(a);
    return after_launch("push_apply_noalloc");
}
