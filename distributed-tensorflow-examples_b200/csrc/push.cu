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
};

__device__ __forceinline__ float4 ld_volatile_v4(const float* p) {
    float4 r;
    asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ float ld_volatile_f32(const float* p) {
    float r;
    asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
    return r;
}

struct OptCtx {
    int kind;
    float lr, h1, h2, eps, lr_t;
};

__device__ __forceinline__ float apply_one(const OptCtx& o, float w, float g, float* s0, float* s1, size_t i) {
    if (o.kind < 0) return w;                 // refresh the compute copies only
    if (o.kind == 0) return w - o.lr * g;
    if (o.kind == 1) {
        float ms = s0[i], mom = s1[i];
        ms = o.h1 * ms + (1.0f - o.h1) * g * g;
        mom = o.h2 * mom + o.lr * g * rsqrtf(ms + o.eps);
        s0[i] = ms;
        s1[i] = mom;
        return w - mom;
    }
    float m = s0[i], v = s1[i];
    m += (1.0f - o.h1) * (g - m);
    v += (1.0f - o.h2) * (g * g - v);
    s0[i] = m;
    s1[i] = v;
    return w - o.lr_t * m / (sqrtf(v) + o.eps);
}

__global__ void __launch_bounds__(kPushThreads) push_apply_kernel(PushArgs a) {
    const int tid = threadIdx.x;
    const bool first_block = blockIdx.x == 0 && blockIdx.y == 0;
    const uint32_t total_blocks = gridDim.x * gridDim.y;
    const bool auto_step = a.step_id == 0;
    if (auto_step) a.step_id = *reinterpret_cast<volatile uint32_t*>(a.local_ctrl + 2) + 1;

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
        const uint32_t stride = gridDim.x * kPushThreads;
        if (vec) {
            for (uint32_t i4 = blockIdx.x * kPushThreads + tid; i4 < s.numel / 4; i4 += stride) {
                const size_t i = static_cast<size_t>(i4) * 4;
                float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int r = 0; r < a.world; ++r) {
                    const float4 t = ld_volatile_v4(a.grad_ptrs[r] + s.grad_offset + i);
                    g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
                }
                g.x *= a.inv_replicas; g.y *= a.inv_replicas; g.z *= a.inv_replicas; g.w *= a.inv_replicas;
                float4 w = *reinterpret_cast<float4*>(master + i);
                w.x = apply_one(o, w.x, g.x, s0, s1, i);
                w.y = apply_one(o, w.y, g.y, s0, s1, i + 1);
                w.z = apply_one(o, w.z, g.z, s0, s1, i + 2);
                w.w = apply_one(o, w.w, g.w, s0, s1, i + 3);
                *reinterpret_cast<float4*>(master + i) = w;
                const int row = static_cast<int>(i / cols), col = static_cast<int>(i % cols);
                const __nv_bfloat162 lo = __floats2bfloat162_rn(w.x, w.y), hi = __floats2bfloat162_rn(w.z, w.w);
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
                        *reinterpret_cast<float4*>(a.farena_ptrs[s.bcast ? d : s.home] + s.f_off + i) = w;
                    if (s.wt_off >= 0) {
                        __nv_bfloat16* p = arena + s.wt_off + static_cast<size_t>(col) * s.ld_wt + row;
                        p[0] = lo.x;
                        p[s.ld_wt] = lo.y;
                        p[2 * static_cast<size_t>(s.ld_wt)] = hi.x;
                        p[3 * static_cast<size_t>(s.ld_wt)] = hi.y;
                    }
                }
            }
        } else {
            for (uint32_t i = blockIdx.x * kPushThreads + tid; i < s.numel; i += stride) {
                float g = 0.0f;
                for (int r = 0; r < a.world; ++r) g += ld_volatile_f32(a.grad_ptrs[r] + s.grad_offset + i);
                g *= a.inv_replicas;
                float w;
                if (a.async_sgd_atomic && o.kind == 0) {
                    w = atomicAdd(master + i, -o.lr * g) - o.lr * g;
                } else {
                    w = apply_one(o, master[i], g, s0, s1, i);
                    master[i] = w;
                }
                const int row = static_cast<int>(i / cols), col = static_cast<int>(i % cols);
                const __nv_bfloat16 wb = __float2bfloat16_rn(w);
                for (int d = 0; d < n_dst; ++d) {
                    __nv_bfloat16* arena = static_cast<__nv_bfloat16*>(a.arena_ptrs[s.bcast ? d : s.home]);
                    if (s.w_off >= 0) arena[s.w_off + static_cast<size_t>(row) * s.ld_w + col] = wb;
                    if (s.f_off >= 0) a.farena_ptrs[s.bcast ? d : s.home][s.f_off + i] = w;
                    if (s.wt_off >= 0) arena[s.wt_off + static_cast<size_t>(col) * s.ld_wt + row] = wb;
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
    if (is_last) {
        if (tid == 0) {
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
    int gx = d->blocks_per_seg > 0 ? d->blocks_per_seg : 16;
    if (gx > 64) gx = 64;
    dim3 grid(gx, d->nseg > 0 ? d->nseg : 1);
    if (grid.x * grid.y > 148) return fail(PSB_ERR_LIMIT, "psb_push_apply: grid %u x %u would not be co-resident", grid.x, grid.y);
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
    push_apply_kernel<<<grid, kPushThreads, 0, st>>>(a);
    return after_launch("push_apply_noalloc");
}
