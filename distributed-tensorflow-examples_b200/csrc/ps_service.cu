// The ps role of a rank as a native service thread.
//
// In the reference a parameter server is a process of its own: its C++ thread pool applies
// whatever gradients arrive, whenever they arrive, no matter what the workers' Python threads
// are doing ("the ps Python thread has no hot loop; all ps work happens in TF's C++
// thread-pool in response to RunGraph RPCs").  On the NVLink transport a rank is worker AND ps
// shard; in the fused sync step both roles share one kernel launch, which is the fast path but
// couples them: a slow worker is then also a slow ps.
//
// This service decouples them again.  One std::thread per rank owns a high-priority stream and
// keeps launching the push kernel in its non-blocking form (PSB_PUSH_TRY_GRADS): the kernel looks
// at the signal pads once, applies if there is something to apply --
//     sync  R of N : the aggregate of the first R gradients that arrived for the current step
//                    (late ones are stale and never read),
//     async        : every newly arrived gradient, one optimizer step each, then an
//                    acknowledgement to its sender --
// publishes the result to the peers' pads and exits.  No kernel of the service ever spins, so
// it cannot dead-lock against the worker kernels it shares the GPU with, and it needs neither
// the worker's host thread nor the Python GIL: a worker that stalls in data loading, or dies,
// leaves its shard serving the others.
//
// ref: tf.train.Server's worker service on a ps task (encoder:96-109); accumulator / token
//      machinery of SyncReplicasOptimizer; Hogwild Apply* on the ps (encoder:137).
#include "../../include/psb200.h"
#include "host_common.h"

#include <atomic>
#include <chrono>
#include <thread>

namespace psb {

struct PsService {
    psb_push_desc_t desc;
    int device = 0;
    int poll_us = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    std::thread th;
    std::atomic<bool> stop{false};
    std::atomic<uint64_t> launches{0};
    std::atomic<int> error{0};
};

static void wait_event(PsService* s, cudaEvent_t e) {
    // poll instead of cudaEventSynchronize: the thread must notice `stop` and must not hold a
    // driver lock while the training thread captures a CUDA graph
    while (!s->stop.load(std::memory_order_relaxed)) {
        cudaError_t q = cudaEventQuery(e);
        if (q == cudaSuccess) return;
        if (q != cudaErrorNotReady) { s->error.store(static_cast<int>(q)); return; }
        std::this_thread::yield();
    }
}

static void service_loop(PsService* s) {
    if (cudaSetDevice(s->device) != cudaSuccess) { s->error.store(-1); return; }
    uint64_t n = 0;
    while (!s->stop.load(std::memory_order_relaxed) && s->error.load() == 0) {
        if (psb_push_apply(&s->desc, s->stream) != 0) { s->error.store(-2); break; }
        cudaEventRecord(s->ev[n & 1], s->stream);
        ++n;
        s->launches.store(n, std::memory_order_relaxed);
        // at most two launches in flight: the stream never runs dry, the queue never grows
        if (n >= 2) wait_event(s, s->ev[n & 1]);
        if (s->poll_us > 0) std::this_thread::sleep_for(std::chrono::microseconds(s->poll_us));
    }
    cudaStreamSynchronize(s->stream);
}

}  // namespace psb

extern "C" int psb_ps_service_start(const psb_push_desc_t* desc, int poll_us, void** out) {
    using namespace psb;
    if (!desc || !out) return fail(PSB_ERR_ARG, "psb_ps_service_start: null argument");
    if (!(desc->phases & PSB_PUSH_TRY_GRADS) || (desc->phases & (PSB_PUSH_WAIT_GRADS | PSB_PUSH_WAIT_PARAMS)))
        return fail(PSB_ERR_ARG, "psb_ps_service_start: the service only issues non-blocking launches "
                                 "(PSB_PUSH_TRY_GRADS without the WAIT phases)");
    PsService* s = new PsService();
    s->desc = *desc;
    s->poll_us = poll_us;
    PSB_CUDA_OK(cudaGetDevice(&s->device));
    int lo = 0, hi = 0;
    PSB_CUDA_OK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    PSB_CUDA_OK(cudaStreamCreateWithPriority(&s->stream, cudaStreamNonBlocking, hi));
    for (int i = 0; i < 2; ++i) PSB_CUDA_OK(cudaEventCreateWithFlags(&s->ev[i], cudaEventDisableTiming));
    s->th = std::thread(service_loop, s);
    *out = s;
    return PSB_OK;
}

extern "C" int psb_ps_service_stop(void* handle, uint64_t* launches) {
    using namespace psb;
    PsService* s = static_cast<PsService*>(handle);
    if (!s) return fail(PSB_ERR_ARG, "psb_ps_service_stop: null handle");
    s->stop.store(true);
    if (s->th.joinable()) s->th.join();
    if (launches) *launches = s->launches.load();
    const int err = s->error.load();
    for (int i = 0; i < 2; ++i)
        if (s->ev[i]) cudaEventDestroy(s->ev[i]);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
    if (err != 0) return fail(PSB_ERR_CUDA, "ps service thread ended with error %d", err);
    return PSB_OK;
}
