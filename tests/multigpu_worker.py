"""Run under torchrun (one rank per GPU): T3 checks of the NVLink store.

sync : after one step, every owner's state == oracle optimizer applied to the MEAN of all
       ranks' device gradients; every rank's arena holds identical copies; global_step
       advanced once per aggregate; a second run from the same seed is bit-identical.
async: global_step == (pushes of all ranks) * steps_per_run, parameters finite and moved.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from psb200.optim import KIND_ADAM, apply_reference  # noqa: E402
from psb200.parallel.trainer import DeviceTrainer  # noqa: E402


def batch_for(name, B, seed):
    g = torch.Generator().manual_seed(seed)
    if name == "encoder":
        return {"x": torch.rand(B, 784, generator=g)}
    if name == "gan":
        return {"x": torch.rand(B, 784, generator=g), "z": torch.rand(B, 100, generator=g) * 2 - 1}
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (B,), generator=g), 10).float()
    return {"x": torch.rand(B, 28, 28, generator=g), "y": y}


def _finish(what, rank, world):
    dist.barrier()
    if rank == 0:
        print("MULTIGPU_OK", what, world)
    dist.destroy_process_group()


def service_test(a, rank, world, group, run_once):
    """The ps role as a service thread (csrc/ps_service.cu).
    sync, R < N: one rank is deliberately slow; the others keep stepping at their own pace on
    aggregates that do not contain its gradient, its late gradients are dropped as stale, and its
    shard keeps serving because the service does not depend on its worker thread.
    async, owner-side apply: every gradient of every rank is applied exactly once by the owners."""
    import time
    tr, values, res = run_once(replicas_to_aggregate=a.replicas, async_apply=a.async_apply)
    assert tr.store.decoupled
    model = tr.model
    b = {k: v.pin_memory() for k, v in batch_for(a.model, tr.batch, 300 + rank).items()}
    steps = a.steps
    mine = 0
    last = res
    if a.mode == "sync" and a.slow_rank >= 0:
        # An aggregate needs R live workers, so the fast ranks keep stepping (a fixed count: with
        # R = N-1 they move in lock-step) for longer than the slow rank needs for its four steps.
        t0 = time.time()
        n_mine = 4 if rank == a.slow_rank else 3000
        for i in range(n_mine):
            if rank == a.slow_rank:
                time.sleep(0.05)
            last = tr.step(b)
            mine += 1
        elapsed = time.time() - t0
        if rank != a.slow_rank:
            # not held back by the slow rank, which needs >= 50 ms per step
            assert elapsed < 3000 * 0.002, (rank, elapsed)
            assert last["global_step"] >= (mine + 1) * model.steps_per_run(), last
    else:
        for i in range(steps):
            last = tr.step(b)
            mine += 1
    torch.cuda.synchronize()
    dist.barrier(group=group)
    time.sleep(0.5)                      # let the services drain what is still on the pads
    dist.barrier(group=group)
    gs = tr.global_step()
    alls = [None] * world
    dist.all_gather_object(alls, (gs, mine), group=group)
    total = sum(m for _, m in alls) + world          # + the step run_once made on every rank
    if a.mode == "async":
        assert gs == total * model.steps_per_run(), (rank, gs, alls)
    else:
        # every owner applied the same aggregates; fewer aggregates than gradients were offered when
        # somebody was late (stale gradients dropped) -- and never more
        assert all(g == gs for g, _ in alls), alls
        assert gs <= total * model.steps_per_run(), (gs, total)
        if a.slow_rank >= 0:
            fast = max(m for _, m in alls) + 1
            assert gs >= fast * model.steps_per_run(), (gs, fast)
            if a.replicas == world - 1:
                # every aggregate was made of the R fast ranks' gradients: the slow rank's were all stale
                assert gs < (fast + 3) * model.steps_per_run(), (gs, fast)
    got = tr.variables()
    for n in model.trainable_names():
        assert torch.isfinite(got[n]).all(), n
        assert (got[n] - values[n]).abs().max().item() > 0, n
    # quiescent: every rank's arena holds the same compute copies, and they follow the masters
    for n in model.trainable_names():
        spec = tr.registry[n]
        mine_c = (tr.store.w(n).t.clone() if len(spec.shape) == 2 else tr.store.vec(n).clone()).cpu()
        allc = [None] * world
        dist.all_gather_object(allc, mine_c, group=group)
        for c in allc:
            assert torch.equal(c, allc[0]), (rank, n)
        if len(spec.shape) == 2:
            m = got[n]
            assert (tr.store.w(n).float().cpu() - m).abs().max().item() <= 8e-3 * m.abs().max().item() + 1e-6, n
    dist.barrier(group=group)
    tr.close()
    _finish("%s service R=%d %s" % (a.model, a.replicas, a.async_apply), rank, world)


def stall_test(a, rank, world, group, run_once):
    """A peer that stops stepping must not hang the others: the push kernel's waits end on their
    deadline (or on the host's abort word) and the trainer raises PeerTimeoutError."""
    import threading
    import time
    from psb200.store.nvl import PeerTimeoutError
    stalled = world - 1
    for phase in ("deadline", "abort"):
        tr, _, _ = run_once(timeout_s=1.5 if phase == "deadline" else 60.0)
        b = {k: v.pin_memory() for k, v in batch_for(a.model, tr.batch, 300 + rank).items()}
        tr.step(b)
        dist.barrier(group=group)
        if rank == stalled:
            time.sleep(6.0)                      # alive, but not stepping
        else:
            if phase == "abort":
                threading.Timer(1.0, tr.store.abort).start()
            t0 = time.time()
            try:
                tr.step(b)
                raise AssertionError("the step should not have completed without rank %d" % stalled)
            except PeerTimeoutError as e:
                dt = time.time() - t0
                assert dt < 5.0, dt
                assert ("aborted" in str(e)) == (phase == "abort"), str(e)
        dist.barrier(group=group)
        tr.close()
        dist.barrier(group=group)
    _finish("stall", rank, world)


def noalloc_test(a, rank, world, group, run_once):
    """psb_push_apply_noalloc with world > 1: the staged pad-pointer table and the kept step token
    are exercised across GPUs.  Two steps through the ordinary entry point on one trainer, the same
    two steps through the no-alloc entry point on a second trainer (explicit, increasing step ids:
    the entry point's own token must not be mixed with psb_push_apply's on the same pads) -- the
    resulting variables must be identical bit for bit."""
    import ctypes
    from psb200.ops import _lib
    import psb200.store.nvl as nvl
    L = _lib.lib()
    fn = L.psb_push_apply_noalloc
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                   ctypes.c_void_p, ctypes.POINTER(nvl.PushDesc), ctypes.c_void_p]
    states = []
    for use_noalloc in (False, True):
        tr = DeviceTrainer(a.model, rank, world, group, mode="sync", seed=5, use_graph=False)
        tr.initialize(tr.registry.initial_values(5))
        st = tr.store
        for k, v in batch_for(a.model, tr.batch, 100 + rank).items():
            tr.plan.inputs[k].copy_(v.reshape(tr.plan.inputs[k].shape))
        for call in range(2):
            tr.plan.forward_backward()
            if not use_noalloc:
                st.push()
            else:
                segs, n, gx = st._build_segs(owned_only=True, peer_state=False, kind_override=None)
                assert n <= 64, n
                host_segs = (nvl.PushSeg * n).from_buffer_copy(bytes(segs.cpu().numpy().tobytes())[:n * ctypes.sizeof(nvl.PushSeg)])
                P = ctypes.c_void_p * world
                h = st.heap
                tabs = [P(*[h.peer_ptr(r, off) for r in range(world)])
                        for off in (st._grad_off, st._arena_off, st._farena_off, st._pad_off)]
                d = st._desc(segs, n, gx, True, 1, True)
                d.step_id = 1000 + call
                d.prof = None
                _lib.check(fn(host_segs, n, tabs[0], tabs[1], tabs[2], tabs[3], ctypes.byref(d),
                              torch.cuda.current_stream().cuda_stream), "psb_push_apply_noalloc")
            torch.cuda.synchronize()
            dist.barrier(group=group)
        states.append(tr.variables())
        tr.close()
        dist.barrier(group=group)
    for k in states[0]:
        assert torch.equal(states[0][k], states[1][k]), (rank, k)
    assert int(states[1][[s.name for s in tr.registry if s.kind == "step"][0]]) == 2
    _finish("noalloc", rank, world)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="lstm")
    ap.add_argument("--mode", default="sync")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--pull_from_owner", type=int, default=0)
    ap.add_argument("--stress", type=int, default=0)
    ap.add_argument("--replicas", type=int, default=0, help="sync: aggregate R < world gradients (ps service)")
    ap.add_argument("--async_apply", default="worker")
    ap.add_argument("--slow_rank", type=int, default=-1)
    ap.add_argument("--stall", type=int, default=0, help="failure containment: the last rank stops stepping")
    ap.add_argument("--noalloc", type=int, default=0, help="psb_push_apply_noalloc across ranks")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    group = dist.new_group(backend="gloo")
    kw = {"device_noise": False} if a.model == "gan" else {}
    if a.model == "lstm" and a.pull_from_owner:
        kw["pull_from_owner"] = True

    def run_once(keep_grads=False, **extra):
        tr = DeviceTrainer(a.model, rank, world, group, mode=a.mode, seed=5, plan_kwargs=kw,
                           bcast=not a.pull_from_owner, keep_grads=keep_grads, **extra)
        values = tr.registry.initial_values(5)
        if a.model == "encoder":
            for s in tr.registry.trainable():
                values[s.name] = values[s.name] * 0.05
        tr.initialize(values)
        b = {k: v.pin_memory() for k, v in batch_for(a.model, tr.batch, 100 + rank).items()}
        res = tr.step(b)
        torch.cuda.synchronize()
        dist.barrier(group=group)
        return tr, values, res

    if a.stall:
        stall_test(a, rank, world, group, run_once)
        return
    if a.noalloc:
        noalloc_test(a, rank, world, group, run_once)
        return
    if a.replicas or a.async_apply == "owner":
        service_test(a, rank, world, group, run_once)
        return
    if a.stress:
        # T4: signal-pad protocol under randomised host delays.  Run A steps in lock-step through the
        # blocking API; run B lets every rank stall for a random time before each submission and
        # mixes the pipelined API in.  The device-side step tokens must make both runs bit-identical.
        import random
        import time
        batches = [{k: v.pin_memory() for k, v in batch_for(a.model, 128 if a.model != "encoder" else 256, 400 + 7 * i + rank).items()}
                   for i in range(4)]
        states = []
        for jitter in (False, True):
            tr, _, _ = run_once()
            rnd = random.Random(1234 + rank)
            for i in range(a.stress):
                if jitter:
                    time.sleep(rnd.random() * 0.003 if rnd.random() < 0.5 else 0.0)
                    if rnd.random() < 0.5:
                        tr.step_async(batches[i % 4])
                    else:
                        tr.drain()
                        tr.step(batches[i % 4])
                else:
                    tr.step(batches[i % 4])
            tr.drain()
            torch.cuda.synchronize()
            dist.barrier(group=group)
            assert tr.global_step() == (a.stress + 1) * tr.model.steps_per_run(), tr.global_step()
            states.append(tr.variables())
            tr.close()
        for k in states[0]:
            assert torch.equal(states[0][k], states[1][k]), (rank, k, "host jitter changed the result")
        dist.barrier()
        if rank == 0:
            print("MULTIGPU_OK", a.model, "stress", world)
        dist.destroy_process_group()
        return

    # first run keeps the gradients readable (the plan clears them itself); the re-run below lets the
    # push kernel clear the consumed gradients in every rank's buffer: both must agree bit for bit
    tr, values, res = run_once(keep_grads=True)
    model = tr.model
    if a.mode == "sync":
        assert res["global_step"] == model.steps_per_run(), res
        # gather every rank's device gradients
        mine = {n: tr.store.grad(n).cpu() for n in model.trainable_names()}
        allg = [None] * world
        dist.all_gather_object(allg, mine, group=group)
        mean = {n: sum(g[n] for g in allg) / world for n in mine}
        want = {k: v.clone() for k, v in values.items()}
        for op in model.train_ops:
            opt = op.optimizer
            bp = (float(values[opt.nonslots[0]]), float(values[opt.nonslots[1]])) if opt.kind == KIND_ADAM else None
            for n in op.var_list:
                apply_reference(opt, want[n], mean[n], [want[s] for s in opt.slots[n]], bp)
            if opt.kind == KIND_ADAM:
                want[opt.nonslots[0]] = values[opt.nonslots[0]] * opt.hyper["beta1"]
                want[opt.nonslots[1]] = values[opt.nonslots[1]] * opt.hyper["beta2"]
        got = tr.variables()
        for s in tr.registry:
            if s.kind == "step":
                assert int(got[s.name]) == model.steps_per_run()
                continue
            x, y = got[s.name].float(), want[s.name].float()
            assert torch.allclose(x, y, rtol=3e-4, atol=1e-7), (rank, s.name, (x - y).abs().max().item())
        # all arenas identical (push-mode all-gather) or owner-only
        if not a.pull_from_owner:
            for n in model.trainable_names():
                spec = tr.registry[n]
                mine_c = (tr.store.w(n).t.clone() if len(spec.shape) == 2 else tr.store.vec(n).clone()).cpu()
                allc = [None] * world
                dist.all_gather_object(allc, mine_c, group=group)
                for c in allc:
                    assert torch.equal(c, allc[0]), (rank, n)
        # more steps, then determinism across a full re-run
        b = {k: v.pin_memory() for k, v in batch_for(a.model, tr.batch, 200 + rank).items()}
        for _ in range(a.steps):
            last = tr.step(b)
        assert last["global_step"] == (a.steps + 1) * model.steps_per_run(), last
        state1 = tr.variables()
        tr.close()
        tr2, _, _ = run_once()
        for _ in range(a.steps):
            tr2.step(b)
        state2 = tr2.variables()
        for k in state1:
            assert torch.equal(state1[k], state2[k]), (rank, k, "sync mode must be bit-deterministic")
        tr2.close()
    else:
        b = {k: v.pin_memory() for k, v in batch_for(a.model, tr.batch, 300 + rank).items()}
        for _ in range(a.steps):
            tr.step(b)
        torch.cuda.synchronize()
        dist.barrier(group=group)
        gs = tr.global_step()
        assert gs == world * (a.steps + 1) * model.steps_per_run(), (rank, gs)
        got = tr.variables()
        for n in model.trainable_names():
            assert torch.isfinite(got[n]).all(), n
            assert (got[n] - values[n]).abs().max().item() > 0, n
        dist.barrier(group=group)
        tr.close()
    dist.barrier()
    if rank == 0:
        print("MULTIGPU_OK", a.model, a.mode, world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
