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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="lstm")
    ap.add_argument("--mode", default="sync")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--pull_from_owner", type=int, default=0)
    ap.add_argument("--stress", type=int, default=0)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    group = dist.new_group(backend="gloo")
    kw = {"device_noise": False} if a.model == "gan" else {}
    if a.model == "lstm" and a.pull_from_owner:
        kw["pull_from_owner"] = True

    def run_once(keep_grads=False):
        tr = DeviceTrainer(a.model, rank, world, group, mode=a.mode, seed=5, plan_kwargs=kw,
                           bcast=not a.pull_from_owner, keep_grads=keep_grads)
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
