"""The shared ``main()`` of the three entrypoints.

The reference's three scripts repeat the same ~110 lines around their model code (flags,
ClusterSpec/Server, ps branch, replica_device_setter scope, Supervisor, session config, training
loop with its log lines, done-queue shutdown: ``encoder:26-42,54-64,88-113,142-176``).  That
boilerplate is the framework; here it exists once.

    ps     : start the server, build its done-queue, dequeue ``--workers`` tokens, print
             ``ps %d received done %d`` each time, then ``ps %d: quitting``       (encoder:100-109)
    worker : build the variable set under the device setter, supervisor bring-up, loop
             ``while not sv.should_stop() and step < num_steps`` with the per-step line
             ``Global step %d Local step %d  AvgTime: %3.2fms``, ``Total Time: %3.2fs``, then
             one done token per ps                                                (encoder:110-176)
"""
from __future__ import annotations

import os
import time
from typing import Dict, Optional

import numpy as np
import torch

from .cluster import ClusterSpec, Server, device_filters
from .checkpoint import Saver
from .control.client import QueueTimeout
from .data import read_data_sets
from .flags import Flags
from .models import make_model
from .placement import VariableRegistry, replica_device_setter
from .store.tcp import TcpStore
from .supervisor import Supervisor


def run_ps(flags: Flags, cluster: ClusterSpec, server: Server) -> None:
    queue = server.done_queue(flags.task_index, capacity=flags.workers)
    for i in range(flags.workers):
        try:
            queue.dequeue(timeout=flags.ps_wait_timeout)
        except QueueTimeout:
            print("ps %d: timed out waiting for done %d" % (flags.task_index, i), flush=True)
            break
        print("ps %d received done %d" % (flags.task_index, i), flush=True)
    print("ps %d: quitting" % (flags.task_index), flush=True)


def _make_batch(model_name: str, mnist, batch_size: int, rng: np.random.RandomState) -> Dict[str, torch.Tensor]:
    bx, by = mnist.train.next_batch(batch_size)
    if model_name == "encoder":
        return {"x": torch.from_numpy(bx)}
    if model_name == "gan":
        # noise drawn on the host each step, float64 then cast at feed time (gan:189,134)
        z = rng.uniform(-1.0, 1.0, size=[batch_size, 100]).astype(np.float32)
        return {"x": torch.from_numpy(bx), "z": torch.from_numpy(z)}
    return {"x": torch.from_numpy(bx.reshape(batch_size, 28, 28)), "y": torch.from_numpy(by)}


def compute_params(params: Dict[str, torch.Tensor], dtype: str) -> Dict[str, torch.Tensor]:
    """What the worker math sees of the pulled variables.  ``--dtype=bf16`` rounds every weight
    matrix to bf16 (the master copy on the ps stays fp32), which is the operand precision of the
    sm_100a kernels: the CPU transport then reproduces the NVLink backend's numerics."""
    if dtype != "bf16":
        return params
    return {k: (v.to(torch.bfloat16).float() if v.dim() == 2 else v) for k, v in params.items()}


def _train_call_cpu(model, store: TcpStore, flags: Flags, batch, local_sync_step: int, sv: Supervisor):
    """One ``sess.run([train ops..., losses..., global_step], feed_dict)`` on the tcp transport."""
    pull = lambda: compute_params(store.pull(), flags.dtype)      # noqa: E731
    if flags.sync:
        params = pull()
        out = model.step_reference(params, batch)
        step = store.push_sync(out.grads, local_sync_step, sv.should_stop)
        return out.losses, step
    if model.name == "gan" and flags.gan_update == "alternating":
        losses = {}
        step = 0
        for op_name in ("train_disc", "train_gen"):
            out = model.step_reference(pull(), batch)
            op = [o for o in model.train_ops if o.name == op_name][0]
            step = store.push_async(op, out.grads[op.name])
            losses[op.loss] = out.losses[op.loss]
        return losses, step
    out = model.step_reference(pull(), batch)
    step = 0
    for op in model.train_ops:
        step = store.push_async(op, out.grads[op.name])
    return out.losses, step


def run_worker_cpu(model_name: str, flags: Flags, cluster: ClusterSpec, server: Server) -> dict:
    setter = replica_device_setter(worker_device="/job:worker/task:%d" % flags.task_index, cluster=cluster)
    registry = VariableRegistry(setter)
    model = make_model(model_name).build(registry)
    num_steps = flags.num_steps or model.num_steps
    batch_size = flags.batch_size or model.batch_size
    store = TcpStore(cluster, registry, model, num_workers=flags.workers)
    _ = device_filters(flags.task_index)        # a worker only ever opens connections to ps tasks
    sv = Supervisor(is_chief=(flags.task_index == 0), logdir=flags.model_dir, store=store,
                    registry=registry, model=model, saver=Saver(bundle_format=flags.checkpoint_format), save_model_secs=flags.save_model_secs,
                    recovery_wait_secs=flags.recovery_wait_secs, seed=flags.seed, sync=flags.sync,
                    replicas_to_aggregate=flags.replicas_to_aggregate, reinit_on_join=flags.reinit_on_join)
    mnist = read_data_sets(flags.data_dir, one_hot=True, synthetic=flags.synthetic, seed=flags.seed + flags.task_index)
    rng = np.random.RandomState(flags.seed * 7919 + flags.task_index)
    begin_time = time.time()
    result = {"local_steps": 0, "global_step": 0}
    with sv.managed_session() as sess:
        step = 0
        local_step = 0
        sync_step = store.global_step() if flags.sync else 0
        while not sv.should_stop() and step < num_steps:
            start_time = time.time()
            batch = _make_batch(model_name, mnist, batch_size, rng)
            losses, step = _train_call_cpu(model, store, flags, batch, sync_step, sv)
            if step < 0:
                break
            sync_step = step
            elapsed_time = time.time() - start_time
            if local_step % max(1, flags.log_every) == 0:
                print("Global step %d" % step, "Local step %d" % local_step,
                      " AvgTime: %3.2fms" % float(elapsed_time * 1000), flush=True)
            if model_name == "gan" and (step % 1000 == 0 or step == 1):
                print('Step %i: Generator Loss: %f, Discriminator Loss: %f'
                      % (step, float(losses["gen_loss"]), float(losses["disc_loss"])), flush=True)
            local_step += 1
        print("Total Time: %3.2fs" % float(time.time() - begin_time), flush=True)
        if model_name == "lstm":
            test_len = 128
            tx = torch.from_numpy(mnist.test.images[:test_len].reshape((-1, 28, 28)))
            ty = torch.from_numpy(mnist.test.labels[:test_len])
            acc = model.accuracy_reference(store.pull(), tx, ty)
            print("Test-Accuracy: %2.4f" % acc, flush=True)
            result["test_accuracy"] = acc
        result.update(local_steps=local_step, global_step=step, restored_from=sv.restored_from)
        if flags.sync:
            sv.request_stop()
        sv.final_checkpoint()       # while the ps tasks still hold the variables
        # Signal to ps shards that we are done
        store.enqueue_done(flags.workers)
    sv.stop()
    store.close()
    return result


def run_worker_nvl(model_name: str, flags: Flags, cluster: ClusterSpec, server: Server) -> dict:
    """Worker on the NVLink transport: rank = task_index, world = number of workers; every rank
    is also a ps shard, so ``ps_hosts`` only matters for the done-queues.  Same Supervisor-driven
    life cycle as the tcp transport (``encoder:147-176``): chief init/restore, non-chief wait,
    ``while not sv.should_stop() and step < num_steps``, checkpoint timer thread, done tokens."""
    import datetime
    import torch.distributed as dist
    from .parallel.session import NvlSession
    from .parallel.trainer import DeviceTrainer
    from .store.nvl import PeerTimeoutError
    world = cluster.num_tasks("worker")
    rank = flags.task_index
    if flags.reinit_on_join:
        raise SystemExit("--reinit_on_join reproduces a tcp-transport race (encoder:160); it has no meaning on "
                         "--backend=nvl, where all ranks initialise collectively")
    replicas = flags.replicas_to_aggregate if flags.sync else 0
    if replicas < 0 or replicas > world:
        raise SystemExit("--replicas_to_aggregate must be between 1 and the number of workers (%d)" % world)
    local = int(os.environ.get("LOCAL_RANK", rank % max(1, torch.cuda.device_count())))
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        host, port = cluster.task_address("worker", 0).rsplit(":", 1)
        rdzv = int(os.environ.get("PSB200_RDZV_PORT", int(port) + 1))
        dist.init_process_group("gloo", init_method="tcp://%s:%d" % (host or "127.0.0.1", rdzv),
                                rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=max(30.0, 4 * flags.peer_timeout_secs)))
        group = dist.group.WORLD
    tr = DeviceTrainer(model_name, rank, world, group, mode="sync" if flags.sync else "async",
                       batch=flags.batch_size or None, seed=flags.seed, gan_update=flags.gan_update,
                       plan_kwargs={"device_noise": False} if model_name == "gan" else None,
                       replicas_to_aggregate=replicas, async_apply=flags.async_apply,
                       timeout_s=flags.peer_timeout_secs)
    model = tr.model
    num_steps = flags.num_steps or model.num_steps
    sess_store = NvlSession(tr, group, is_chief=(rank == 0))
    sv = Supervisor(is_chief=(rank == 0), logdir=flags.model_dir, store=sess_store, registry=tr.registry,
                    model=model, saver=Saver(bundle_format=flags.checkpoint_format),
                    save_model_secs=flags.save_model_secs, recovery_wait_secs=flags.recovery_wait_secs,
                    seed=flags.seed, sync=False)        # sync aggregation lives in the push kernel, not in a chief thread
    mnist = read_data_sets(flags.data_dir, one_hot=True, synthetic=flags.synthetic, seed=flags.seed + rank)
    rng = np.random.RandomState(flags.seed * 7919 + rank)
    begin_time = time.time()
    result = {"local_steps": 0, "global_step": 0}
    failed = None
    with sv.managed_session():
        step, local_step = sess_store.global_step(), 0
        pinned = {k: torch.zeros_like(v, device="cpu").pin_memory() for k, v in tr.plan.inputs.items()}
        try:
            while not sv.should_stop() and step < num_steps:
                start_time = time.time()
                batch = _make_batch(model_name, mnist, tr.batch, rng)
                for k, v in batch.items():
                    pinned[k].copy_(v.reshape(pinned[k].shape))
                out = tr.step(pinned)
                step = out["global_step"]
                sess_store.note_step(step)
                elapsed_time = time.time() - start_time
                if local_step % max(1, flags.log_every) == 0:
                    print("Global step %d" % step, "Local step %d" % local_step,
                          " AvgTime: %3.2fms" % float(elapsed_time * 1000), flush=True)
                if model_name == "gan" and (step % 1000 == 0 or step == 1):
                    print('Step %i: Generator Loss: %f, Discriminator Loss: %f' % (step, out["gen_loss"], out["disc_loss"]),
                          flush=True)
                local_step += 1
        except PeerTimeoutError as e:
            # a peer is gone or stuck: stop cleanly instead of hanging in the kernel
            failed = str(e)
            print("worker %d: %s -- stopping" % (rank, failed), flush=True)
            sv.request_stop()
        print("Total Time: %3.2fs" % float(time.time() - begin_time), flush=True)
        result.update(local_steps=local_step, global_step=step, restored_from=sv.restored_from, failed=failed)
        if model_name == "lstm" and failed is None:
            tx = torch.from_numpy(mnist.test.images[:128].reshape((-1, 28, 28))).cuda()
            ty = torch.from_numpy(mnist.test.labels[:128]).cuda()
            acc = tr.accuracy(tx, ty)
            print("Test-Accuracy: %2.4f" % acc, flush=True)
            result["test_accuracy"] = acc
        if world > 1 and failed is None:
            try:
                dist.barrier(group=group)          # everybody has stopped stepping: the final state is final
            except Exception as e:                 # a dead peer must not keep the survivors here forever
                print("worker %d: final barrier failed (%s)" % (rank, type(e).__name__), flush=True)
        sess_store.note_step(tr.global_step() if failed is None else step)
        sv.final_checkpoint()
        if flags.ps_hosts:
            store = TcpStore(cluster, tr.registry, model, num_workers=flags.workers)
            store.enqueue_done(flags.workers)
            store.close()
    sv.stop()
    if world > 1 and failed is None:
        try:
            dist.barrier(group=group)
        except Exception:
            pass
    tr.close()
    if world > 1:
        dist.destroy_process_group()
    return result


def main(model_name: str, flags: Flags) -> Optional[dict]:
    """``main(_)`` of the reference scripts."""
    ps_hosts = flags.ps_hosts.split(",")
    worker_hosts = flags.worker_hosts.split(",")
    cluster = ClusterSpec({"ps": ps_hosts, "worker": worker_hosts})
    server = Server(cluster, job_name=flags.job_name, task_index=flags.task_index)
    try:
        if flags.job_name == "ps":
            run_ps(flags, cluster, server)
            return None
        elif flags.job_name == "worker":
            if flags.backend == "nvl":
                return run_worker_nvl(model_name, flags, cluster, server)
            return run_worker_cpu(model_name, flags, cluster, server)
    finally:
        server.stop()
    return None
