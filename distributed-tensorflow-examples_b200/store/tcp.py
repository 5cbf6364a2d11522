"""TCP parameter store: variables live in ps processes (native runtime), workers reach them
over sockets.  This is the transport of the "1 ps + 2 workers on CPU localhost" configuration
and the one that maps one-to-one onto the reference's process model.

Per training call a worker does what TF's partitioned graph does (``encoder:165``):
pull every variable (X1), compute gradients locally, push each gradient to the ps that owns
the variable where the optimizer update runs (X2/O1-O3), bump ``global_step`` (O4).
"""
from __future__ import annotations

import threading
import time
from typing import Dict, Iterable, List, Optional, Sequence

import torch

from ..cluster import ClusterSpec
from ..control.client import PsClient, PsError, QueueTimeout
from ..models.base import Model, TrainOp
from ..optim import KIND_ADAM
from ..placement import VariableRegistry, VariableSpec

SYNC_TOKEN_QUEUE = "sync_token_q"


class TcpStore:
    def __init__(self, cluster: ClusterSpec, registry: VariableRegistry, model: Model,
                 num_workers: Optional[int] = None):
        self.cluster, self.registry, self.model = cluster, registry, model
        self.ps_addrs = cluster.job_tasks("ps")
        self.num_workers = num_workers or cluster.num_tasks("worker")
        self._local = threading.local()

    # one connection per (thread, ps task)
    def _client(self, ps_task: int) -> PsClient:
        cache = getattr(self._local, "clients", None)
        if cache is None:
            cache = self._local.clients = {}
        if ps_task not in cache:
            cache[ps_task] = PsClient(self.ps_addrs[ps_task])
        return cache[ps_task]

    def _of(self, spec: VariableSpec) -> PsClient:
        return self._client(spec.ps_task)

    # ------------------------------------------------------------------ graph materialisation
    def declare_all(self) -> None:
        for s in self.registry:
            self._of(s).var_declare(s.name, s.shape, s.dtype)

    def initialize(self, values: Dict[str, torch.Tensor]) -> None:
        """``init_op`` / ``saver.restore``: assign every variable on its ps."""
        for s in self.registry:
            if s.name in values:
                self._of(s).var_assign(s.name, values[s.name].to(s.dtype))

    def uninitialized(self) -> int:
        """``ready_op``: number of declared variables without a value, over all ps tasks."""
        return sum(self._client(i).uninitialized_count() for i in range(len(self.ps_addrs)))

    # ------------------------------------------------------------------ pull
    def read(self, name: str) -> torch.Tensor:
        s = self.registry[name]
        return self._of(s).var_read(s.name, s.shape, s.dtype)

    def pull(self, names: Optional[Iterable[str]] = None) -> Dict[str, torch.Tensor]:
        names = list(names) if names is not None else [s.name for s in self.registry.trainable()]
        return {n: self.read(n) for n in names}

    def read_all(self) -> Dict[str, torch.Tensor]:
        return {s.name: self.read(s.name) for s in self.registry}

    def global_step(self) -> int:
        return int(self.read(self.model.global_step.name))

    # ------------------------------------------------------------------ async push (Hogwild)
    def _beta_powers(self, op: TrainOp):
        if op.optimizer.kind != KIND_ADAM:
            return (0.0, 0.0)
        return tuple(float(self.read(n)) for n in op.optimizer.nonslots)

    def push_async(self, op: TrainOp, grads: Dict[str, torch.Tensor]) -> int:
        """One ``minimize`` train op: apply each gradient on its ps without locking, run the
        optimizer's finish step, ``AssignAdd(global_step, 1)``.  Returns the new step."""
        opt = op.optimizer
        hv = opt.hyper_vector()
        bp = self._beta_powers(op)
        for n in op.var_list:
            s = self.registry[n]
            self._of(s).apply(n, grads[n], opt.kind, hv, opt.slots[n], bp)
        if opt.kind == KIND_ADAM:
            b1, b2 = self.registry[opt.nonslots[0]], self.registry[opt.nonslots[1]]
            self._of(b1).var_scale_f32(b1.name, opt.hyper["beta1"])
            self._of(b2).var_scale_f32(b2.name, opt.hyper["beta2"])
        gs = self.model.global_step
        if op.increments_global_step:
            return self._of(gs).var_add_i32(gs.name, 1)
        return self.global_step()

    # ------------------------------------------------------------------ sync push
    def sync_setup(self, is_chief: bool) -> None:
        for op in self.model.train_ops:
            for n in op.var_list:
                self._of(self.registry[n]).acc_declare(n)
        gs = self.model.global_step
        self._of(gs).queue_create(SYNC_TOKEN_QUEUE, max(4 * self.num_workers, 16))

    def push_sync(self, grads_by_op: Dict[str, Dict[str, torch.Tensor]], local_step: int,
                  should_stop=None) -> int:
        """Send (grad, local_step) to every accumulator, then block on a token; the token is
        the new ``global_step`` and becomes the worker's ``local_step``."""
        for op in self.model.train_ops:
            for n in op.var_list:
                self._of(self.registry[n]).acc_apply(n, grads_by_op[op.name][n], local_step)
        gs = self.model.global_step
        while True:
            try:
                return self._of(gs).queue_dequeue(SYNC_TOKEN_QUEUE, timeout=0.5)
            except QueueTimeout:
                if should_stop is not None and should_stop():
                    return -1

    def chief_aggregate_once(self, replicas_to_aggregate: int) -> int:
        """The chief's queue-runner body: wait for R gradients per variable, apply their mean
        once, advance ``global_step`` once per train op, release one token per worker."""
        new_step = None
        for op in self.model.train_ops:
            opt = op.optimizer
            hv = opt.hyper_vector()
            bp = self._beta_powers(op)
            for n in op.var_list:
                self._of(self.registry[n]).acc_take_apply(n, replicas_to_aggregate, opt.kind, hv,
                                                          opt.slots[n], bp)
            if opt.kind == KIND_ADAM:
                b1, b2 = self.registry[opt.nonslots[0]], self.registry[opt.nonslots[1]]
                self._of(b1).var_scale_f32(b1.name, opt.hyper["beta1"])
                self._of(b2).var_scale_f32(b2.name, opt.hyper["beta2"])
            gs = self.model.global_step
            new_step = self._of(gs).var_add_i32(gs.name, 1) if op.increments_global_step else new_step
        for op in self.model.train_ops:
            for n in op.var_list:
                self._of(self.registry[n]).acc_set_step(n, new_step)
        gs = self.model.global_step
        self._of(gs).queue_enqueue_many(SYNC_TOKEN_QUEUE, [new_step] * self.num_workers)
        return new_step

    def release_sync_waiters(self) -> None:
        """Chief shutdown in sync mode: no aggregate will ever be applied again, so hand every worker
        a sentinel token (-1).  A worker blocked in :meth:`push_sync` returns -1, leaves its loop and
        sends its done token -- otherwise the ps tasks would wait for it forever."""
        gs = self.model.global_step
        try:
            self._of(gs).queue_enqueue_many(SYNC_TOKEN_QUEUE, [-1] * self.num_workers)
        except Exception:
            pass

    # ------------------------------------------------------------------ done queues (C6)
    def enqueue_done(self, capacity: int) -> None:
        """``for op in enq_ops: sess.run(op)`` — one token onto every ps task's queue
        (``encoder:142-145,172-174``)."""
        for i in range(len(self.ps_addrs)):
            c = self._client(i)
            name = "done_queue%d" % i
            c.queue_create(name, capacity)
            c.queue_enqueue(name, 1)

    def stats(self) -> List[dict]:
        return [self._client(i).stats() for i in range(len(self.ps_addrs))]

    def close(self) -> None:
        cache = getattr(self._local, "clients", {})
        for c in cache.values():
            c.close()
        cache.clear()
