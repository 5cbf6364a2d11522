"""GPU training engine: one process per B200, the whole step as one CUDA graph.

``DeviceTrainer`` is the public API a user (and ``bench.py``) drives:

    tr = DeviceTrainer("lstm", rank, world, group, mode="sync")
    tr.initialize()                       # chief init_op (or restore) + first parameter pull
    out = tr.step({"x": x_pinned, "y": y_pinned})     # H2D, graph replay, loss D2H
    prev = tr.step_async(next_batch)      # same, pipelined: returns the PREVIOUS step's result

One step = the model's device plan (forward, losses, backward: hand-written sm_100a
kernels) followed by the fused gradient push (reduce-scatter + optimizer + all-gather in
one kernel over NVLink peer memory).  It plays the role of one ``sess.run([train_op, loss,
global_step], feed_dict=...)`` of the reference (``encoder:165``, ``gan:193``,
``lstm:128``).
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from ..models import make_model
from ..models.device_plans import make_plan
from ..ops import _lib
from ..placement import VariableRegistry, replica_device_setter
from ..store.nvl import CTRL_GLOBAL_STEP, CTRL_STATUS, CTRL_WORDS, NvlStore


class DeviceTrainer:
    def __init__(self, model_name: str, rank: int = 0, world: int = 1, group=None, mode: str = "sync",
                 batch: Optional[int] = None, seed: int = 0, use_graph: bool = True,
                 gan_update: str = "simultaneous", bcast: bool = True, plan_kwargs: Optional[dict] = None,
                 model_kwargs: Optional[dict] = None, partition: bool = True, timeout_s: float = 30.0,
                 keep_grads: bool = False, replicas_to_aggregate: int = 0, async_apply: str = "worker"):
        if not torch.cuda.is_available():
            raise RuntimeError("DeviceTrainer needs a CUDA device; the CPU path is psb200.session.Session")
        _lib.lib()   # fail loudly if the kernel library is missing
        self.rank, self.world, self.group, self.mode, self.seed = rank, world, group, mode, seed
        self.registry = VariableRegistry(replica_device_setter(
            worker_device="/job:worker/task:%d" % rank, ps_tasks=world))
        self.model = make_model(model_name, **(model_kwargs or {})).build(self.registry)
        self.store = NvlStore(self.model, self.registry, rank, world, group, mode=mode, bcast=bcast,
                              partition=partition, timeout_s=timeout_s, keep_grads=keep_grads,
                              replicas=replicas_to_aggregate, async_apply=async_apply)
        kw = dict(plan_kwargs or {})
        if self.model.name == "gan":
            kw.setdefault("seed", seed)
        self.plan = make_plan(self.model, self.store, batch, **kw)
        self.batch = self.plan.B
        self.gan_update = gan_update if self.model.name == "gan" else "simultaneous"
        self.use_graph = use_graph
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._loss_keys = list(self.plan.losses.keys())
        self._host_losses = torch.zeros(len(self._loss_keys), dtype=torch.float32).pin_memory()
        # one D2H brings back the push kernel's control block: status word and global_step
        # (ctrl and the worker-side wctrl are adjacent in the heap: 2 x CTRL_WORDS)
        self._host_ctrl = torch.zeros(2 * CTRL_WORDS, dtype=torch.int32).pin_memory()
        # one D2H per step for all losses: when the plan already keeps them next to each other the
        # copy reads them in place, otherwise they are gathered first
        self._dev_losses = self._adjacent_losses()
        self._gather_losses = self._dev_losses is None
        if self._gather_losses:
            self._dev_losses = torch.zeros(len(self._loss_keys), device=self.store.device)
        self.local_step = 0
        self._launches_per_step = 0
        self._pipe = None           # step_async state, built on first use

    def _adjacent_losses(self) -> Optional[torch.Tensor]:
        ts = [self.plan.losses[k] for k in self._loss_keys]
        base = ts[0]
        for i, t in enumerate(ts):
            if (t.dtype != torch.float32 or t.dim() != 0
                    or t.untyped_storage().data_ptr() != base.untyped_storage().data_ptr()
                    or t.storage_offset() != base.storage_offset() + i):
                return None
        return base.as_strided((len(ts),), (1,), base.storage_offset())

    # ------------------------------------------------------------------ lifecycle
    def initialize(self, values: Optional[Dict[str, torch.Tensor]] = None) -> None:
        """``init_op`` (or a restored checkpoint) -> owners' state, then the first pull.
        Collective: every rank calls it; all ranks derive identical values from ``seed``."""
        if self.graph is None and self.use_graph:
            self._warm_up()
        if values is None:
            values = self.registry.initial_values(self.seed)
        self.store.assign(values)
        self._barrier()
        self.store.refresh_copies()
        self._barrier()
        if self.use_graph and self.graph is None:
            self._capture()
        # the ps service thread (R-of-N sync, owner-side async) starts last: after the capture, with
        # every rank's compute copies in place
        self.store.start_service()

    def _barrier(self) -> None:
        torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.group)

    def _enqueue_step(self, push: bool = True) -> None:
        if not push:
            self.plan.forward_backward()
            return
        if self.gan_update == "alternating":
            self.plan.forward_backward(disc=True, gen=False)
            self.store.push(["train_disc"])
            self.plan.forward_backward(disc=False, gen=True)
            self.store.push(["train_gen"])
        else:
            self.plan.forward_backward()
            self.store.push()
        if self._gather_losses:
            for i, k in enumerate(self._loss_keys):
                self._dev_losses[i:i + 1].copy_(self.plan.losses[k].reshape(1))

    def _warm_up(self) -> None:
        """One eager step on scratch state: sets kernel attributes, fills the TMA descriptor
        cache, and proves the launch sequence before it is captured."""
        before = _lib.launch_count()
        # with a decoupled ps role the push is a wait for the ps services, which are not running yet
        self._enqueue_step(push=not self.store.decoupled)
        self._launches_per_step = _lib.launch_count() - before + (1 if self.store.decoupled else 0)
        self._barrier()

    def _capture(self) -> None:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            self._enqueue_step()
        self.graph = g
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ stepping
    def launches_per_step(self) -> int:
        """Number of this library's kernels in one training step."""
        return self._launches_per_step

    def step_device(self) -> None:
        """Run one step on whatever is in the plan's static input buffers (no H2D, no sync)."""
        if self.graph is not None:
            self.graph.replay()
        else:
            self._enqueue_step()
        self.local_step += 1

    def step(self, batch: Dict[str, torch.Tensor]) -> Dict[str, float]:
        """One training step end to end: H2D of the batch, the step, D2H of losses and
        ``global_step``."""
        for k, v in batch.items():
            if k in self.plan.inputs:
                self.plan.inputs[k].copy_(v.reshape(self.plan.inputs[k].shape), non_blocking=True)
        self.step_device()
        self._host_losses.copy_(self._dev_losses, non_blocking=True)
        self._host_ctrl.copy_(self._ctrl_src(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self._result(self._host_losses, self._host_ctrl)

    def _ctrl_src(self) -> torch.Tensor:
        """The control block that holds the authoritative global_step: this rank's in sync mode
        (lock-step replicas), the step owner's in async mode."""
        gs_rank = self.store.gs_owner if self.mode == "async" else self.rank
        return self.store.heap.tensor(self.store._ctrl_off, (2 * CTRL_WORDS,), torch.int32, gs_rank)

    def _result(self, losses: torch.Tensor, ctrl: torch.Tensor) -> Dict[str, float]:
        if self.mode == "sync" or self.store.gs_owner == self.rank:
            # a peer died / stalled: raise, do not hang
            self.store.check_status(max(int(ctrl[CTRL_STATUS]), int(ctrl[CTRL_WORDS + CTRL_STATUS])))
        out = {k: float(losses[i]) for i, k in enumerate(self._loss_keys)}
        out["global_step"] = int(ctrl[CTRL_GLOBAL_STEP])
        return out

    # ---- pipelined stepping: input prefetch on a copy stream, results read one step late
    def _pipe_state(self):
        if self._pipe is None:
            dev = self.store.device
            self._pipe = {
                "seq": 0, "stream": torch.cuda.Stream(device=dev),
                "stage": [{k: torch.empty_like(v) for k, v in self.plan.inputs.items()} for _ in range(2)],
                "losses": [torch.zeros(len(self._loss_keys), dtype=torch.float32).pin_memory() for _ in range(2)],
                "gstep": [torch.zeros(2 * CTRL_WORDS, dtype=torch.int32).pin_memory() for _ in range(2)],
                "h2d": [torch.cuda.Event() for _ in range(2)], "consumed": [torch.cuda.Event() for _ in range(2)],
                "done": [torch.cuda.Event() for _ in range(2)], "pending": [False, False],
            }
        return self._pipe

    def _pipe_result(self, k: int) -> Dict[str, float]:
        p = self._pipe
        p["done"][k].synchronize()
        p["pending"][k] = False
        return self._result(p["losses"][k], p["gstep"][k])

    def step_async(self, batch: Dict[str, torch.Tensor]) -> Optional[Dict[str, float]]:
        """Pipelined :meth:`step`: enqueue this batch's H2D (copy stream, into one of two staging
        sets), the step and the D2H of its results WITHOUT waiting for them, then return the result
        of the previously submitted step (``None`` on the first call; :meth:`drain` returns the
        last).  The host never idles the GPU between steps: the next batch's copy overlaps the
        running step.  ``batch`` must be pinned and stay unmodified until its result is returned."""
        p = self._pipe_state()
        k = p["seq"] & 1
        main = torch.cuda.current_stream()
        with torch.cuda.stream(p["stream"]):
            if p["seq"] >= 2:
                p["stream"].wait_event(p["consumed"][k])          # staging set k was read by step seq-2
            for name, v in batch.items():
                if name in self.plan.inputs:
                    p["stage"][k][name].copy_(v.reshape(self.plan.inputs[name].shape), non_blocking=True)
            p["h2d"][k].record(p["stream"])
        main.wait_event(p["h2d"][k])
        for name, t in p["stage"][k].items():
            if name in batch:
                self.plan.inputs[name].copy_(t, non_blocking=True)
        p["consumed"][k].record(main)
        self.step_device()
        p["losses"][k].copy_(self._dev_losses, non_blocking=True)
        p["gstep"][k].copy_(self._ctrl_src(), non_blocking=True)
        p["done"][k].record(main)
        p["pending"][k] = True
        p["seq"] += 1
        return self._pipe_result(1 - k) if p["pending"][1 - k] else None

    def drain(self) -> Optional[Dict[str, float]]:
        """Wait for the last :meth:`step_async` submission and return its result."""
        if self._pipe is None:
            return None
        k = (self._pipe["seq"] - 1) & 1
        return self._pipe_result(k) if self._pipe["pending"][k] else None

    def trace_step(self) -> Dict[str, float]:
        """One EAGER (un-graphed) step with CUDA-event timing per phase: forward+backward
        kernels vs the fused push/pull kernel.  Milliseconds per phase."""
        from ..utils.trace import PhaseTimer
        with PhaseTimer() as pt:
            with pt.phase("forward_backward"):
                self.plan.forward_backward()
            with pt.phase("push_pull"):
                self.store.push()
        self.local_step += 1
        return pt.report()

    def h2d_bytes_per_step(self, batch: Dict[str, torch.Tensor]) -> int:
        return sum(v.numel() * v.element_size() for k, v in batch.items() if k in self.plan.inputs)

    def d2h_bytes_per_step(self) -> int:
        return self._host_losses.numel() * 4 + self._host_ctrl.numel() * 4

    def global_step(self) -> int:
        return self.store.global_step()

    # ------------------------------------------------------------------ evaluation / state
    def accuracy(self, x: torch.Tensor, y: torch.Tensor) -> float:
        """LSTM: fraction of ``x`` classified correctly (``lstm:99-100,134-138``)."""
        if self.model.name != "lstm":
            raise ValueError("accuracy is defined for the LSTM classifier only")
        self.plan.inputs["x"].copy_(x.reshape(self.plan.inputs["x"].shape))
        self.plan.inputs["y"].copy_(y)
        self.plan.evaluate()
        torch.cuda.synchronize()
        return float(self.plan.correct.item()) / x.shape[0]

    def variables(self) -> Dict[str, torch.Tensor]:
        """Every checkpoint variable, read from its owner (CPU tensors keyed by SURVEY §2.4 names)."""
        return self.store.read_all()

    def close(self) -> None:
        self.graph = None
        self.store.close()
