"""NVLink parameter store: every GPU is a ps shard.

Replaces, on one 8xB200 box, the ps processes of the reference and the gRPC traffic to
them (``replica_device_setter`` placement ``encoder:111-113``; per-step variable reads and
``Apply*`` updates ``encoder:137,165``, ``gan:158-159,193``, ``lstm:97,128``):

* variable ``i`` (creation order) is owned by rank ``i mod world`` — fp32 master weights,
  optimizer slots, ``global_step`` and Adam's beta powers live in the owner's HBM;
* every rank keeps its gradients in one flat fp32 buffer inside a CUDA-IPC symmetric heap;
* **push** = one ``psb_push_apply`` kernel per step: the owner P2P-loads all ranks'
  gradient slices, averages, applies SGD/RMSProp/Adam and stores the new bf16/fp32 compute
  copies (``sync``), or a worker applies its own gradient straight onto the owner's master
  copy through peer pointers / peer atomics (``async``);
* **pull** = the compute copies are either written into every rank's arena by the push
  kernel itself (push-mode all-gather, ``bcast=True``) or read from the owner's arena by
  the consuming kernel (``bcast=False``, e.g. the LSTM kernel's weight prologue).
"""
from __future__ import annotations

import ctypes
from ctypes import POINTER, Structure, c_float, c_int, c_int32, c_int64, c_uint32, c_uint64, c_void_p
from typing import Dict, List, Optional

import torch

from ..models.base import Model
from ..ops import _lib
from ..ops.cuda import BMat, round_up
from ..optim import KIND_ADAM, Optimizer
from ..parallel.peer import SymmetricHeap
from ..placement import VariableRegistry, VariableSpec


class PushSeg(Structure):
    _fields_ = [("grad_offset", c_uint64), ("numel", c_uint32), ("cols", c_uint32),
                ("opt_kind", c_int32), ("lr", c_float), ("h1", c_float), ("h2", c_float),
                ("eps", c_float), ("master", c_void_p), ("slot0", c_void_p), ("slot1", c_void_p),
                ("beta_pow", c_void_p), ("w_off", c_int64), ("ld_w", c_int32), ("wt_off", c_int64),
                ("ld_wt", c_int32), ("f_off", c_int64), ("home", c_int32), ("bcast", c_int32)]


class BetaUpdate(Structure):
    _fields_ = [("beta_pow", c_void_p), ("beta1", c_float), ("beta2", c_float)]


class PushDesc(Structure):
    _fields_ = [("segs", c_void_p), ("nseg", c_int), ("grad_ptrs", c_void_p), ("arena_ptrs", c_void_p),
                ("farena_ptrs", c_void_p), ("pad_ptrs", c_void_p), ("world", c_int), ("rank", c_int),
                ("n_arenas", c_int), ("inv_replicas", c_float), ("step_id", c_uint32),
                ("local_ctrl", c_void_p), ("beta", BetaUpdate * 2), ("nbeta", c_int),
                ("global_step", c_void_p), ("step_delta", c_int), ("async_sgd_atomic", c_int),
                ("signal_params", c_int), ("blocks_per_seg", c_int), ("prof", c_void_p)]


_lib.register_signature("psb_push_apply", [POINTER(PushDesc), c_void_p])
_lib.register_signature("psb_wait_params", [c_void_p, c_int, c_uint32, c_void_p])


class _VarLayout:
    __slots__ = ("spec", "rows", "cols", "grad_off", "master_off", "slot_offs", "w_off", "ld_w",
                 "wt_off", "ld_wt", "f_off", "owner")


class NvlStore:
    """Parameter shards + gradient buffers of one rank, and the push launcher."""

    PAD_WORDS = 64

    def __init__(self, model: Model, registry: VariableRegistry, rank: int = 0, world: int = 1,
                 group=None, mode: str = "sync", bcast: bool = True, device=None):
        if mode not in ("sync", "async"):
            raise ValueError("mode must be 'sync' or 'async'")
        self.model, self.registry = model, registry
        self.rank, self.world, self.group, self.mode, self.bcast = rank, world, group, mode, bcast
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.layout: Dict[str, _VarLayout] = {}
        self._opt_of: Dict[str, Optimizer] = {}
        for op in model.train_ops:
            for n in op.var_list:
                self._opt_of[n] = op.optimizer
        self._plan_heap()
        self.heap = SymmetricHeap(self._heap_bytes, rank, world, group, self.device)
        assert self.heap.alloc(self._heap_bytes, 256) == 0
        h = self.heap
        self.pad = h.tensor(self._pad_off, (self.PAD_WORDS,), torch.int32)
        self.ctrl = h.tensor(self._ctrl_off, (4,), torch.int32)
        self.grads = h.tensor(self._grad_off, (self._grad_elems,), torch.float32)
        self.global_step_t = h.tensor(self._gs_off, (1,), torch.int32)
        self.gs_owner = model.global_step.ps_task % world
        self.prof = torch.zeros(8, dtype=torch.int64, device=self.device)
        self._tables()
        self._seg_cache = {}
        self._segs_refresh = self._build_segs(owned_only=True, peer_state=False, kind_override=-1)
        self._push_count = 0

    # ------------------------------------------------------------------ layout
    def _plan_heap(self) -> None:
        cur = 0

        def take(nbytes: int, align: int = 256) -> int:
            nonlocal cur
            off = (cur + align - 1) // align * align
            cur = off + nbytes
            return off

        self._pad_off = take(self.PAD_WORDS * 4)
        self._ctrl_off = take(16)
        self._gs_off = take(16)
        # Adam beta powers: one {b1p, b2p} pair per Adam optimizer
        self._beta_off: Dict[int, int] = {}
        for k, op in enumerate(self.model.train_ops):
            if op.optimizer.kind == KIND_ADAM:
                self._beta_off[k] = take(16)
        vars_ = [self.registry[n] for op in self.model.train_ops for n in op.var_list]
        vars_.sort(key=lambda s: s.index)
        # flat gradient buffer
        g = 0
        for s in vars_:
            L = _VarLayout()
            L.spec = s
            L.rows, L.cols = (s.shape[0], s.shape[1]) if len(s.shape) == 2 else (1, s.numel)
            L.owner = s.ps_task % self.world
            L.grad_off = g
            g += round_up(s.numel, 4)
            self.layout[s.name] = L
        self._grad_elems = g
        self._grad_off = take(g * 4)
        # fp32 copies of 1-D variables, bf16 copies of 2-D variables
        f = 0
        b = 0
        for s in vars_:
            L = self.layout[s.name]
            L.w_off = L.wt_off = L.f_off = -1
            L.ld_w = L.ld_wt = 0
            if len(s.shape) == 2:
                L.ld_w, L.ld_wt = round_up(L.cols, 8), round_up(L.rows, 8)
                L.w_off = b
                b += round_up(L.rows * L.ld_w, 128)
                L.wt_off = b
                b += round_up(L.cols * L.ld_wt, 128)
            else:
                L.f_off = f
                f += round_up(s.numel, 4)
        self._farena_elems, self._arena_elems = max(f, 4), max(b, 128)
        self._farena_off = take(self._farena_elems * 4)
        self._arena_off = take(self._arena_elems * 2, 1024)
        for s in vars_:
            L = self.layout[s.name]
            L.master_off = take(s.numel * 4)
            L.slot_offs = [take(s.numel * 4) for _ in self._opt_of[s.name].slots.get(s.name, [])]
        self._heap_bytes = (cur + 4095) // 4096 * 4096

    def _tables(self) -> None:
        h = self.heap
        self._t_grads = h.ptr_table(self._grad_off)
        self._t_arena = h.ptr_table(self._arena_off)
        self._t_farena = h.ptr_table(self._farena_off)
        self._t_pad = h.ptr_table(self._pad_off)
        # single-entry tables for the async launch (world == 1 from the kernel's viewpoint)
        self._t_grads_self = torch.tensor([h.ptr(self._grad_off)], dtype=torch.int64, device=self.device)

    def _build_segs(self, owned_only: bool, peer_state: bool, kind_override: Optional[int],
                    op_names: Optional[List[str]] = None):
        h = self.heap
        segs = []
        wanted = None
        if op_names is not None:
            wanted = {n for op in self.model.train_ops if op.name in op_names for n in op.var_list}
        for name, L in self.layout.items():
            if owned_only and L.owner != self.rank:
                continue
            if wanted is not None and name not in wanted:
                continue
            opt = self._opt_of[name]
            hv = opt.hyper_vector()
            s = PushSeg()
            s.grad_offset, s.numel, s.cols = L.grad_off, L.spec.numel, L.cols
            s.opt_kind = opt.kind if kind_override is None else kind_override
            s.lr, s.h1, s.h2, s.eps = hv
            owner = L.owner if peer_state else self.rank
            s.master = h.peer_ptr(owner, L.master_off)
            s.slot0 = h.peer_ptr(owner, L.slot_offs[0]) if len(L.slot_offs) > 0 else None
            s.slot1 = h.peer_ptr(owner, L.slot_offs[1]) if len(L.slot_offs) > 1 else None
            if opt.kind == KIND_ADAM:
                k = [i for i, op in enumerate(self.model.train_ops) if op.optimizer is opt][0]
                brank = self._beta_owner(k) if peer_state else self.rank
                s.beta_pow = h.peer_ptr(brank, self._beta_off[k])
            s.w_off, s.ld_w, s.wt_off, s.ld_wt, s.f_off = L.w_off, L.ld_w, L.wt_off, L.ld_wt, L.f_off
            s.home = L.owner if peer_state else self.rank
            s.bcast = 1 if self.bcast else 0
            segs.append(s)
        arr = (PushSeg * max(len(segs), 1))(*segs)
        raw = bytes(arr) if segs else b"\0" * ctypes.sizeof(PushSeg)
        dev = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.device)
        return dev, len(segs)

    def _beta_owner(self, op_index: int) -> int:
        opt = self.model.train_ops[op_index].optimizer
        return self.registry[opt.nonslots[0]].ps_task % self.world

    # ------------------------------------------------------------------ views for the model plans
    def w(self, name: str, rank: Optional[int] = None) -> BMat:
        """Natural-layout bf16 copy ``[rows, cols]`` (this rank's arena, or a peer's)."""
        L = self.layout[name]
        m = BMat.__new__(BMat)
        m.rows, m.cols = L.rows, L.cols
        m.remote = rank is not None and rank != self.rank
        m.t = self.heap.tensor(self._arena_off + 2 * L.w_off, (L.rows, L.ld_w), torch.bfloat16, rank)
        return m

    def wt(self, name: str, rank: Optional[int] = None) -> BMat:
        """Transposed bf16 copy ``[cols, rows]``."""
        L = self.layout[name]
        m = BMat.__new__(BMat)
        m.rows, m.cols = L.cols, L.rows
        m.remote = rank is not None and rank != self.rank
        m.t = self.heap.tensor(self._arena_off + 2 * L.wt_off, (L.cols, L.ld_wt), torch.bfloat16, rank)
        return m

    def vec(self, name: str, rank: Optional[int] = None) -> torch.Tensor:
        """fp32 compute copy of a 1-D variable."""
        L = self.layout[name]
        return self.heap.tensor(self._farena_off + 4 * L.f_off, (L.spec.numel,), torch.float32, rank)

    # where a consumer kernel should read a compute copy from: this rank's arena when the push
    # kernel broadcasts (bcast=True), otherwise the owner's arena through its peer mapping
    def _src_rank(self, name: str) -> Optional[int]:
        return None if self.bcast else self.layout[name].owner

    def w_src(self, name: str) -> BMat:
        return self.w(name, self._src_rank(name))

    def wt_src(self, name: str) -> BMat:
        return self.wt(name, self._src_rank(name))

    def vec_src(self, name: str) -> torch.Tensor:
        return self.vec(name, self._src_rank(name))

    def grad(self, name: str) -> torch.Tensor:
        L = self.layout[name]
        return self.heap.tensor(self._grad_off + 4 * L.grad_off, L.spec.shape, torch.float32)

    def owner(self, name: str) -> int:
        return self.layout[name].owner

    def master(self, name: str, rank: Optional[int] = None) -> torch.Tensor:
        L = self.layout[name]
        return self.heap.tensor(L.master_off, L.spec.shape, torch.float32, L.owner if rank is None else rank)

    # ------------------------------------------------------------------ variable access by checkpoint names
    def _locate(self, name: str):
        """(offset, shape, dtype, owner rank) of any registry entry that has storage here."""
        spec = self.registry[name]
        if name in self.layout:
            L = self.layout[name]
            return L.master_off, spec.shape, torch.float32, L.owner
        if spec.kind == "step":
            return self._gs_off, (), torch.int32, self.gs_owner
        if spec.kind == "slot":
            L = self.layout[spec.primary]
            idx = self._opt_of[spec.primary].slots[spec.primary].index(name)
            return L.slot_offs[idx], spec.shape, torch.float32, L.owner
        if spec.kind == "nonslot":
            for k, op in enumerate(self.model.train_ops):
                if name in op.optimizer.nonslots:
                    j = op.optimizer.nonslots.index(name)
                    return self._beta_off[k] + 4 * j, (), torch.float32, self._beta_owner(k)
        raise KeyError(name)

    def read(self, name: str) -> torch.Tensor:
        """Authoritative value (from the owner's HBM) as a CPU tensor."""
        off, shape, dtype, owner = self._locate(name)
        n = 1
        for d in shape:
            n *= d
        return self.heap.tensor(off, (n,), dtype, owner).cpu().reshape(shape).clone()

    def read_all(self) -> Dict[str, torch.Tensor]:
        torch.cuda.synchronize()
        return {s.name: self.read(s.name) for s in self.registry}

    def assign(self, values: Dict[str, torch.Tensor], everywhere: bool = True) -> None:
        """Write values into the owners' state (the chief's init_op / checkpoint restore).
        With ``everywhere`` each rank writes its local replica too (lock-step replicas of
        global_step / beta powers in sync mode)."""
        for name, v in values.items():
            if name not in self.registry:
                continue
            off, shape, dtype, owner = self._locate(name)
            n = max(1, v.numel())
            src = v.to(dtype).reshape(n).to(self.device)
            targets = {owner}
            if everywhere:
                targets.add(self.rank)
            for r in targets:
                self.heap.tensor(off, (n,), dtype, r).copy_(src)
        torch.cuda.synchronize()

    def global_step(self) -> int:
        r = self.gs_owner if self.mode == "async" else self.rank
        return int(self.heap.tensor(self._gs_off, (1,), torch.int32, r).item())

    # ------------------------------------------------------------------ push
    def _desc(self, segs, nseg: int, sync: bool, step_delta: int, with_beta: bool,
              op_names: Optional[List[str]] = None) -> PushDesc:
        d = PushDesc()
        d.segs, d.nseg = segs.data_ptr(), nseg
        d.arena_ptrs, d.farena_ptrs = self._t_arena.data_ptr(), self._t_farena.data_ptr()
        d.pad_ptrs = self._t_pad.data_ptr()
        d.n_arenas = self.world
        d.local_ctrl = self.ctrl.data_ptr()
        d.step_id = 0
        big = self._grad_elems > (4 << 20)          # large shards: two CTAs per SM to keep NVLink loads in flight
        cap = 296 if big else 148
        d.blocks_per_seg = max(1, min(cap, cap // max(1, nseg))) if (sync and self.world > 1) else (cap // max(1, nseg) if big else 48)
        d.prof = self.prof.data_ptr()
        if sync:
            d.grad_ptrs, d.world, d.rank = self._t_grads.data_ptr(), self.world, self.rank
            d.inv_replicas = 1.0 / self.world
            d.signal_params = 1
            d.global_step = self.heap.ptr(self._gs_off)
        else:
            d.grad_ptrs, d.world, d.rank = self._t_grads_self.data_ptr(), 1, 0
            d.inv_replicas = 1.0
            d.signal_params = 0
            d.step_id = 1            # explicit: async pushes must not advance the barrier token
            d.async_sgd_atomic = 1
            d.global_step = self.heap.peer_ptr(self.gs_owner, self._gs_off)
        d.step_delta = step_delta
        nb = 0
        if with_beta:
            for k, op in enumerate(self.model.train_ops):
                if op_names is not None and op.name not in op_names:
                    continue
                if op.optimizer.kind == KIND_ADAM:
                    r = self.rank if sync else self._beta_owner(k)
                    d.beta[nb].beta_pow = self.heap.peer_ptr(r, self._beta_off[k])
                    d.beta[nb].beta1 = op.optimizer.hyper["beta1"]
                    d.beta[nb].beta2 = op.optimizer.hyper["beta2"]
                    nb += 1
        d.nbeta = nb
        return d

    def push(self, op_names: Optional[List[str]] = None) -> None:
        """Enqueue this step's gradient push on the current stream (graph-capturable).
        ``op_names`` restricts it to some train ops (GAN alternating updates)."""
        stream = _lib.current_stream_ptr()
        sync = self.mode == "sync"
        key = (sync, tuple(op_names) if op_names is not None else None)
        if key not in self._seg_cache:
            self._seg_cache[key] = self._build_segs(owned_only=sync, peer_state=not sync,
                                                    kind_override=None, op_names=op_names)
        segs, n = self._seg_cache[key]
        delta = sum(1 for op in self.model.train_ops if op.increments_global_step and
                    (op_names is None or op.name in op_names))
        d = self._desc(segs, n, sync, delta, True, op_names)
        _lib.check(_lib.lib().psb_push_apply(ctypes.byref(d), stream), "psb_push_apply")
        self._push_count += 1

    def refresh_copies(self) -> None:
        """Re-derive every bf16/fp32 compute copy from the owners' master weights (after the
        chief's init or a checkpoint restore).  Collective in sync geometry."""
        segs, n = self._segs_refresh
        d = self._desc(segs, n, True, 0, False)
        _lib.check(_lib.lib().psb_push_apply(ctypes.byref(d), _lib.current_stream_ptr()), "psb_push_apply(refresh)")
        torch.cuda.synchronize()

    def push_profile(self, reset: bool = True) -> Dict[str, float]:
        """Average microseconds per push spent (a) waiting for every rank's gradients, (b)
        reducing + applying + storing the copies, (c) waiting for every owner's parameters —
        measured inside the kernel with %globaltimer.  (a)+(c) is the exposed synchronisation
        cost of the step, (b) the exposed push/pull work."""
        torch.cuda.synchronize()
        p = self.prof.cpu().tolist()
        n = max(1, p[7])
        out = {"calls": p[7], "wait_grads_us": p[4] / n / 1e3, "apply_us": p[5] / n / 1e3,
               "wait_params_us": p[6] / n / 1e3}
        if reset:
            self.prof.zero_()
        return out

    def zero_grads(self) -> None:
        self.grads.zero_()

    def close(self) -> None:
        self.heap.close()
