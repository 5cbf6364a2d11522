"""NVLink parameter store: every GPU is a ps shard.

Replaces, on one 8xB200 box, the ps processes of the reference and the gRPC traffic to
them (``replica_device_setter`` placement ``encoder:111-113``; per-step variable reads and
``Apply*`` updates ``encoder:137,165``, ``gan:158-159,193``, ``lstm:97,128``):

* variable ``i`` (creation order) has its home on rank ``i mod world``; small variables live there
  whole, big ones are cut into ``world`` runs of rows owned by ranks ``home, home+1, ...``
  (:func:`row_blocks`) -- fp32 master weights, optimizer slots, ``global_step`` and Adam's beta
  powers live in the owner's HBM, checkpoint names and shapes are those of the whole variable;
* every rank keeps its gradients in one flat fp32 buffer inside a CUDA-IPC symmetric heap;
* **push** = one ``psb_push_apply`` kernel per step: every owner P2P-loads all ranks' slices of its
  segments, averages, applies SGD/RMSProp/Adam and stores the new bf16/fp32 compute copies
  (``sync``), or a worker applies its own gradient straight onto the owner's master copy through
  peer pointers / peer atomics (``async``, ``async_apply="worker"``);
* the ps role can be **decoupled** from the worker role (``replicas`` < world in sync mode,
  ``async_apply="owner"``): a native service thread per rank (``csrc/ps_service.cu``) keeps launching
  the push kernel in its non-blocking form -- first-R aggregation with stale gradients dropped, or
  one optimizer step per arrived gradient -- and the workers' step only signals and waits
  (``psb_push_commit``);
* **pull** = the compute copies are either written into every rank's arena by the push kernel itself
  (push-mode all-gather, ``bcast=True``) or stay in the home rank's arena and are fetched by the
  consuming GEMMs with TMA over NVLink, tile by tile as the owners publish per-row-block readiness
  flags (``bcast=False``);
* every cross-GPU wait is bounded (``timeout_s``) and can be aborted from the host
  (:meth:`NvlStore.abort`); a step that gave up raises :class:`PeerTimeoutError`.
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
                ("ld_wt", c_int32), ("f_off", c_int64), ("home", c_int32), ("bcast", c_int32),
                ("flags", c_uint32), ("nblocks", ctypes.c_uint16), ("block_x", ctypes.c_uint16),
                ("ready_slot", c_uint32), ("reserved_", c_uint32)]


class BetaUpdate(Structure):
    _fields_ = [("beta_pow", c_void_p), ("beta1", c_float), ("beta2", c_float)]


class PushDesc(Structure):
    _fields_ = [("segs", c_void_p), ("nseg", c_int), ("grad_ptrs", c_void_p), ("arena_ptrs", c_void_p),
                ("farena_ptrs", c_void_p), ("pad_ptrs", c_void_p), ("world", c_int), ("rank", c_int),
                ("n_arenas", c_int), ("inv_replicas", c_float), ("step_id", c_uint32),
                ("local_ctrl", c_void_p), ("beta", BetaUpdate * 2), ("nbeta", c_int),
                ("global_step", c_void_p), ("step_delta", c_int), ("async_sgd_atomic", c_int),
                ("signal_params", c_int), ("blocks_per_seg", c_int), ("prof", c_void_p),
                ("phases", c_int), ("replicas", c_int), ("timeout_ns", c_uint64),
                ("ready_ptrs", c_void_p), ("seg_done", c_void_p), ("big_tiles", c_int)]


SEG_ZERO_GRAD = 1
SEG_BIG_TILES = 2
CTRL_WORDS = 64                      # PSB_PUSH_CTRL_WORDS
CTRL_STEP, CTRL_STATUS, CTRL_ABORT, CTRL_APPLIED = 2, 3, 4, 7
CTRL_GLOBAL_STEP = 48                # global_step lives in a spare word of the control block: one D2H reads both
PH_SIGNAL_GRADS, PH_WAIT_GRADS, PH_TRY_GRADS, PH_APPLY, PH_SIGNAL_PARAMS, PH_WAIT_PARAMS, PH_EACH = 1, 2, 4, 8, 16, 32, 64
STATUS_TEXT = {1: "timed out waiting for a peer's gradients", 2: "timed out waiting for a peer's parameters",
               3: "aborted by the host"}


class ReadyInfo:
    """Readiness flags of one variable as a consuming GEMM needs them (``psb_epilogue_t.ready*``)."""
    __slots__ = ("flags", "step", "rows", "dim", "nblk", "first", "status", "timeout_ns")

    def __init__(self, flags, step, rows, dim, nblk, first, status, timeout_ns):
        self.flags, self.step, self.rows, self.dim, self.nblk = flags, step, rows, dim, nblk
        self.first, self.status, self.timeout_ns = first, status, timeout_ns


class PeerTimeoutError(RuntimeError):
    """A cross-GPU wait inside the push kernel ended on its deadline or on the abort word: a peer
    rank is dead or stuck.  The step was not applied on this rank."""


_lib.register_signature("psb_push_apply", [POINTER(PushDesc), c_void_p])
_lib.register_signature("psb_wait_params", [c_void_p, c_int, c_uint32, c_void_p])
_lib.register_signature("psb_push_commit", [c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_uint64, c_void_p])
_lib.register_signature("psb_wait_params_token", [c_void_p, c_int, c_void_p, c_uint64, c_void_p])
_lib.register_signature("psb_ps_service_start", [POINTER(PushDesc), c_int, POINTER(c_void_p)])
_lib.register_signature("psb_ps_service_stop", [c_void_p, POINTER(c_uint64)])


class _VarLayout:
    __slots__ = ("spec", "rows", "cols", "grad_off", "master_off", "slot_offs", "w_off", "ld_w",
                 "wt_off", "ld_wt", "f_off", "owner", "blocks", "ready_base")


def row_blocks(rows: int, cols: int, home: int, world: int, partition: bool = True, min_elems: int = 8192):
    """Row blocks ``(r0, r1, owner)`` of one variable.  A variable stays whole on its home rank
    (``replica_device_setter``'s round-robin, ``encoder:111-113``) unless it is big enough for the
    reduction to matter; then it is cut into ``world`` runs of rows (multiples of four, so both bf16
    copies keep their 8-byte stores) owned by ranks ``home, home+1, ...`` -- a partitioned variable in
    TF terms, same checkpoint name and shape."""
    if not partition or world <= 1 or rows * cols < min_elems or rows < 4 * world:
        return [(0, rows, home)]
    per = (rows + world - 1) // world
    per = (per + 3) // 4 * 4
    out, r0, b = [], 0, 0
    while r0 < rows:
        r1 = min(rows, r0 + per)
        out.append((r0, r1, (home + b) % world))
        r0, b = r1, b + 1
    return out


def share_blocks(rounds: List[int], budget: int = 296) -> List[int]:
    """Blocks per segment for one push launch.  ``rounds[j]`` = loop iterations segment j needs if one
    block did it alone; the result uses the smallest number of iterations per block such that the whole
    launch is at most ``budget`` blocks (one wave of two blocks per SM): small models then finish in a
    single iteration per block, and no block is idle whatever the mix of segment sizes."""
    R = 1
    while sum((r + R - 1) // R for r in rounds) > budget:
        R += 1
    return [max(1, (r + R - 1) // R) for r in rounds]


class NvlStore:
    """Parameter shards + gradient buffers of one rank, and the push launcher."""

    PAD_WORDS = 64
    PARTITION_MIN_ELEMS = 8192
    # row blocks at least this big are pushed in 128x128 tiles (256-byte runs); PSB200_BIG_TILES_MIN overrides (tests)
    BIG_TILES_MIN_ELEMS = int(__import__("os").environ.get("PSB200_BIG_TILES_MIN", 1 << 20))

    def __init__(self, model: Model, registry: VariableRegistry, rank: int = 0, world: int = 1,
                 group=None, mode: str = "sync", bcast: bool = True, device=None, partition: bool = True,
                 timeout_s: float = 30.0, keep_grads: bool = False, replicas: int = 0,
                 async_apply: str = "worker"):
        if mode not in ("sync", "async"):
            raise ValueError("mode must be 'sync' or 'async'")
        if async_apply not in ("worker", "owner"):
            raise ValueError("async_apply must be 'worker' (remote read-modify-write) or 'owner' (ps service)")
        if replicas < 0 or replicas > world:
            raise ValueError("replicas_to_aggregate must be in 1..%d" % world)
        self.model, self.registry = model, registry
        self.rank, self.world, self.group, self.mode, self.bcast = rank, world, group, mode, bcast
        self.partition = partition and world > 1
        self.timeout_ns = int(timeout_s * 1e9)
        self.keep_grads = keep_grads          # tests/debugging: leave the gradients readable after the push
        # The ps role runs as its own native service thread (csrc/ps_service.cu) when the aggregate does
        # not wait for everybody (sync, R < N: backup workers, stale gradients dropped) or when async
        # updates are applied by the owner; otherwise worker and ps role share the fused push kernel.
        self.replicas = replicas if 0 < replicas < world else 0
        self.async_apply = async_apply if (mode == "async" and world > 1) else "worker"
        self.decoupled = world > 1 and ((mode == "sync" and self.replicas > 0) or
                                        (mode == "async" and self.async_apply == "owner"))
        self._service = None
        # bcast=False (pull mode): the copies stay with the variable's home rank, consumers fetch them
        # over NVLink inside their GEMMs, gated row block by row block on readiness flags; the push then
        # does not wait for the other owners at its end
        self.pull_flags = (not bcast) and world > 1 and mode == "sync" and not self.decoupled
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.layout: Dict[str, _VarLayout] = {}
        self._opt_of: Dict[str, Optimizer] = {}
        for op in model.train_ops:
            for n in op.var_list:
                self._opt_of[n] = op.optimizer
        self._zero_after_push = set()
        self._plan_heap()
        self.heap = SymmetricHeap(self._heap_bytes, rank, world, group, self.device)
        assert self.heap.alloc(self._heap_bytes, 256) == 0
        h = self.heap
        self.pad = h.tensor(self._pad_off, (self.PAD_WORDS,), torch.int32)
        self.ctrl = h.tensor(self._ctrl_off, (CTRL_WORDS,), torch.int32)
        self.wctrl = h.tensor(self._wctrl_off, (CTRL_WORDS,), torch.int32)
        self.grads = h.tensor(self._grad_off, (self._grad_elems,), torch.float32)
        self.global_step_t = h.tensor(self._gs_off, (1,), torch.int32)
        self.gs_owner = model.global_step.ps_task % world
        self.prof = torch.zeros(8, dtype=torch.int64, device=self.device)
        self._tables()
        self._seg_cache = {}
        self._push_count = 0
        self._abort_src = torch.ones(1, dtype=torch.int32).pin_memory()
        self._abort_stream = None

    # ------------------------------------------------------------------ layout
    def _row_blocks(self, rows: int, cols: int, home: int):
        return row_blocks(rows, cols, home, self.world, self.partition, self.PARTITION_MIN_ELEMS)

    def _plan_heap(self) -> None:
        cur = 0

        def take(nbytes: int, align: int = 256) -> int:
            nonlocal cur
            off = (cur + align - 1) // align * align
            cur = off + nbytes
            return off

        self._pad_off = take(self.PAD_WORDS * 4)
        self._ctrl_off = take(CTRL_WORDS * 4)
        self._wctrl_off = take(CTRL_WORDS * 4)          # worker-side words of the decoupled protocol
        # readiness words: one per (variable, row block) -- written by the block's owner on every rank,
        # watched by consumers that pull the copies (bcast=False); and the owners' per-block counters
        nv = sum(len(op.var_list) for op in self.model.train_ops)
        self._ready_words = max(1, nv * max(1, self.world))
        self._ready_off = take(self._ready_words * 4)
        self._segdone_off = take(self._ready_words * 4)
        self._gs_off = self._ctrl_off + 4 * CTRL_GLOBAL_STEP
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
            L.blocks = self._row_blocks(L.rows, L.cols, L.owner) if len(s.shape) == 2 else [(0, 1, L.owner)]
            L.grad_off = g
            g += round_up(s.numel, 4)
            L.ready_base = len(self.layout) * max(1, self.world)      # first readiness word of this variable
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
        self._t_ready = h.ptr_table(self._ready_off)
        # single-entry tables for the async launch (world == 1 from the kernel's viewpoint)
        self._t_grads_self = torch.tensor([h.ptr(self._grad_off)], dtype=torch.int64, device=self.device)

    def zero_after_push(self, names) -> None:
        """Variables whose gradient producers ACCUMULATE (bias column sums, atomically added split-K
        partials): the push kernel clears their gradients in every rank's buffer right after reading
        them, which replaces a separate clear of the gradient buffer in every step."""
        self._zero_after_push.update(names)
        self._seg_cache = {}

    @property
    def owner_zeroes_grads(self) -> bool:
        """True: the push kernel clears the accumulated gradients it has consumed (PSB_SEG_ZERO_GRAD);
        False: the plans clear the gradient buffer themselves at the start of a step."""
        return not self.keep_grads and not self.decoupled

    @property
    def zero_in_before_grads(self) -> bool:
        """Async with owner-side apply: the gradient buffer may only be cleared once every owner has
        acknowledged the previous gradient, so :meth:`before_grads` does it."""
        return (self.decoupled and self.mode == "async") or (self.pull_flags and not self.owner_zeroes_grads)

    def _build_segs(self, owned_only: bool, peer_state: bool, kind_override: Optional[int],
                    op_names: Optional[List[str]] = None):
        """Segment table (device) + grid.x.  One segment per owned row block."""
        h = self.heap
        segs, rounds = [], []
        wanted = None
        if op_names is not None:
            wanted = {n for op in self.model.train_ops if op.name in op_names for n in op.var_list}
        for name, L in self.layout.items():
            refresh_only = False
            if wanted is not None and name not in wanted:
                if not (self.pull_flags and kind_override is None):
                    continue
                # pull mode: consumers wait for EVERY variable's flag to reach the new step token, so the
                # variables this push does not update are re-published unchanged
                refresh_only = True
            opt = self._opt_of[name]
            hv = opt.hyper_vector()
            for bi, (r0, r1, owner) in enumerate(L.blocks):
                if owned_only and owner != self.rank:
                    continue
                two_d = len(L.spec.shape) == 2
                e0 = r0 * L.cols if two_d else 0
                numel = (r1 - r0) * L.cols if two_d else L.spec.numel
                s = PushSeg()
                s.grad_offset, s.numel, s.cols = L.grad_off + e0, numel, L.cols
                s.opt_kind = -1 if refresh_only else (opt.kind if kind_override is None else kind_override)
                s.lr, s.h1, s.h2, s.eps = hv
                st_rank = owner if peer_state else self.rank
                s.master = h.peer_ptr(st_rank, L.master_off + 4 * e0)
                s.slot0 = h.peer_ptr(st_rank, L.slot_offs[0] + 4 * e0) if len(L.slot_offs) > 0 else None
                s.slot1 = h.peer_ptr(st_rank, L.slot_offs[1] + 4 * e0) if len(L.slot_offs) > 1 else None
                if opt.kind == KIND_ADAM:
                    k = [i for i, op in enumerate(self.model.train_ops) if op.optimizer is opt][0]
                    brank = self._beta_owner(k) if peer_state else self.rank
                    s.beta_pow = h.peer_ptr(brank, self._beta_off[k])
                s.w_off = L.w_off + r0 * L.ld_w if L.w_off >= 0 else -1
                s.wt_off = L.wt_off + r0 if L.wt_off >= 0 else -1
                s.ld_w, s.ld_wt, s.f_off = L.ld_w, L.ld_wt, L.f_off
                s.home = L.owner            # bcast=0: every row block lands in the arena of the variable's home rank
                s.bcast = 1 if self.bcast else 0
                s.flags = SEG_ZERO_GRAD if (name in self._zero_after_push and kind_override is None
                                            and self.owner_zeroes_grads and not refresh_only) else 0
                s.ready_slot = (L.ready_base + bi + 1) if self.pull_flags else 0
                # rounds of work: one block iteration covers TU 32x32 tiles, or 256 threads x U groups of
                # four (TU, U as in csrc/push.cu for this world's template bucket)
                kw = 1 if peer_state else self.world        # legacy async launches run with world == 1
                maxw = 1 << max(0, (kw - 1).bit_length())
                if two_d and numel >= self.BIG_TILES_MIN_ELEMS and (r1 - r0) >= 128 and L.cols >= 128 and not peer_state:
                    s.flags |= SEG_BIG_TILES
                    rounds.append((((r1 - r0) + 127) // 128) * ((L.cols + 127) // 128))
                elif two_d:
                    units = (((r1 - r0) + 31) // 32) * ((L.cols + 31) // 32)
                    tu = 1 if maxw >= 8 else 2
                    rounds.append(max(1, (units + tu - 1) // tu))
                else:
                    per_iter = 256 * 4 * (1 if maxw >= 8 else (2 if maxw >= 4 else 4))
                    rounds.append(max(1, (numel + per_iter - 1) // per_iter))
                segs.append(s)
        # `rounds[j]` = block iterations segment j needs.  Smallest number of iterations per block R such
        # that the whole launch fits in one wave of 2 blocks per SM: small models then finish in a single
        # iteration per block.  The table lists every segment once per block that works on it, and the
        # launch is one block per entry: no idle blocks, whatever the mix of segment sizes.
        items = []
        for s, nb in zip(segs, share_blocks(rounds)):
            for x in range(nb):
                it = PushSeg.from_buffer_copy(bytes(s))
                it.nblocks, it.block_x = nb, x
                items.append(it)
        arr = (PushSeg * max(len(items), 1))(*items)
        raw = bytes(arr) if items else b"\0" * ctypes.sizeof(PushSeg)
        dev = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.device)
        dev._psb_big_tiles = any(it.flags & SEG_BIG_TILES for it in items)
        return dev, len(items), 1

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
        m.ready = self._ready_info(L, dim=1)
        return m

    def wt(self, name: str, rank: Optional[int] = None) -> BMat:
        """Transposed bf16 copy ``[cols, rows]``."""
        L = self.layout[name]
        m = BMat.__new__(BMat)
        m.rows, m.cols = L.cols, L.rows
        m.remote = rank is not None and rank != self.rank
        m.t = self.heap.tensor(self._arena_off + 2 * L.wt_off, (L.cols, L.ld_wt), torch.bfloat16, rank)
        m.ready = self._ready_info(L, dim=0)
        return m

    def _ready_info(self, L: _VarLayout, dim: int) -> Optional[ReadyInfo]:
        """Flags of ``L``'s row blocks for a consuming GEMM: ``dim`` 0 when the blocks cut the GEMM's K
        (the transposed copy is B), 1 when they cut its N (the natural copy is B)."""
        if not self.pull_flags:
            return None
        nblk = len(L.blocks)
        rows = L.blocks[0][1] - L.blocks[0][0]
        first = 0
        for i, (_, _, owner) in enumerate(L.blocks):
            if owner == self.rank:
                first = i
        return ReadyInfo(self.heap.ptr(self._ready_off + 4 * L.ready_base), self.heap.ptr(self._ctrl_off + 4 * CTRL_STEP),
                         max(1, rows), dim, nblk, first, self.heap.ptr(self._ctrl_off + 4 * CTRL_STATUS), self.timeout_ns)

    def ready_for(self, name: str) -> Optional[ReadyInfo]:
        """Readiness flag of a 1-D variable (bias): the consuming GEMM's epilogue waits for it."""
        return self._ready_info(self.layout[name], dim=0)

    def wait_all_params(self) -> None:
        """Pull mode: block the stream until every owner has published this step's parameters -- for
        consumers that do not watch readiness flags themselves (the LSTM kernels' weight prologue), and
        once before the first gradient of a step is written (the owners have then also finished reading
        the previous gradients)."""
        if self.pull_flags:
            _lib.check(_lib.lib().psb_wait_params_token(self.pad.data_ptr(), self.world, self.ctrl.data_ptr(),
                                                        self.timeout_ns, _lib.current_stream_ptr()), "psb_wait_params_token")

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
        """Home rank of a variable (of its first row block when it is partitioned)."""
        return self.layout[name].owner

    def master(self, name: str, rank: Optional[int] = None) -> torch.Tensor:
        L = self.layout[name]
        return self.heap.tensor(L.master_off, L.spec.shape, torch.float32, L.owner if rank is None else rank)

    # ------------------------------------------------------------------ variable access by checkpoint names
    def _locate(self, name: str):
        """(offset, shape, dtype, [(elem0, elem1, owner rank)]) of any registry entry that has storage here."""
        spec = self.registry[name]
        if name in self.layout:
            L = self.layout[name]
            return L.master_off, spec.shape, torch.float32, self._runs(L)
        if spec.kind == "step":
            return self._gs_off, (), torch.int32, [(0, 1, self.gs_owner)]
        if spec.kind == "slot":
            L = self.layout[spec.primary]
            idx = self._opt_of[spec.primary].slots[spec.primary].index(name)
            return L.slot_offs[idx], spec.shape, torch.float32, self._runs(L)
        if spec.kind == "nonslot":
            for k, op in enumerate(self.model.train_ops):
                if name in op.optimizer.nonslots:
                    j = op.optimizer.nonslots.index(name)
                    return self._beta_off[k] + 4 * j, (), torch.float32, [(0, 1, self._beta_owner(k))]
        raise KeyError(name)

    @staticmethod
    def _runs(L: _VarLayout):
        if len(L.spec.shape) != 2:
            return [(0, max(1, L.spec.numel), L.owner)]
        return [(r0 * L.cols, r1 * L.cols, owner) for (r0, r1, owner) in L.blocks]

    def read(self, name: str) -> torch.Tensor:
        """Authoritative value (every row block from its owner's HBM) as a CPU tensor."""
        off, shape, dtype, runs = self._locate(name)
        parts = [self.heap.tensor(off + 4 * e0, (e1 - e0,), dtype, owner).cpu() for (e0, e1, owner) in runs]
        return torch.cat(parts).reshape(shape).clone()

    def read_all(self) -> Dict[str, torch.Tensor]:
        self.flush()
        torch.cuda.synchronize()
        return {s.name: self.read(s.name) for s in self.registry}

    def snapshot_async(self, stream: "torch.cuda.Stream") -> Dict[str, torch.Tensor]:
        """Every variable copied owner-HBM -> pinned host memory on ``stream`` without stopping the
        training stream (the checkpoint timer thread's read: like the reference's Saver it may see a
        mixture of two consecutive steps, ``encoder:138,151``).  Synchronise ``stream`` before use."""
        out = {}
        if not hasattr(self, "_snap_bufs"):
            self._snap_bufs = {}
        with torch.cuda.stream(stream):
            for s in self.registry:
                off, shape, dtype, runs = self._locate(s.name)
                n = runs[-1][1]
                host = self._snap_bufs.get(s.name)
                if host is None:
                    host = self._snap_bufs[s.name] = torch.empty(n, dtype=dtype).pin_memory()
                for (e0, e1, owner) in runs:
                    host[e0:e1].copy_(self.heap.tensor(off + 4 * e0, (e1 - e0,), dtype, owner), non_blocking=True)
                out[s.name] = host.reshape(shape)
        return out

    def assign(self, values: Dict[str, torch.Tensor], everywhere: bool = True) -> None:
        """Write values into the owners' state (the chief's init_op / checkpoint restore).
        With ``everywhere`` each rank writes its local replica too (lock-step replicas of
        global_step / beta powers in sync mode)."""
        for name, v in values.items():
            if name not in self.registry:
                continue
            off, shape, dtype, runs = self._locate(name)
            n = max(1, v.numel())
            src = v.to(dtype).reshape(n).to(self.device)
            for (e0, e1, owner) in runs:
                targets = {owner}
                if everywhere:
                    targets.add(self.rank)
                for r in targets:
                    self.heap.tensor(off + 4 * e0, (e1 - e0,), dtype, r).copy_(src[e0:e1])
        torch.cuda.synchronize()

    def global_step(self) -> int:
        self.flush()
        r = self.gs_owner if self.mode == "async" else self.rank
        return int(self.heap.tensor(self._gs_off, (1,), torch.int32, r).item())

    # ------------------------------------------------------------------ failure containment
    def status(self) -> int:
        return max(int(self.ctrl[CTRL_STATUS].item()), int(self.wctrl[CTRL_STATUS].item()))

    def check_status(self, status: Optional[int] = None) -> None:
        st = self.status() if status is None else status
        if st != 0:
            raise PeerTimeoutError("rank %d: push kernel %s (status %d)" % (self.rank, STATUS_TEXT.get(st, "failed"), st))

    def abort(self) -> None:
        """Make every cross-GPU wait of a running or future push kernel on this rank give up (a
        supervisor calls this when a peer is known to be gone).  Safe while a kernel is spinning:
        the write goes through a separate stream."""
        if self._abort_stream is None:
            self._abort_stream = torch.cuda.Stream(device=self.device)
        with torch.cuda.stream(self._abort_stream):
            self.ctrl[CTRL_ABORT:CTRL_ABORT + 1].copy_(self._abort_src, non_blocking=True)
            self.wctrl[CTRL_ABORT:CTRL_ABORT + 1].copy_(self._abort_src, non_blocking=True)
        self._abort_stream.synchronize()

    # ------------------------------------------------------------------ push
    def _desc(self, segs, nseg: int, gx: int, sync: bool, step_delta: int, with_beta: bool,
              op_names: Optional[List[str]] = None) -> PushDesc:
        d = PushDesc()
        d.segs, d.nseg = segs.data_ptr(), nseg
        d.big_tiles = 1 if getattr(segs, "_psb_big_tiles", False) else 0
        d.arena_ptrs, d.farena_ptrs = self._t_arena.data_ptr(), self._t_farena.data_ptr()
        d.pad_ptrs = self._t_pad.data_ptr()
        d.n_arenas = self.world
        d.local_ctrl = self.ctrl.data_ptr()
        d.step_id = 0
        d.blocks_per_seg = gx
        d.prof = self.prof.data_ptr()
        d.timeout_ns = self.timeout_ns
        if self.pull_flags:
            d.ready_ptrs, d.seg_done = self._t_ready.data_ptr(), self.heap.ptr(self._segdone_off)
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

    # ---- decoupled ps role -------------------------------------------------------------------
    def _commit(self, wait: int) -> None:
        _lib.check(_lib.lib().psb_push_commit(self._t_pad.data_ptr(), self.world, self.rank, self.wctrl.data_ptr(),
                                              1 if self.mode == "async" else 0, wait, self.timeout_ns,
                                              _lib.current_stream_ptr()), "psb_push_commit")

    def before_grads(self) -> None:
        """Called by the plans right before the first kernel that writes a gradient.  With owner-side
        async apply the gradient buffer is still owned by the ps shards until each of them has
        acknowledged the previous gradient: wait for that (normally long done: it overlapped the
        forward pass), then clear what the backward kernels accumulate into."""
        if self.decoupled and self.mode == "async":
            self._commit(1)
        self.wait_all_params()
        if self.zero_in_before_grads and self._zero_after_push:
            self.grads.zero_()

    def start_service(self, poll_us: int = 0) -> None:
        """Start this rank's ps service thread (collective: call on every rank after the compute
        copies are in place and after any CUDA-graph capture)."""
        if not self.decoupled or self._service is not None:
            return
        sync = self.mode == "sync"
        if "service" not in self._seg_cache:
            self._seg_cache["service"] = self._build_segs(owned_only=True, peer_state=False, kind_override=None)
        segs, n, gx = self._seg_cache["service"]
        delta = sum(1 for op in self.model.train_ops if op.increments_global_step)
        d = self._desc(segs, n, gx, True, delta, True)
        d.phases = PH_TRY_GRADS | PH_APPLY | (PH_SIGNAL_PARAMS if sync else PH_EACH)
        d.replicas = self.replicas if sync else 1
        d.inv_replicas = 1.0
        d.signal_params = 0
        d.prof = None
        if not sync and self.rank != self.gs_owner:
            d.global_step = None            # one increment per gradient: only the step's owner counts
        torch.cuda.synchronize()
        if sync:
            # the workers' first token is the owners' next aggregate (the parameter refresh consumed some)
            self.wctrl[CTRL_STEP] = self.ctrl[CTRL_STEP] + 1
        torch.cuda.synchronize()
        if self.world > 1 and self.group is not None:
            import torch.distributed as dist
            dist.barrier(group=self.group)
        h = c_void_p()
        _lib.check(_lib.lib().psb_ps_service_start(ctypes.byref(d), int(poll_us), ctypes.byref(h)), "psb_ps_service_start")
        self._service = h
        self._service_desc = d               # keeps the ctypes struct (and what it points to) alive
        import atexit
        atexit.register(self.stop_service)   # never leave the native thread running into interpreter teardown

    def stop_service(self) -> int:
        if self._service is None:
            return 0
        n = c_uint64(0)
        h, self._service = self._service, None
        _lib.check(_lib.lib().psb_ps_service_stop(h, ctypes.byref(n)), "psb_ps_service_stop")
        return int(n.value)

    def flush(self) -> None:
        """Async with owner-side apply: block until every owner has consumed this rank's last gradient."""
        if self.decoupled and self.mode == "async" and self._service is not None:
            self._commit(1)
            torch.cuda.synchronize()
            self.check_status()

    def push(self, op_names: Optional[List[str]] = None) -> None:
        """Enqueue this step's gradient push on the current stream (graph-capturable).
        ``op_names`` restricts it to some train ops (GAN alternating updates)."""
        if self.decoupled:
            if op_names is not None:
                raise ValueError("per-op pushes (GAN alternating updates) need the fused push: use "
                                 "replicas_to_aggregate == workers / async_apply='worker'")
            self._commit(0)
            self._push_count += 1
            return
        stream = _lib.current_stream_ptr()
        sync = self.mode == "sync"
        key = (sync, tuple(op_names) if op_names is not None else None)
        if key not in self._seg_cache:
            self._seg_cache[key] = self._build_segs(owned_only=sync, peer_state=not sync,
                                                    kind_override=None, op_names=op_names)
        segs, n, gx = self._seg_cache[key]
        delta = sum(1 for op in self.model.train_ops if op.increments_global_step and
                    (op_names is None or op.name in op_names))
        d = self._desc(segs, n, gx, sync, delta, True, op_names)
        if self.pull_flags:
            d.phases = PH_SIGNAL_GRADS | PH_WAIT_GRADS | PH_APPLY | PH_SIGNAL_PARAMS     # no wait for the other owners
        _lib.check(_lib.lib().psb_push_apply(ctypes.byref(d), stream), "psb_push_apply")
        self._push_count += 1

    def refresh_copies(self) -> None:
        """Re-derive every bf16/fp32 compute copy from the owners' master weights (after the
        chief's init or a checkpoint restore).  Collective in sync geometry."""
        key = "refresh"
        if key not in self._seg_cache:
            self._seg_cache[key] = self._build_segs(owned_only=True, peer_state=False, kind_override=-1)
        segs, n, gx = self._seg_cache[key]
        d = self._desc(segs, n, gx, True, 0, False)
        _lib.check(_lib.lib().psb_push_apply(ctypes.byref(d), _lib.current_stream_ptr()), "psb_push_apply(refresh)")
        torch.cuda.synchronize()
        self.check_status()

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
        self.stop_service()
        self.heap.close()
