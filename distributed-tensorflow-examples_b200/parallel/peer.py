"""Symmetric heap over CUDA IPC: the NVLink address space the fused push/pull kernels use.

Every rank (one process per GPU) cudaMalloc's one heap of the same size, exports its
64-byte IPC handle, and maps every peer's heap.  Sub-allocations are made in the same
order on every rank, so an object lives at the same offset everywhere and
``peer_ptr(r, off)`` is the NVLink-visible address of rank r's copy.

This replaces the gRPC tensor transport between ``/job:ps`` and ``/job:worker`` processes
that ``tf.train.Server`` provides in the reference (``encoder:96-98,157``).
"""
from __future__ import annotations

import ctypes
from ctypes import POINTER, c_int, c_size_t, c_ubyte, c_void_p
from typing import List, Optional, Sequence

import torch

from ..ops import _lib

_lib.register_signature("psb_heap_alloc", [c_size_t, POINTER(c_void_p)])
_lib.register_signature("psb_heap_free", [c_void_p])
_lib.register_signature("psb_ipc_export", [c_void_p, POINTER(c_ubyte)])
_lib.register_signature("psb_ipc_open", [POINTER(c_ubyte), POINTER(c_void_p)])
_lib.register_signature("psb_ipc_close", [c_void_p])

_DT_STR = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1", torch.int64: "<i8",
           torch.uint32: "<u4"}


class _Raw:
    """Minimal ``__cuda_array_interface__`` carrier so torch can view raw device memory."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 3, "strides": None}


def view_device_memory(ptr: int, shape: Sequence[int], dtype: torch.dtype, device) -> torch.Tensor:
    """Zero-copy torch tensor over ``ptr`` (bf16 goes through an int16 view)."""
    shape = tuple(int(s) for s in shape)
    if dtype == torch.bfloat16:
        t = torch.as_tensor(_Raw(ptr, shape, "<i2"), device=device)
        return t.view(torch.bfloat16)
    return torch.as_tensor(_Raw(ptr, shape, _DT_STR[dtype]), device=device)


class SymmetricHeap:
    def __init__(self, nbytes: int, rank: int = 0, world: int = 1, group=None, device=None):
        self.nbytes = int(nbytes)
        self.rank, self.world, self.group = rank, world, group
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())
        self._cursor = 0
        lib = _lib.lib()
        p = c_void_p()
        _lib.check(lib.psb_heap_alloc(self.nbytes, ctypes.byref(p)), "psb_heap_alloc")
        self.base = int(p.value)
        self.peer_bases: List[int] = [self.base] * world
        self._opened: List[int] = []
        if world > 1:
            self._exchange()

    def _exchange(self) -> None:
        import torch.distributed as dist
        lib = _lib.lib()
        buf = (c_ubyte * 64)()
        _lib.check(lib.psb_ipc_export(self.base, buf), "psb_ipc_export")
        mine = bytes(buf)
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, mine, group=self.group)
        for r, h in enumerate(handles):
            if r == self.rank:
                continue
            hb = (c_ubyte * 64).from_buffer_copy(h)
            p = c_void_p()
            _lib.check(lib.psb_ipc_open(hb, ctypes.byref(p)), "psb_ipc_open(rank %d)" % r)
            self.peer_bases[r] = int(p.value)
            self._opened.append(int(p.value))
        dist.barrier(group=self.group)

    # ---- sub-allocation (identical sequence on every rank) -----------------------
    def alloc(self, nbytes: int, align: int = 256) -> int:
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise MemoryError("symmetric heap exhausted: need %d more bytes" % (off + nbytes - self.nbytes))
        self._cursor = off + int(nbytes)
        return off

    def ptr(self, off: int) -> int:
        return self.base + off

    def peer_ptr(self, rank: int, off: int) -> int:
        return self.peer_bases[rank] + off

    def tensor(self, off: int, shape, dtype: torch.dtype, rank: Optional[int] = None) -> torch.Tensor:
        base = self.base if rank is None else self.peer_bases[rank]
        return view_device_memory(base + off, shape, dtype, self.device)

    def ptr_table(self, off: int) -> torch.Tensor:
        """int64 device tensor [world] with every rank's address of the object at ``off``."""
        return torch.tensor([b + off for b in self.peer_bases], dtype=torch.int64, device=self.device)

    def close(self) -> None:
        lib = _lib.lib()
        for p in self._opened:
            lib.psb_ipc_close(p)
        self._opened = []
        if self.base:
            lib.psb_heap_free(self.base)
            self.base = 0
