"""Client of the native parameter-server runtime (``csrc/ps_server.cpp``).

One :class:`PsClient` = one TCP connection to one task's server.  It is the RPC surface the
reference gets implicitly from TF's Send/Recv and queue ops: variable declare/assign/read
(``tf.Variable`` + ``init_op``, ``encoder:114-125,140``), optimizer apply on the ps
(``encoder:137``), ``global_step`` AssignAdd, shared-name FIFO queues (``encoder:54-63``) and
the sync-mode accumulators.  A client is not thread-safe: one per thread.
"""
from __future__ import annotations

import ctypes
import os
import socket
import struct
import time
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import PACKAGE_DIR

(OP_PING, OP_VAR_DECLARE, OP_VAR_ASSIGN, OP_VAR_READ, OP_VAR_UNINIT, OP_APPLY, OP_VAR_ADD_I32,
 OP_VAR_SCALE_F32, OP_Q_CREATE, OP_Q_ENQUEUE, OP_Q_DEQUEUE, OP_Q_SIZE, OP_ACC_DECLARE, OP_ACC_APPLY,
 OP_ACC_TAKE_APPLY, OP_ACC_SET_STEP, OP_SHUTDOWN, OP_VAR_LIST, OP_STATS) = range(1, 20)

ST_OK, ST_BAD_REQUEST, ST_UNINITIALIZED, ST_TIMEOUT, ST_NOT_FOUND, ST_SHAPE, ST_STOPPED = 0, -1, -2, -3, -4, -5, -6
_STATUS = {ST_BAD_REQUEST: "bad request", ST_UNINITIALIZED: "variable not initialized", ST_TIMEOUT: "timeout",
           ST_NOT_FOUND: "not found", ST_SHAPE: "shape/dtype mismatch", ST_STOPPED: "server stopping"}
MAGIC = 0x31425350

PS_LIB_PATH = os.path.join(PACKAGE_DIR, "lib", "libpsb200_ps.so")
_pslib = None


class PsError(RuntimeError):
    def __init__(self, status: int, what: str):
        super().__init__("%s: %s (%d)" % (what, _STATUS.get(status, "error"), status))
        self.status = status


class QueueTimeout(PsError):
    pass


def ps_lib():
    """The native server library (built by ``python -m psb200.build ps``)."""
    global _pslib
    if _pslib is None:
        if not os.path.exists(PS_LIB_PATH):
            from ..build import build_ps
            build_ps()
        lib = ctypes.CDLL(PS_LIB_PATH)
        lib.psb_ps_start.restype = ctypes.c_int
        lib.psb_ps_start.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        lib.psb_ps_stop.restype = ctypes.c_int
        lib.psb_ps_stop.argtypes = [ctypes.c_int]
        lib.psb_ps_port.restype = ctypes.c_int
        lib.psb_ps_port.argtypes = [ctypes.c_int]
        _pslib = lib
    return _pslib


def split_hostport(hp: str) -> Tuple[str, int]:
    host, _, port = hp.rpartition(":")
    return host or "127.0.0.1", int(port)


def _np_dtype(dtype: torch.dtype):
    return {torch.float32: np.float32, torch.int32: np.int32}[dtype]


def _buffer(t: torch.Tensor) -> memoryview:
    """The tensor's bytes as a flat buffer (no copy for contiguous CPU tensors)."""
    return memoryview(t.detach().contiguous().reshape(-1).numpy()).cast("B")


class PsClient:
    def __init__(self, hostport: str, connect_timeout: float = 60.0):
        self.hostport = hostport
        host, port = split_hostport(hostport)
        deadline = time.time() + connect_timeout
        last = None
        while True:
            try:
                self.sock = socket.create_connection((host, port), timeout=5.0)
                break
            except OSError as e:       # the peer task may not have started yet
                last = e
                if time.time() > deadline:
                    raise ConnectionError("cannot reach %s: %s" % (hostport, last))
                time.sleep(0.1)
        self.sock.settimeout(None)
        self.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)

    # ---- transport ---------------------------------------------------------------
    def _call(self, op: int, name: str = "", payload=b"", what: str = "", tail=None, out=None) -> bytes:
        """One request/response.  ``tail`` is an optional second payload part (a tensor's buffer,
        sent straight from its memory); ``out`` an optional writable buffer that receives a response
        body of exactly its size (a tensor read lands in place, no intermediate copies)."""
        nb = name.encode()
        tail_len = tail.nbytes if tail is not None else 0
        self.sock.sendall(struct.pack("<IIIQ", MAGIC, op, len(nb), len(payload) + tail_len) + nb + payload)
        if tail_len:
            self.sock.sendall(tail)
        status, n = struct.unpack("<iQ", self._recv(12))
        if out is not None and status == ST_OK and n == out.nbytes:
            self._recv_into(out)
            body = b""
        else:
            body = self._recv(n) if n else b""
            if out is not None and status == ST_OK:
                raise PsError(ST_SHAPE, "%s(%s)@%s: %d bytes, expected %d" % (what, name, self.hostport, n, out.nbytes))
        if status == ST_TIMEOUT:
            raise QueueTimeout(status, what or name)
        if status != ST_OK:
            raise PsError(status, "%s(%s)@%s" % (what or "op%d" % op, name, self.hostport))
        return body

    def _recv_into(self, view: memoryview) -> None:
        view = view.cast("B")
        got, n = 0, view.nbytes
        while got < n:
            r = self.sock.recv_into(view[got:], n - got)
            if r == 0:
                raise ConnectionError("connection to %s closed" % self.hostport)
            got += r

    def _recv(self, n: int) -> bytes:
        buf = bytearray(n)
        self._recv_into(memoryview(buf))
        return bytes(buf)

    def close(self) -> None:
        try:
            self.sock.close()
        except OSError:
            pass

    # ---- variables ---------------------------------------------------------------
    def ping(self) -> None:
        self._call(OP_PING, what="ping")

    def var_declare(self, name: str, shape: Sequence[int], dtype: torch.dtype) -> bool:
        """Create the variable if absent; returns whether it already holds a value."""
        code = 0 if dtype == torch.float32 else 1
        p = struct.pack("<II", code, len(shape)) + b"".join(struct.pack("<Q", int(d)) for d in shape)
        return struct.unpack("<I", self._call(OP_VAR_DECLARE, name, p, "declare"))[0] == 1

    def var_assign(self, name: str, value: torch.Tensor) -> None:
        self._call(OP_VAR_ASSIGN, name, b"", "assign", tail=_buffer(value))

    def var_read(self, name: str, shape: Sequence[int], dtype: torch.dtype) -> torch.Tensor:
        arr = np.empty(tuple(shape), dtype=_np_dtype(dtype))
        self._call(OP_VAR_READ, name, b"", "read", out=memoryview(arr.reshape(-1)))
        return torch.from_numpy(arr)

    def uninitialized_count(self) -> int:
        return struct.unpack("<I", self._call(OP_VAR_UNINIT, what="uninit"))[0]

    def var_list(self) -> List[str]:
        return [n for n in self._call(OP_VAR_LIST, what="list").decode().split("\n") if n]

    def var_add_i32(self, name: str, delta: int) -> int:
        return struct.unpack("<i", self._call(OP_VAR_ADD_I32, name, struct.pack("<i", delta), "add"))[0]

    def var_scale_f32(self, name: str, factor: float) -> None:
        self._call(OP_VAR_SCALE_F32, name, struct.pack("<f", factor), "scale")

    # ---- optimizer apply ---------------------------------------------------------
    @staticmethod
    def _apply_header(kind: int, hyper: Sequence[float], beta_powers: Sequence[float],
                      slot_names: Sequence[str]) -> bytes:
        lr, h1, h2, eps = hyper
        b = struct.pack("<IffffffI", kind, lr, h1, h2, eps, beta_powers[0], beta_powers[1], len(slot_names))
        for s in slot_names:
            sb = s.encode()
            b += struct.pack("<I", len(sb)) + sb
        return b

    def apply(self, name: str, grad: torch.Tensor, kind: int, hyper: Sequence[float],
              slot_names: Sequence[str], beta_powers: Sequence[float] = (0.0, 0.0)) -> None:
        """Async (Hogwild) dense update of ``name`` on its ps."""
        self._call(OP_APPLY, name, self._apply_header(kind, hyper, beta_powers, slot_names), "apply",
                   tail=_buffer(grad.to(torch.float32)))

    # ---- queues --------------------------------------------------------------------
    def queue_create(self, name: str, capacity: int) -> None:
        self._call(OP_Q_CREATE, name, struct.pack("<I", capacity), "queue_create")

    def queue_enqueue(self, name: str, value: int = 1) -> None:
        self._call(OP_Q_ENQUEUE, name, struct.pack("<i", value), "enqueue")

    def queue_enqueue_many(self, name: str, values: Sequence[int]) -> None:
        self._call(OP_Q_ENQUEUE, name, b"".join(struct.pack("<i", int(v)) for v in values), "enqueue_many")

    def queue_dequeue(self, name: str, timeout: float = 0.0) -> int:
        return struct.unpack("<i", self._call(OP_Q_DEQUEUE, name, struct.pack("<d", timeout), "dequeue"))[0]

    def queue_size(self, name: str) -> int:
        return struct.unpack("<I", self._call(OP_Q_SIZE, name, b"", "queue_size"))[0]

    # ---- sync-mode accumulators ------------------------------------------------------
    def acc_declare(self, name: str) -> None:
        self._call(OP_ACC_DECLARE, name, b"", "acc_declare")

    def acc_apply(self, name: str, grad: torch.Tensor, local_step: int) -> bool:
        """Returns False when the gradient was stale and dropped."""
        body = self._call(OP_ACC_APPLY, name, struct.pack("<q", local_step), "acc_apply",
                          tail=_buffer(grad.to(torch.float32)))
        return struct.unpack("<I", body)[0] == 1

    def acc_take_apply(self, name: str, required: int, kind: int, hyper: Sequence[float],
                       slot_names: Sequence[str], beta_powers: Sequence[float] = (0.0, 0.0)) -> int:
        """Block until ``required`` gradients arrived, apply their mean; returns how many were used."""
        p = struct.pack("<I", required) + self._apply_header(kind, hyper, beta_powers, slot_names)
        return struct.unpack("<I", self._call(OP_ACC_TAKE_APPLY, name, p, "acc_take_apply"))[0]

    def acc_set_step(self, name: str, step: int) -> None:
        self._call(OP_ACC_SET_STEP, name, struct.pack("<q", step), "acc_set_step")

    # ---- misc ------------------------------------------------------------------------
    def stats(self) -> dict:
        r, a, bi, bo, dropped = struct.unpack("<5Q", self._call(OP_STATS, what="stats"))
        return {"requests": r, "applies": a, "bytes_in": bi, "bytes_out": bo, "stale_dropped": dropped}

    def shutdown_server(self) -> None:
        try:
            self._call(OP_SHUTDOWN, what="shutdown")
        except (PsError, ConnectionError):
            pass
