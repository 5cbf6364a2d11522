"""Python face of the sm_100a kernels (the product compute path).

Every function here launches a hand-written kernel from ``csrc/`` on the current CUDA
stream through the C ABI in ``include/psb200.h``.  Operands are bf16 matrices with a
padded leading dimension (:class:`BMat`), accumulators/gradients/master weights are fp32.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from ._lib import Epilogue, check, current_stream_ptr

ACT_NONE, ACT_SIGMOID, ACT_RELU = 0, 1, 2


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class BMat:
    """bf16 row-major matrix ``[rows, cols]`` stored with ``ld = round_up(cols, 8)``.

    The pad columns are kept at zero so the matrix can be the K-padded operand of the
    next GEMM (TMA needs 16-byte row strides: K=100 -> 104, K=156 -> 160, N=1 -> 8).
    """

    __slots__ = ("t", "rows", "cols")

    def __init__(self, rows: int, cols: int, device="cuda", data: Optional[torch.Tensor] = None):
        self.rows, self.cols = int(rows), int(cols)
        self.t = torch.zeros(self.rows, round_up(self.cols, 8), dtype=torch.bfloat16, device=device)
        if data is not None:
            self.t[:, : self.cols].copy_(data)

    @property
    def ld(self) -> int:
        return self.t.shape[1]

    @property
    def ptr(self) -> int:
        return self.t.data_ptr()

    def float(self) -> torch.Tensor:
        return self.t[:, : self.cols].float()

    @staticmethod
    def from_tensor(x: torch.Tensor) -> "BMat":
        return BMat(x.shape[0], x.shape[1], x.device, x)


def _ptr(t) -> Optional[int]:
    if t is None:
        return None
    if isinstance(t, BMat):
        return t.ptr
    return t.data_ptr()


def gemm(a: BMat, b: BMat, *, alpha: float = 1.0, bias: Optional[torch.Tensor] = None,
         act: int = ACT_NONE, yprev: Optional[BMat] = None, dact: int = 0,
         out: Optional[BMat] = None, out_t: Optional[BMat] = None,
         out_f32: Optional[torch.Tensor] = None, colsum: Optional[torch.Tensor] = None,
         block_n: int = 0, k: Optional[int] = None) -> None:
    """``acc = alpha * a @ b.T`` (a: [M,K], b: [N,K]) with the fused epilogue of
    ``psb_gemm_bf16``.  ``k`` overrides the reduction length (defaults to ``a.cols``)."""
    M, N = a.rows, b.rows
    K = int(k) if k is not None else a.cols
    if k is None and a.cols != b.cols:
        raise ValueError("gemm: reduction lengths differ (%d vs %d)" % (a.cols, b.cols))
    e = Epilogue()
    e.alpha = alpha
    e.act = act
    e.dact = dact
    e.bias = _ptr(bias)
    e.yprev = _ptr(yprev)
    e.ld_yprev = yprev.ld if yprev is not None else 0
    e.out = _ptr(out)
    e.ld_out = out.ld if out is not None else 0
    e.out_t = _ptr(out_t)
    e.ld_out_t = out_t.ld if out_t is not None else 0
    e.out_f32 = _ptr(out_f32)
    e.ld_f32 = out_f32.stride(0) if out_f32 is not None and out_f32.dim() == 2 else (
        N if out_f32 is not None else 0)
    e.colsum = _ptr(colsum)
    if out is not None and (out.rows != M or out.cols != N):
        raise ValueError("gemm: out must be [%d,%d]" % (M, N))
    if out_t is not None and (out_t.rows != N or out_t.cols != M):
        raise ValueError("gemm: out_t must be [%d,%d]" % (N, M))
    check(_lib.lib().psb_gemm_bf16(a.ptr, a.ld, b.ptr, b.ld, M, N, K, e, block_n,
                                   current_stream_ptr()), "psb_gemm_bf16")
