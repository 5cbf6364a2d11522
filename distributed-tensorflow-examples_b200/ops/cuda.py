"""Python face of the sm_100a kernels (the product compute path).

Every function here launches a hand-written kernel from ``csrc/`` on the current CUDA
stream through the C ABI in ``include/psb200.h``.  Operands are bf16 matrices with a
padded leading dimension (:class:`BMat`), accumulators/gradients/master weights are fp32.
"""
from __future__ import annotations

import ctypes
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

    __slots__ = ("t", "rows", "cols", "remote", "ready")

    def __init__(self, rows: int, cols: int, device="cuda", data: Optional[torch.Tensor] = None):
        self.rows, self.cols = int(rows), int(cols)
        self.remote = False          # True: the storage is another GPU's memory (peer mapping)
        self.ready = None            # readiness flags of the variable behind this copy (store/nvl.py: ReadyInfo)
        self.t = torch.zeros(self.rows, round_up(self.cols, 8), dtype=torch.bfloat16, device=device)
        if data is not None:
            self.t[:, : self.cols].copy_(data)

    @property
    def ld(self) -> int:
        return self.t.stride(0)

    def rows_view(self, r0: int, r1: int) -> "BMat":
        """Rows ``r0:r1`` as a matrix sharing this storage."""
        m = BMat.__new__(BMat)
        m.rows, m.cols, m.t, m.remote, m.ready = r1 - r0, self.cols, self.t[r0:r1], self.remote, None
        return m

    def cols_view(self, c0: int, c1: int) -> "BMat":
        """Columns ``c0:c1`` (c0 a multiple of 8) as a matrix sharing this storage."""
        if c0 % 8:
            raise ValueError("column views must start on a 16-byte boundary")
        m = BMat.__new__(BMat)
        m.rows, m.cols, m.t, m.remote, m.ready = self.rows, c1 - c0, self.t[:, c0:], self.remote, None
        return m

    @property
    def ptr(self) -> int:
        return self.t.data_ptr()

    def float(self) -> torch.Tensor:
        return self.t[:, : self.cols].float()

    @staticmethod
    def wrap(t: torch.Tensor, rows: int, cols: int, remote: bool = False) -> "BMat":
        m = BMat.__new__(BMat)
        m.rows, m.cols, m.t, m.remote, m.ready = rows, cols, t, remote, None
        return m

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
         block_n: int = 0, k: Optional[int] = None, split_k: int = 1, split_stride: int = 0,
         b_static: bool = False, bias_ready=None, pull_mode: int = 0) -> None:
    """``acc = alpha * a @ b.T`` (a: [M,K], b: [N,K]) with the fused epilogue of
    ``psb_gemm_bf16``.  ``k`` overrides the reduction length (defaults to ``a.cols``).
    ``b_static=True`` promises that ``b`` (a weight copy) was not written by the kernel launched
    just before this one, which lets the kernel prefetch it under that kernel's tail."""
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
    e.split_k = int(split_k)
    e.split_stride = int(split_stride)
    e.b_static = 1 if b_static else 0
    e.pull_mode = int(pull_mode)
    rd = getattr(b, "ready", None)
    if rd is not None:
        # the copy is refreshed row block by row block by the owners' push kernels: gate the tiles on
        # their flags (parameter pull fused into this GEMM, overlapped tile by tile)
        e.ready, e.ready_step, e.ready_rows, e.ready_dim = rd.flags, rd.step, rd.rows, rd.dim
        e.ready_nblk, e.ready_first, e.ready_status, e.ready_timeout_ns = rd.nblk, rd.first, rd.status, rd.timeout_ns
    if bias_ready is not None:
        e.bias_ready, e.ready_step = bias_ready.flags, bias_ready.step
        e.ready_status, e.ready_timeout_ns = bias_ready.status, bias_ready.timeout_ns
    if out is not None and (out.rows != M or out.cols != N):
        raise ValueError("gemm: out must be [%d,%d]" % (M, N))
    if out_t is not None and (out_t.rows != N or out_t.cols != M):
        raise ValueError("gemm: out_t must be [%d,%d]" % (N, M))
    if getattr(b, "remote", False):
        # the weight copy lives in its owner's HBM: pull it inside the GEMM (no TMA on peer memory)
        check(_lib.lib().psb_gemm_bf16_pull(a.ptr, a.ld, b.ptr, b.ld, M, N, K, e, block_n,
                                            current_stream_ptr()), "psb_gemm_bf16_pull")
    else:
        check(_lib.lib().psb_gemm_bf16(a.ptr, a.ld, b.ptr, b.ld, M, N, K, e, block_n,
                                       current_stream_ptr()), "psb_gemm_bf16")


# --------------------------------------------------------------------------- LSTM
from ctypes import c_float, c_int, c_void_p  # noqa: E402

_lib.register_signature("psb_lstm_fwd", [c_void_p, c_void_p, c_int, c_void_p, c_float, c_int, c_int,
                                         c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_void_p,
                                         c_void_p, c_void_p, c_int, c_void_p])
_lib.register_signature("psb_lstm_bwd", [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int,
                                         c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p])


class LstmHead(ctypes.Structure):
    """Mirror of ``psb_lstm_head_t``."""
    _fields_ = [("h_last", c_void_p), ("ld_h", c_int), ("w_out", c_void_p), ("ld_wo", c_int),
                ("b_out", c_void_p), ("labels", c_void_p), ("ld_l", c_int), ("C", c_int),
                ("logits", c_void_p), ("ld_logits", c_int), ("loss", c_void_p), ("correct", c_void_p),
                ("d_w_out", c_void_p), ("ld_dwo", c_int), ("d_b_out", c_void_p), ("scratch", c_void_p),
                ("accumulate", c_int), ("zero_ptr", c_void_p), ("zero_bytes", ctypes.c_size_t)]


_lib.register_signature("psb_lstm_bwd_head", [c_void_p, c_int, ctypes.POINTER(LstmHead), c_void_p, c_void_p, c_int,
                                              c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p])
_lib.register_signature("psb_lstm_head_scratch_floats", [c_int])      # small: fits the int return
_lib.register_signature("psb_cast_transpose", [c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                                               c_void_p, c_int, c_void_p])
_lib.register_signature("psb_cast_transpose_clear", [c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                                                     c_void_p, c_int, c_void_p, c_int, c_void_p])
_lib.register_signature("psb_mse_sigmoid_grad", [c_void_p, c_int, c_void_p, c_int, c_int, c_int,
                                                 c_void_p, c_void_p, c_int, c_void_p, c_int,
                                                 c_void_p, c_void_p])
_lib.register_signature("psb_gan_loss_grad", [c_void_p, c_int, c_int, c_void_p, c_void_p, c_int,
                                              c_void_p, c_void_p, c_int, c_void_p, c_void_p])
_lib.register_signature("psb_softmax_xent_grad", [c_void_p, c_int, c_void_p, c_int, c_int, c_int,
                                                  c_void_p, c_void_p, c_int, c_void_p, c_int,
                                                  c_void_p, c_void_p, c_void_p])


class LstmSaved:
    """Buffers the forward kernel fills for BPTT (preallocated once per plan)."""

    def __init__(self, B: int, T: int, I: int, H: int, device="cuda"):
        self.B, self.T, self.I, self.H = B, T, I, H
        self.ctas = (B + 15) // 16
        L = _lib.lib()
        L.psb_lstm_saved_bytes.restype = ctypes.c_size_t
        L.psb_lstm_saved_bytes.argtypes = [c_int, c_int, c_int, c_int]
        # opaque blobs in the kernels' private layout (see csrc/lstm.cu)
        self.gates = torch.zeros(L.psb_lstm_saved_bytes(B, T, H, 0) // 2, dtype=torch.bfloat16, device=device)
        self.c_all = torch.zeros(L.psb_lstm_saved_bytes(B, T, H, 1) // 4, dtype=torch.float32, device=device)
        # operands of the weight-gradient GEMM; column order (cta, t, b16) — see csrc/lstm.cu
        self.xh_t = BMat(I + H, self.ctas * T * 16, device)
        self.dg_t = BMat(4 * H, self.ctas * T * 16, device)
        self.dbias_scratch = torch.zeros(self.ctas * 4 * H + 1, dtype=torch.float32, device=device)
        self.head_scratch = None      # allocated by lstm_bwd_head on first use

    def gates_tbh(self) -> torch.Tensor:
        """Activated gates as ``[T, 4, B, H]`` (debug/tests)."""
        # [t][cta][g][half][u][8] -> [t][g][cta*16 + half*8 + j][u]
        g = self.gates.view(self.T, self.ctas, 4, 2, self.H, 8).permute(0, 2, 1, 3, 5, 4)
        return g.reshape(self.T, 4, self.ctas * 16, self.H)[:, :, : self.B]

    def xh_tb(self) -> torch.Tensor:
        """``xh_t`` un-permuted to ``[I+H, T*B]`` with column ``t*B + b`` (debug/tests)."""
        v = self.xh_t.float().view(self.I + self.H, self.ctas, self.T, 16).permute(0, 2, 1, 3)
        return v.reshape(self.I + self.H, self.T, self.ctas * 16)[:, :, : self.B].reshape(self.I + self.H, self.T * self.B)

    def c_tbh(self) -> torch.Tensor:
        """Cell state after every step as ``[T, B, H]`` (debug/tests)."""
        # [t][cta][half][2][u][4] -> [t][cta*16 + half*8 + q*4 + j][u]
        c = self.c_all.view(self.T, self.ctas, 2, 2, self.H, 4).permute(0, 1, 2, 3, 5, 4)
        return c.reshape(self.T, self.ctas * 16, self.H)[:, : self.B]


def lstm_fwd(x: torch.Tensor, wt: BMat, bias: torch.Tensor, forget_bias: float,
             h_last: Optional[BMat] = None, h_last_t: Optional[BMat] = None,
             saved: Optional[LstmSaved] = None, wt_ptr: Optional[int] = None) -> None:
    """``x`` fp32 [B,T,I]; ``wt`` = cell kernel transposed, bf16 [4H, I+H].  ``wt_ptr`` overrides
    the weight address (a peer GPU's copy) while ``wt`` still supplies the geometry."""
    B, T, I = x.shape
    H = wt.rows // 4
    check(_lib.lib().psb_lstm_fwd(
        x.data_ptr(), wt_ptr if wt_ptr is not None else wt.ptr, wt.ld, bias.data_ptr(),
        float(forget_bias), B, T, I, H,
        _ptr(h_last), h_last.ld if h_last is not None else 0,
        _ptr(h_last_t), h_last_t.ld if h_last_t is not None else 0,
        saved.gates.data_ptr() if saved else None, saved.c_all.data_ptr() if saved else None,
        saved.xh_t.ptr if saved else None, saved.xh_t.ld if saved else 0,
        current_stream_ptr()), "psb_lstm_fwd")


def lstm_bwd(w: BMat, dh_last: torch.Tensor, saved: LstmSaved, dbias: torch.Tensor,
             w_ptr: Optional[int] = None) -> None:
    """``w`` = cell kernel natural layout bf16 [I+H, 4H]; ``dh_last`` fp32 [B,H]; fills
    ``saved.dg_t`` and atomically accumulates ``dbias`` fp32 [4H]."""
    check(_lib.lib().psb_lstm_bwd(
        w_ptr if w_ptr is not None else w.ptr, w.ld, dh_last.data_ptr(), dh_last.stride(0),
        saved.gates.data_ptr(), saved.c_all.data_ptr(), saved.B, saved.T, saved.I, saved.H,
        saved.dg_t.ptr, saved.dg_t.ld, dbias.data_ptr(), saved.dbias_scratch.data_ptr(),
        current_stream_ptr()), "psb_lstm_bwd")


def lstm_bwd_head(w: BMat, saved: LstmSaved, dbias: torch.Tensor, h_last: BMat, w_out: BMat, b_out: torch.Tensor,
                  labels: torch.Tensor, d_w_out: torch.Tensor, d_b_out: torch.Tensor,
                  loss: Optional[torch.Tensor] = None, correct: Optional[torch.Tensor] = None,
                  logits: Optional[torch.Tensor] = None, w_ptr: Optional[int] = None, accumulate: bool = True,
                  zero: Optional[torch.Tensor] = None) -> None:
    """:func:`lstm_bwd` with the classifier head fused in front of it: logits, softmax-CE, their
    gradients (``d_w_out`` fp32 [H,C] and ``d_b_out`` [C] accumulate) and ``dh_last`` are computed
    inside the BPTT kernel from ``h_last`` (what :func:`lstm_fwd` wrote) and ``labels`` [B,C].
    ``accumulate=False`` overwrites the three gradients instead; ``zero`` is a tensor the kernel
    clears in its prologue (the cell-kernel gradient the split-K wgrad GEMM then accumulates into)."""
    C = labels.shape[1]
    if saved.head_scratch is None:
        n = _lib.lib().psb_lstm_head_scratch_floats(saved.B)
        saved.head_scratch = torch.zeros(n, dtype=torch.float32, device=saved.gates.device)
    hd = LstmHead(h_last.ptr, h_last.ld, w_out.ptr, w_out.ld, b_out.data_ptr(), labels.data_ptr(), labels.stride(0), C,
                  logits.data_ptr() if logits is not None else None, logits.stride(0) if logits is not None else 0,
                  loss.data_ptr() if loss is not None else None, correct.data_ptr() if correct is not None else None,
                  d_w_out.data_ptr(), d_w_out.stride(0), d_b_out.data_ptr(), saved.head_scratch.data_ptr(),
                  1 if accumulate else 0, zero.data_ptr() if zero is not None else None,
                  zero.numel() * zero.element_size() if zero is not None else 0)
    check(_lib.lib().psb_lstm_bwd_head(
        w_ptr if w_ptr is not None else w.ptr, w.ld, ctypes.byref(hd), saved.gates.data_ptr(), saved.c_all.data_ptr(),
        saved.B, saved.T, saved.I, saved.H, saved.dg_t.ptr, saved.dg_t.ld, dbias.data_ptr(),
        saved.dbias_scratch.data_ptr(), current_stream_ptr()), "psb_lstm_bwd_head")


# --------------------------------------------------------------------------- cast / losses
def cast_transpose(x: torch.Tensor, out: Optional[BMat] = None, out_t: Optional[BMat] = None,
                   out_ptr: Optional[int] = None, out_t_ptr: Optional[int] = None,
                   clear: Optional[torch.Tensor] = None) -> None:
    """fp32 [M,N] -> bf16 ``out`` [M,N] and/or ``out_t`` [N,M].  ``out_ptr``/``out_t_ptr``
    redirect the write into a sub-block of a larger matrix (same leading dimensions).  ``clear``
    is a small fp32 tensor (<= 256 elements, e.g. a loss accumulator) the kernel also zeroes."""
    M, N = x.shape
    if clear is not None:
        check(_lib.lib().psb_cast_transpose_clear(
            x.data_ptr(), x.stride(0), M, N,
            out_ptr if out_ptr is not None else _ptr(out), out.ld if out is not None else 0,
            out_t_ptr if out_t_ptr is not None else _ptr(out_t), out_t.ld if out_t is not None else 0,
            clear.data_ptr(), clear.numel(), current_stream_ptr()), "psb_cast_transpose_clear")
        return
    check(_lib.lib().psb_cast_transpose(
        x.data_ptr(), x.stride(0), M, N,
        out_ptr if out_ptr is not None else _ptr(out), out.ld if out is not None else 0,
        out_t_ptr if out_t_ptr is not None else _ptr(out_t), out_t.ld if out_t is not None else 0,
        current_stream_ptr()), "psb_cast_transpose")


def mse_sigmoid_grad(y: torch.Tensor, x: torch.Tensor, loss: torch.Tensor, dz: BMat, dz_t: BMat,
                     colsum: torch.Tensor) -> None:
    M, N = x.shape
    check(_lib.lib().psb_mse_sigmoid_grad(
        y.data_ptr(), y.stride(0), x.data_ptr(), x.stride(0), M, N, loss.data_ptr(),
        _ptr(dz), dz.ld if dz is not None else 0, _ptr(dz_t), dz_t.ld if dz_t is not None else 0,
        _ptr(colsum), current_stream_ptr()), "psb_mse_sigmoid_grad")


def gan_loss_grad(logit: torch.Tensor, B: int, losses: torch.Tensor, d_disc: BMat, d_disc_t: BMat,
                  d_gen: BMat, db_disc_out: Optional[torch.Tensor]) -> None:
    check(_lib.lib().psb_gan_loss_grad(
        logit.data_ptr(), logit.stride(0), B, losses.data_ptr(), d_disc.ptr, d_disc.ld, d_disc_t.ptr,
        d_gen.ptr, d_gen.ld, _ptr(db_disc_out), current_stream_ptr()), "psb_gan_loss_grad")


def softmax_xent_grad(logits: torch.Tensor, labels: torch.Tensor, loss: torch.Tensor,
                      dz: Optional[BMat], dz_t: Optional[BMat], colsum: Optional[torch.Tensor],
                      correct: Optional[torch.Tensor] = None) -> None:
    B, C = logits.shape
    check(_lib.lib().psb_softmax_xent_grad(
        logits.data_ptr(), logits.stride(0), labels.data_ptr(), labels.stride(0), B, C,
        loss.data_ptr(), _ptr(dz), dz.ld if dz is not None else 0,
        _ptr(dz_t), dz_t.ld if dz_t is not None else 0, _ptr(colsum), _ptr(correct),
        current_stream_ptr()), "psb_softmax_xent_grad")


_lib.register_signature("psb_uniform_noise", [c_void_p, c_int, c_int, c_int, c_float, c_float,
                                              ctypes.c_uint32, c_void_p, c_void_p, c_int, c_void_p,
                                              c_int, c_void_p])


def uniform_noise(rows: int, cols: int, lo: float, hi: float, seed: int,
                  step_counter: Optional[torch.Tensor], out: Optional[torch.Tensor] = None,
                  out_bf16: Optional[BMat] = None, out_t: Optional[BMat] = None) -> None:
    """U(lo,hi) noise generated on the device (``np.random.uniform`` of ``gan:189``)."""
    check(_lib.lib().psb_uniform_noise(
        _ptr(out), rows, cols, out.stride(0) if out is not None else 0, lo, hi, seed & 0xFFFFFFFF,
        _ptr(step_counter), _ptr(out_bf16), out_bf16.ld if out_bf16 is not None else 0,
        _ptr(out_t), out_t.ld if out_t is not None else 0, current_stream_ptr()), "psb_uniform_noise")


# --------------------------------------------------------------------------- large-H LSTM path
_lib.register_signature("psb_lstm_cell_fwd", [c_void_p, c_int, c_void_p, c_int, c_void_p, c_float, c_void_p,
                                              c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p])
_lib.register_signature("psb_lstm_cell_bwd", [c_void_p, c_int, c_int, ctypes.c_longlong, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p])
_lib.register_signature("psb_transpose_bf16", [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_void_p])


def lstm_cell_fwd(gx: torch.Tensor, gh: torch.Tensor, bias: torch.Tensor, forget_bias: float,
                  c: torch.Tensor, h_out: BMat, gates_save: Optional[torch.Tensor] = None,
                  c_save: Optional[torch.Tensor] = None) -> None:
    B, H = c.shape
    check(_lib.lib().psb_lstm_cell_fwd(gx.data_ptr(), gx.stride(0), gh.data_ptr(), gh.stride(0),
                                       bias.data_ptr(), float(forget_bias), c.data_ptr(), h_out.ptr, h_out.ld,
                                       _ptr(gates_save), _ptr(c_save), B, H, current_stream_ptr()),
          "psb_lstm_cell_fwd")


def lstm_cell_bwd(dh_rec: Optional[torch.Tensor], dy: Optional[torch.Tensor], gates: torch.Tensor,
                  c_new: torch.Tensor, c_prev: Optional[torch.Tensor], dc: torch.Tensor, dg: BMat,
                  dbias: Optional[torch.Tensor]) -> None:
    """``dh_rec`` may be ``[B,H]`` or ``[parts,B,H]`` (split-K partial sums, added here)."""
    B, H = dc.shape
    parts, pstride = 1, 0
    if dh_rec is not None and dh_rec.dim() == 3:
        parts, pstride = dh_rec.shape[0], dh_rec.stride(0)
    check(_lib.lib().psb_lstm_cell_bwd(_ptr(dh_rec), dh_rec.stride(-2) if dh_rec is not None else 0, parts, pstride,
                                       _ptr(dy), dy.stride(0) if dy is not None else 0, gates.data_ptr(),
                                       c_new.data_ptr(), _ptr(c_prev), dc.data_ptr(), dg.ptr, dg.ld,
                                       _ptr(dbias), B, H, current_stream_ptr()), "psb_lstm_cell_bwd")


def transpose_bf16(src: BMat, dst: BMat) -> None:
    check(_lib.lib().psb_transpose_bf16(src.ptr, src.ld, src.rows, src.cols, dst.ptr, dst.ld,
                                        current_stream_ptr()), "psb_transpose_bf16")


_lib.register_signature("psb_lstm_step_fwd", [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                                              c_void_p, c_float, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p])


def lstm_step_fwd(h_prev: BMat, wh_t: BMat, gx: torch.Tensor, bias: torch.Tensor, forget_bias: float,
                  c: torch.Tensor, h_out: BMat, gates_save: Optional[torch.Tensor] = None,
                  c_save: Optional[torch.Tensor] = None) -> None:
    """Recurrent GEMM + cell pointwise of one timestep fused in one kernel (large-H path)."""
    B, H = c.shape
    check(_lib.lib().psb_lstm_step_fwd(h_prev.ptr, h_prev.ld, wh_t.ptr, wh_t.ld, B, H, gx.data_ptr(), gx.stride(0),
                                       bias.data_ptr(), float(forget_bias), c.data_ptr(), h_out.ptr, h_out.ld,
                                       _ptr(gates_save), _ptr(c_save), current_stream_ptr()), "psb_lstm_step_fwd")


_lib.register_signature("psb_memset_zero", [c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t, c_void_p])


def zero_(t: torch.Tensor) -> None:
    """Clear a contiguous tensor or a 2-D view with contiguous rows on the current stream with a memset
    node (no kernel launch)."""
    es = t.element_size()
    if t.is_contiguous():
        check(_lib.lib().psb_memset_zero(t.data_ptr(), t.numel() * es, t.numel() * es, 1, current_stream_ptr()), "psb_memset_zero")
        return
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError("zero_: need a contiguous tensor or a 2-D view with contiguous rows")
    check(_lib.lib().psb_memset_zero(t.data_ptr(), t.stride(0) * es, t.shape[1] * es, t.shape[0], current_stream_ptr()),
          "psb_memset_zero")
