"""fp32 torch implementation of every op the models use, with hand-written backward.

Two jobs:

* it is the compute path of the CPU plumbing configuration ("1 ps + 2 workers on CPU
  localhost", BASELINE.json) where no GPU exists;
* it is the numerical oracle for the sm_100a kernels (``tests/`` compare each kernel with
  the function of the same name here) — and its own backward formulas are checked against
  ``torch.autograd`` in ``tests/test_reference_ops.py``.

Op list = SURVEY §2.6 K1-K9: dense+bias+activation (``encoder:69-84``, ``gan:77-92``,
``lstm:53``), BasicLSTMCell/static_rnn (``lstm:50-52``), MSE (``encoder:135``), GAN
log-losses (``gan:142-143``), softmax cross-entropy + accuracy (``lstm:92-100``).
"""
from __future__ import annotations

from typing import Dict, NamedTuple, Optional, Tuple

import torch
import torch.nn.functional as F

ACT_NONE, ACT_SIGMOID, ACT_RELU = 0, 1, 2


# ------------------------------------------------------------------ dense (K1-K4)
def dense_fwd(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], act: int) -> torch.Tensor:
    """``act(x @ w + b)`` with ``w`` stored ``[in, out]`` like the reference variables."""
    z = x @ w
    if b is not None:
        z = z + b
    if act == ACT_SIGMOID:
        return torch.sigmoid(z)
    if act == ACT_RELU:
        return torch.relu(z)
    return z


def act_grad(y: torch.Tensor, dy: torch.Tensor, act: int) -> torch.Tensor:
    """``dZ = dY * act'(Z)`` expressed through the saved output ``y``."""
    if act == ACT_SIGMOID:
        return dy * y * (1.0 - y)
    if act == ACT_RELU:
        return dy * (y > 0).to(dy.dtype)
    return dy


def dense_bwd(x: torch.Tensor, w: torch.Tensor, dz: torch.Tensor,
              need_dx: bool = True) -> Tuple[Optional[torch.Tensor], torch.Tensor, torch.Tensor]:
    """Given ``dZ`` return ``(dX, dW, db)``: ``dX = dZ Wᵀ``, ``dW = Xᵀ dZ``, ``db = Σ_rows dZ``."""
    dx = dz @ w.t() if need_dx else None
    return dx, x.t() @ dz, dz.sum(0)


# ------------------------------------------------------------------ losses (K6-K9)
def mse_loss(y_pred: torch.Tensor, y_true: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``mean((y_true - y_pred)^2)`` over all elements and its gradient w.r.t. ``y_pred``."""
    diff = y_pred - y_true
    return (diff * diff).mean(), diff * (2.0 / diff.numel())


def gan_losses(logit_real: torch.Tensor, logit_fake: torch.Tensor):
    """GAN log-losses computed from the discriminator *logits* (stable form of
    ``-mean(log D_fake)`` and ``-mean(log D_real + log(1 - D_fake))``, ``gan:142-143``).

    Returns ``gen_loss, disc_loss, d_gen/d_logit_fake, d_disc/d_logit_real,
    d_disc/d_logit_fake``.
    """
    n = logit_real.numel()
    gen_loss = F.softplus(-logit_fake).mean()
    disc_loss = (F.softplus(-logit_real) + F.softplus(logit_fake)).mean()
    g_fake_gen = -(1.0 - torch.sigmoid(logit_fake)) / n
    g_real_disc = -(1.0 - torch.sigmoid(logit_real)) / n
    g_fake_disc = torch.sigmoid(logit_fake) / n
    return gen_loss, disc_loss, g_fake_gen, g_real_disc, g_fake_disc


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``mean(softmax_cross_entropy_with_logits(logits, labels))`` with one-hot (or soft)
    ``labels`` and its gradient w.r.t. ``logits``."""
    logp = torch.log_softmax(logits, dim=1)
    loss = -(labels * logp).sum(1).mean()
    grad = (torch.softmax(logits, dim=1) * labels.sum(1, keepdim=True) - labels) / logits.shape[0]
    return loss, grad


def accuracy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    return (logits.argmax(1) == labels.argmax(1)).float().mean()


# ------------------------------------------------------------------ LSTM (K5)
class LstmSaved(NamedTuple):
    xh: torch.Tensor     # [T, B, I+H]  concat([x_t, h_{t-1}])
    gates: torch.Tensor  # [T, B, 4H]   activated: σ(i), tanh(j), σ(f+fb), σ(o)
    c: torch.Tensor      # [T+1, B, H]  c[0] = 0
    h: torch.Tensor      # [T+1, B, H]  h[0] = 0


def lstm_fwd(x: torch.Tensor, kernel: torch.Tensor, bias: torch.Tensor,
             forget_bias: float = 1.0) -> Tuple[torch.Tensor, LstmSaved]:
    """``BasicLSTMCell`` unrolled by ``static_rnn`` from a zero state.

    ``x``: [B, T, I]; ``kernel``: [I+H, 4H]; gate order i, j, f, o;
    ``c' = c·σ(f + forget_bias) + σ(i)·tanh(j)``; ``h' = tanh(c')·σ(o)``.
    Returns the last hidden state [B, H] and what backward needs.
    """
    B, T, I = x.shape
    H = kernel.shape[1] // 4
    h = x.new_zeros(B, H)
    c = x.new_zeros(B, H)
    xh_all, gates_all, c_all, h_all = [], [], [c], [h]
    for t in range(T):
        xh = torch.cat([x[:, t, :], h], dim=1)
        g = xh @ kernel + bias
        i, j, f, o = g.split(H, dim=1)
        i, j, f, o = torch.sigmoid(i), torch.tanh(j), torch.sigmoid(f + forget_bias), torch.sigmoid(o)
        c = c * f + i * j
        h = torch.tanh(c) * o
        xh_all.append(xh)
        gates_all.append(torch.cat([i, j, f, o], dim=1))
        c_all.append(c)
        h_all.append(h)
    saved = LstmSaved(torch.stack(xh_all), torch.stack(gates_all), torch.stack(c_all), torch.stack(h_all))
    return h, saved


def lstm_bwd(dh_last: Optional[torch.Tensor], saved: LstmSaved, kernel: torch.Tensor,
             dh_seq: Optional[torch.Tensor] = None, need_dx: bool = False):
    """BPTT for :func:`lstm_fwd`.  ``dh_last`` [B,H] is the gradient reaching the last hidden
    state (the reference's case, ``outputs[-1]``); ``dh_seq`` [T,B,H] optionally adds a gradient
    at every step (a stacked layer above).  Returns ``(d_kernel [I+H,4H], d_bias [4H])`` and,
    with ``need_dx``, also ``d_x [B,T,I]``."""
    T, B, IH = saved.xh.shape
    H = kernel.shape[1] // 4
    I = IH - H
    dh = torch.zeros(B, H, dtype=kernel.dtype, device=kernel.device) if dh_last is None else dh_last
    dc = torch.zeros_like(dh)
    dk = torch.zeros_like(kernel)
    db = kernel.new_zeros(4 * H)
    wh_t = kernel[I:, :].t()        # [4H, H]
    wx_t = kernel[:I, :].t()        # [4H, I]
    dxs = []
    for t in range(T - 1, -1, -1):
        if dh_seq is not None:
            dh = dh + dh_seq[t]
        i, j, f, o = saved.gates[t].split(H, dim=1)
        c_new, c_old = saved.c[t + 1], saved.c[t]
        tc = torch.tanh(c_new)
        do = dh * tc
        dc = dc + dh * o * (1.0 - tc * tc)
        di, dj, df = dc * j, dc * i, dc * c_old
        dg = torch.cat([di * i * (1 - i), dj * (1 - j * j), df * f * (1 - f), do * o * (1 - o)], dim=1)
        dk += saved.xh[t].t() @ dg
        db += dg.sum(0)
        dh = dg @ wh_t
        if need_dx:
            dxs.append(dg @ wx_t)
        dc = dc * f
    if need_dx:
        return dk, db, torch.stack(dxs[::-1], dim=1)
    return dk, db
