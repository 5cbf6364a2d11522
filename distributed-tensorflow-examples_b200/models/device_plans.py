"""Device execution plans: the three models as fixed sequences of sm_100a kernels.

A plan owns every activation / gradient-of-activation buffer (allocated once), reads the
bf16/fp32 compute copies of the parameters from the :class:`~psb200.store.nvl.NvlStore`
arena and leaves the fp32 parameter gradients in the store's flat gradient buffer.  No
allocation and no host synchronisation happens inside ``forward_backward`` — the trainer
captures it (plus the gradient push) into one CUDA graph.

Layout rules every plan follows (all GEMM operands are K-major bf16):
  forward   Y[B,N]  = act(X[B,K] Wt[N,K]^T + b)       -> Y and Y^T
  dgrad     dX[B,K] = (dZ[B,N] W[K,N]^T) * act'(X)    -> dZ_prev and dZ_prev^T, db via colsum
  wgrad     dW[K,N] = X^T[K,B] dZ^T[N,B]^T            -> fp32 into the gradient buffer
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from ..ops import cuda as K
from ..ops.cuda import ACT_NONE, ACT_RELU, ACT_SIGMOID, BMat
from ..store.nvl import NvlStore
from . import encoder as enc_mod
from . import gan as gan_mod
from . import lstm as lstm_mod


class DevicePlan:
    """Common surface: static input tensors, ``forward_backward()``, result tensors."""

    inputs: Dict[str, torch.Tensor]
    losses: Dict[str, torch.Tensor]

    def __init__(self, model, store: NvlStore, batch: int):
        self.model, self.store, self.B = model, store, batch
        self.dev = store.device
        self.inputs = {}
        self.losses = {}
        # independent branches of the backward pass (weight gradients vs. the dgrad chain) are
        # issued on a side stream; inside the captured graph they become parallel branches
        self.side = torch.cuda.Stream(device=self.dev)

    def _fork(self) -> None:
        """Side stream may start once everything issued so far on the main stream is done."""
        self.side.wait_stream(torch.cuda.current_stream())

    def _join(self) -> None:
        torch.cuda.current_stream().wait_stream(self.side)

    def _zero_grads_on_side(self) -> None:
        """Clear the gradient buffer on the side stream: the forward kernels do not touch it, so the
        fill leaves the critical path; callers ``_join()`` before the first gradient is produced."""
        self._fork()
        with torch.cuda.stream(self.side):
            self.store.zero_grads()

    def _name(self, key: str) -> str:
        return self.model.var[key].name

    def forward_backward(self) -> None:
        raise NotImplementedError

    def forward(self) -> None:
        """Inference-only pass (LSTM accuracy)."""
        raise NotImplementedError

    def kernels_per_step(self) -> int:
        raise NotImplementedError


class EncoderPlan(DevicePlan):
    """``encoder:67-85,128-137`` — 4 fused dense+sigmoid forward GEMMs, MSE kernel, 4 wgrad and
    3 dgrad GEMMs (sigmoid' and bias-gradient fused in their epilogues)."""

    def __init__(self, model, store, batch=enc_mod.batch_size):
        super().__init__(model, store, batch)
        B, dev = batch, self.dev
        dims = [enc_mod.num_input, enc_mod.num_hidden_1, enc_mod.num_hidden_2, enc_mod.num_hidden_1,
                enc_mod.num_input]
        self.dims = dims
        self.inputs["x"] = torch.zeros(B, dims[0], device=dev)
        self.act = [BMat(B, dims[0], dev)] + [BMat(B, d, dev) for d in dims[1:4]]
        self.act_t = [BMat(d, B, dev) for d in dims[0:4]]
        self.y = torch.zeros(B, dims[4], device=dev)
        self.dz = [None] + [BMat(B, d, dev) for d in dims[1:]]        # dz[l] = grad at output of layer l
        self.dz_t = [None] + [BMat(d, B, dev) for d in dims[1:]]
        self.losses["loss"] = torch.zeros((), device=dev)
        self.wkeys = [w for w, _ in enc_mod.LAYERS]
        self.bkeys = [b for _, b in enc_mod.LAYERS]
        # the bias gradients are accumulated (column sums): the push kernel clears them after use
        store.zero_after_push([self._name(b) for b in self.bkeys])

    def forward(self, need_t: bool = True, clear: Optional[torch.Tensor] = None) -> None:
        s, n = self.store, self._name
        K.cast_transpose(self.inputs["x"], self.act[0], self.act_t[0] if need_t else None, clear=clear)
        for l in range(3):
            K.gemm(self.act[l], s.wt_src(n(self.wkeys[l])), bias=s.vec_src(n(self.bkeys[l])), bias_ready=s.ready_for(n(self.bkeys[l])), act=ACT_SIGMOID,
                   out=self.act[l + 1], out_t=self.act_t[l + 1] if need_t else None, b_static=True)
        K.gemm(self.act[3], s.wt_src(n(self.wkeys[3])), bias=s.vec_src(n(self.bkeys[3])), bias_ready=s.ready_for(n(self.bkeys[3])), act=ACT_SIGMOID,
               out_f32=self.y, b_static=True)

    def forward_backward(self) -> None:
        s, n = self.store, self._name
        # no fill kernels in the step: the input cast clears the loss accumulator, the previous push
        # cleared the accumulated (bias) gradients, every other gradient is overwritten
        local_zero = not s.owner_zeroes_grads and not s.zero_in_before_grads
        if local_zero:
            self._zero_grads_on_side()
        self.forward(clear=self.losses["loss"])
        if local_zero:
            self._join()
        s.before_grads()
        K.mse_sigmoid_grad(self.y, self.inputs["x"], self.losses["loss"], self.dz[4], self.dz_t[4],
                           s.grad(n(self.bkeys[3])))
        for l in range(3, -1, -1):
            # dW_l = act_{l}^T dz_{l+1}  (side stream: only the dgrad chain is on the critical path)
            self._fork()
            with torch.cuda.stream(self.side):
                K.gemm(self.act_t[l], self.dz_t[l + 1], out_f32=s.grad(n(self.wkeys[l])))
            if l > 0:
                K.gemm(self.dz[l + 1], s.w_src(n(self.wkeys[l])), yprev=self.act[l], dact=1,
                       out=self.dz[l], out_t=self.dz_t[l], colsum=s.grad(n(self.bkeys[l - 1])), b_static=True)
        self._join()

    def kernels_per_step(self) -> int:
        return 1 + 4 + 1 + 4 + 3


class GanPlan(DevicePlan):
    """``gan:76-93,139-159`` — G forward, D forward on [real; fake] in one batch of 2B rows, the
    fused log-loss kernel, then the discriminator-variable path and the generator-variable path
    (through D) as dgrad/wgrad GEMMs."""

    def __init__(self, model, store, batch=gan_mod.batch_size, device_noise: bool = True, seed: int = 0):
        super().__init__(model, store, batch)
        B, dev = batch, self.dev
        nz, gh, im, dh = gan_mod.noise_dim, gan_mod.gen_hidden_dim, gan_mod.image_dim, gan_mod.disc_hidden_dim
        self.device_noise, self.seed = device_noise, seed
        self.noise_counter = torch.zeros(1, dtype=torch.int32, device=dev)
        self.inputs["x"] = torch.zeros(B, im, device=dev)
        self.inputs["z"] = torch.zeros(B, nz, device=dev)
        self.Z, self.ZT = BMat(B, nz, dev), BMat(nz, B, dev)
        self.GH, self.GHT = BMat(B, gh, dev), BMat(gh, B, dev)
        self.Din, self.DinT = BMat(2 * B, im, dev), BMat(im, 2 * B, dev)
        self.DH, self.DHT = BMat(2 * B, dh, dev), BMat(dh, 2 * B, dev)
        self.logit = torch.zeros(2 * B, 1, device=dev)
        self.losses_t = torch.zeros(2, device=dev)
        self.losses["gen_loss"], self.losses["disc_loss"] = self.losses_t[0], self.losses_t[1]
        self.d_disc, self.d_disc_t, self.d_gen = BMat(2 * B, 1, dev), BMat(1, 2 * B, dev), BMat(B, 1, dev)
        self.dZ1, self.dZ1T = BMat(2 * B, dh, dev), BMat(dh, 2 * B, dev)
        self.dZ1f = BMat(B, dh, dev)
        self.dZo, self.dZoT = BMat(B, im, dev), BMat(im, B, dev)
        self.dZh, self.dZhT = BMat(B, gh, dev), BMat(gh, B, dev)
        # views of the fake half
        self.Din_fake, self.DinT_fake = self.Din.rows_view(B, 2 * B), self.DinT.cols_view(B, 2 * B)
        self.Din_real, self.DinT_real = self.Din.rows_view(0, B), self.DinT.cols_view(0, B)
        self.DH_fake = self.DH.rows_view(B, 2 * B)
        # accumulated column sums; b_disc_out is overwritten by the loss kernel
        store.zero_after_push([self._name(k) for k in ("b_disc_hidden1", "b_gen_out", "b_gen_hidden1")])

    def forward(self) -> None:
        s, n, B = self.store, self._name, self.B
        K.cast_transpose(self.inputs["x"], self.Din_real, self.DinT_real)
        if self.device_noise:
            K.uniform_noise(B, gan_mod.noise_dim, -1.0, 1.0, self.seed + 7919 * s.rank,
                            self.noise_counter, out=self.inputs["z"], out_bf16=self.Z, out_t=self.ZT)
        else:
            K.cast_transpose(self.inputs["z"], self.Z, self.ZT)
        K.gemm(self.Z, s.wt_src(n("w_gen_hidden1")), bias=s.vec_src(n("b_gen_hidden1")), bias_ready=s.ready_for(n("b_gen_hidden1")), act=ACT_RELU,
               out=self.GH, out_t=self.GHT, b_static=True)
        K.gemm(self.GH, s.wt_src(n("w_gen_out")), bias=s.vec_src(n("b_gen_out")), bias_ready=s.ready_for(n("b_gen_out")), act=ACT_SIGMOID,
               out=self.Din_fake, out_t=self.DinT_fake, b_static=True)
        K.gemm(self.Din, s.wt_src(n("w_disc_hidden1")), bias=s.vec_src(n("b_disc_hidden1")), bias_ready=s.ready_for(n("b_disc_hidden1")), act=ACT_RELU,
               out=self.DH, out_t=self.DHT, b_static=True)
        K.gemm(self.DH, s.wt_src(n("w_disc_out")), bias=s.vec_src(n("b_disc_out")), bias_ready=s.ready_for(n("b_disc_out")), act=ACT_NONE,
               out_f32=self.logit, b_static=True)

    def forward_backward(self, disc: bool = True, gen: bool = True) -> None:
        s, n, B = self.store, self._name, self.B
        local_zero = not s.owner_zeroes_grads and not s.zero_in_before_grads
        if local_zero:
            self._zero_grads_on_side()
        self.forward()
        if local_zero:
            self._join()
        s.before_grads()
        K.gan_loss_grad(self.logit, B, self.losses_t, self.d_disc, self.d_disc_t, self.d_gen,
                        s.grad(n("b_disc_out")) if disc else None)
        if disc:
            # discriminator-variable path on the side stream, generator-variable path on the main one
            self._fork()
            with torch.cuda.stream(self.side):
                K.gemm(self.DHT, self.d_disc_t, out_f32=s.grad(n("w_disc_out")))
                K.gemm(self.d_disc, s.w_src(n("w_disc_out")), yprev=self.DH, dact=2, out=self.dZ1, out_t=self.dZ1T,
                       colsum=s.grad(n("b_disc_hidden1")), b_static=True)
                K.gemm(self.DinT, self.dZ1T, out_f32=s.grad(n("w_disc_hidden1")))
        if gen:
            K.gemm(self.d_gen, s.w_src(n("w_disc_out")), yprev=self.DH_fake, dact=2, out=self.dZ1f, b_static=True)
            K.gemm(self.dZ1f, s.w_src(n("w_disc_hidden1")), yprev=self.Din_fake, dact=1, out=self.dZo,
                   out_t=self.dZoT, colsum=s.grad(n("b_gen_out")), b_static=True)
            K.gemm(self.GHT, self.dZoT, out_f32=s.grad(n("w_gen_out")))
            K.gemm(self.dZo, s.w_src(n("w_gen_out")), yprev=self.GH, dact=2, out=self.dZh, out_t=self.dZhT,
                   colsum=s.grad(n("b_gen_hidden1")), b_static=True)
            K.gemm(self.ZT, self.dZhT, out_f32=s.grad(n("w_gen_hidden1")))
        if disc:
            self._join()

    def kernels_per_step(self) -> int:
        return 2 + 4 + 1 + 3 + 5


class LstmPlan(DevicePlan):
    """``lstm:49-53,89-100`` — persistent LSTM forward kernel, persistent BPTT kernel with the
    classifier head (logits, softmax-CE, head gradients) fused into its prologue, one big wgrad
    GEMM for the cell kernel.  ``fuse_head=False`` runs the head as separate kernels instead (head
    GEMM, softmax-CE kernel, head wgrad/dgrad GEMMs): that is also the evaluation path."""

    def __init__(self, model, store, batch=lstm_mod.batch_size, pull_from_owner: bool = False,
                 fuse_head: bool = True):
        super().__init__(model, store, batch)
        self.fuse_head = fuse_head and model.C <= 16
        B, dev = batch, self.dev
        T, I, H, C = model.T, model.I, model.H, model.C
        self.inputs["x"] = torch.zeros(B, T, I, device=dev)
        self.inputs["y"] = torch.zeros(B, C, device=dev)
        self.h_last, self.h_last_t = BMat(B, H, dev), BMat(H, B, dev)
        self.saved = K.LstmSaved(B, T, I, H, dev)
        self.logits = torch.zeros(B, C, device=dev)
        self.losses["loss"] = torch.zeros((), device=dev)
        self.correct = torch.zeros((), device=dev)
        self.dlogits, self.dlogits_t = BMat(B, C, dev), BMat(C, B, dev)
        self.dh = torch.zeros(B, H, device=dev)
        # bcast=False stores keep the cell kernel only in its owner's arena: the LSTM kernels
        # then read it through the owner's peer pointer (pull fused into the consumer)
        pull_from_owner = pull_from_owner or not store.bcast
        self.pull_from_owner = pull_from_owner
        kname = self._name("kernel")
        self._k_owner = store.owner(kname) if pull_from_owner else None

    def _kernel_ptrs(self):
        s, kname = self.store, self._name("kernel")
        if self._k_owner is None:
            return None, None
        return s.wt(kname, self._k_owner).ptr, s.w(kname, self._k_owner).ptr

    def forward(self, train: bool = False) -> None:
        s, n = self.store, self._name
        s.wait_all_params()          # pull mode: the LSTM kernels read the owner's copies without flags
        wt_ptr, _ = self._kernel_ptrs()
        K.lstm_fwd(self.inputs["x"], s.wt(n("kernel")), s.vec_src(n("bias")), lstm_mod.forget_bias,
                   self.h_last, self.h_last_t if train else None, self.saved if train else None,
                   wt_ptr=wt_ptr)
        if not (train and self.fuse_head):
            K.gemm(self.h_last, s.wt_src(n("out_w")), bias=s.vec_src(n("out_b")), bias_ready=s.ready_for(n("out_b")), out_f32=self.logits, b_static=True)

    def evaluate(self) -> None:
        """Fills ``losses['loss']`` and ``correct`` for the current inputs without touching
        gradients (``sess.run(accuracy, ...)``, ``lstm:134-138``)."""
        self.forward(train=False)
        K.softmax_xent_grad(self.logits, self.inputs["y"], self.losses["loss"], None, None, None,
                            self.correct)

    def forward_backward(self) -> None:
        s, n = self.store, self._name
        _, w_ptr = self._kernel_ptrs()
        if self.fuse_head:
            # no separate clear of the gradient buffer: the BPTT kernel overwrites the three small
            # gradients and zeroes the cell-kernel gradient (what the split-K GEMM below adds into)
            self.forward(train=True)
            s.before_grads()
            K.lstm_bwd_head(s.w(n("kernel")), self.saved, s.grad(n("bias")), self.h_last, s.w_src(n("out_w")),
                            s.vec_src(n("out_b")), self.inputs["y"], s.grad(n("out_w")), s.grad(n("out_b")),
                            loss=self.losses["loss"], correct=self.correct, logits=self.logits, w_ptr=w_ptr,
                            accumulate=False, zero=s.grad(n("kernel")))
        else:
            self.forward(train=True)
            s.before_grads()
            s.zero_grads()
            K.softmax_xent_grad(self.logits, self.inputs["y"], self.losses["loss"], self.dlogits,
                                self.dlogits_t, s.grad(n("out_b")), self.correct)
            self._fork()
            with torch.cuda.stream(self.side):
                K.gemm(self.h_last_t, self.dlogits_t, out_f32=s.grad(n("out_w")))
            K.gemm(self.dlogits, s.w_src(n("out_w")), out_f32=self.dh, b_static=True)
            K.lstm_bwd(s.w(n("kernel")), self.dh, self.saved, s.grad(n("bias")), w_ptr=w_ptr)
        # reduction over all (t, b) = 3584 columns: two-way split-K; the two partial sums are added
        # atomically into the zeroed gradient buffer (two addends commute: still deterministic)
        K.gemm(self.saved.xh_t, self.saved.dg_t, out_f32=s.grad(n("kernel")), split_k=2)
        if not self.fuse_head:
            self._join()

    def kernels_per_step(self) -> int:
        return 3 if self.fuse_head else 7


class ScaledLstmPlan(DevicePlan):
    """Stacked large-H LSTM: per layer one big GEMM for the input projections of all timesteps,
    then per timestep a full-chip GEMM ``h_{t-1} Wh`` + the pointwise cell kernel; backward mirrors
    it and batches both weight gradients and the gradient to the layer below into large GEMMs.
    Input ``x`` is time-major fp32 ``[T, B, I]``."""

    def __init__(self, model, store, batch=None, fuse_fwd: bool = True):
        super().__init__(model, store, batch or model.batch_size)
        B, dev = self.B, self.dev
        T, I, H, L, C = model.T, model.I, model.H, model.L, model.C
        TB = T * B
        self.inputs["x"] = torch.zeros(T, B, I, device=dev)
        self.inputs["y"] = torch.zeros(B, C, device=dev)
        self.X0, self.X0T = BMat(TB, I, dev), BMat(I, TB, dev)
        self.Hall = [BMat(TB, H, dev) for _ in range(L)]             # h_t of every step (time-major rows)
        self.HprevT = BMat(H, TB, dev)                               # [h_{t-1}]^T, reused per layer
        self.XinT = BMat(H, TB, dev) if L > 1 else None              # layer input transposed (layers >= 1)
        self.gx = torch.zeros(TB, 4 * H, device=dev)                 # x-part pre-activations (reused per layer)
        self.gh = torch.zeros(B, 4 * H, device=dev)
        self.gates = [torch.zeros(TB, 4 * H, dtype=torch.bfloat16, device=dev) for _ in range(L)]
        self.c_all = [torch.zeros(TB, H, device=dev) for _ in range(L)]
        self.c = torch.zeros(B, H, device=dev)
        self.h0 = BMat(B, H, dev)                                    # zero initial state
        self.dG, self.dGT = BMat(TB, 4 * H, dev), BMat(4 * H, TB, dev)
        # The recurrent GEMMs are skinny (M = B).  Measured in situ at H=4096, T=512, B=64 (ms/step):
        # auto tiles 156.1 | fwd BN=128 151.1 | + bwd split-K 2, BN=64 127.5 | bwd BN=128 133.0.
        # One wave of <= 148 CTAs keeps programmatic dependent launch on (see gemm.cu); dh = dG Wh^T
        # has a 4H-long reduction, so it is split in two partial buffers that the cell kernel adds.
        big = H >= 2048
        self.split_k = 2 if big else 1
        self.bn_fwd = 128 if big else 0
        self.bn_bwd = 64 if big else 0
        # forward timestep = ONE kernel (recurrent GEMM with the cell pointwise in its epilogue)
        self.fuse_fwd = fuse_fwd and H % 32 == 0
        self.dh_rec = torch.zeros(self.split_k, B, H, device=dev)
        self.dc = torch.zeros(B, H, device=dev)
        self.dY = torch.zeros(TB, H, device=dev) if L > 1 else None  # gradient arriving from the layer above
        self.logits = torch.zeros(B, C, device=dev)
        self.losses["loss"] = torch.zeros((), device=dev)
        self.correct = torch.zeros((), device=dev)
        self.dlogits, self.dlogits_t = BMat(B, C, dev), BMat(C, B, dev)
        self.dh_top = torch.zeros(B, H, device=dev)
        self.h_last_t = BMat(H, B, dev)
        # accumulated gradients (bias column sums); the big kernels' gradients are overwritten by their GEMMs
        store.zero_after_push([self._name("out_b")] + [self._name("bias%d" % l) for l in range(L)])

    def _rows(self, m: BMat, t: int) -> BMat:
        return m.rows_view(t * self.B, (t + 1) * self.B)

    def forward(self, train: bool = True) -> None:
        s, n, mdl, B = self.store, self._name, self.model, self.B
        T, H = mdl.T, mdl.H
        s.wait_all_params()          # pull mode: column views of the kernels carry no readiness flags
        K.cast_transpose(self.inputs["x"].view(T * B, mdl.I), self.X0, self.X0T if train else None)
        inp = self.X0
        for l in range(mdl.L):
            I = mdl.in_dim(l)
            wt = s.wt_src(n("kernel%d" % l))                  # [4H, I+H]
            wx_t, wh_t = wt.cols_view(0, I), wt.cols_view(I, I + H)
            bias = s.vec_src(n("bias%d" % l))
            K.gemm(inp, wx_t, out_f32=self.gx, b_static=True)                # all timesteps at once
            K.zero_(self.c)
            for t in range(T):
                hprev = self.h0 if t == 0 else self._rows(self.Hall[l], t - 1)
                if self.fuse_fwd:
                    K.lstm_step_fwd(hprev, wh_t, self.gx[t * B:(t + 1) * B], bias, 1.0, self.c,
                                    self._rows(self.Hall[l], t),
                                    self.gates[l][t * B:(t + 1) * B] if train else None,
                                    self.c_all[l][t * B:(t + 1) * B] if train else None)
                    continue
                K.gemm(hprev, wh_t, out_f32=self.gh, block_n=self.bn_fwd, b_static=True)
                K.lstm_cell_fwd(self.gx[t * B:(t + 1) * B], self.gh, bias, 1.0, self.c, self._rows(self.Hall[l], t),
                                self.gates[l][t * B:(t + 1) * B] if train else None,
                                self.c_all[l][t * B:(t + 1) * B] if train else None)
            inp = self.Hall[l]
        h_last = self._rows(self.Hall[-1], T - 1)
        K.gemm(h_last, s.wt_src(n("out_w")), bias=s.vec_src(n("out_b")), bias_ready=s.ready_for(n("out_b")), out_f32=self.logits, b_static=True)

    def forward_backward(self) -> None:
        s, n, mdl, B = self.store, self._name, self.model, self.B
        T, H = mdl.T, mdl.H
        TB = T * B
        self.forward(train=True)
        s.before_grads()
        if not s.owner_zeroes_grads and not s.zero_in_before_grads:
            s.zero_grads()
        K.softmax_xent_grad(self.logits, self.inputs["y"], self.losses["loss"], self.dlogits, self.dlogits_t,
                            s.grad(n("out_b")), self.correct)
        h_last = self._rows(self.Hall[-1], T - 1)
        K.transpose_bf16(h_last, self.h_last_t)
        K.gemm(self.h_last_t, self.dlogits_t, out_f32=s.grad(n("out_w")))
        K.gemm(self.dlogits, s.w_src(n("out_w")), out_f32=self.dh_top, b_static=True)
        for l in range(mdl.L - 1, -1, -1):
            I = mdl.in_dim(l)
            w = s.w_src(n("kernel%d" % l))                    # [I+H, 4H]
            wx, wh = w.rows_view(0, I), w.rows_view(I, I + H)
            gk = s.grad(n("kernel%d" % l))
            K.zero_(self.dc)
            top = l == mdl.L - 1
            for t in range(T - 1, -1, -1):
                dy = None
                if top:
                    dy = self.dh_top if t == T - 1 else None
                else:
                    dy = self.dY[t * B:(t + 1) * B]
                K.lstm_cell_bwd(self.dh_rec if t < T - 1 else None, dy, self.gates[l][t * B:(t + 1) * B],
                                self.c_all[l][t * B:(t + 1) * B],
                                self.c_all[l][(t - 1) * B:t * B] if t > 0 else None, self.dc,
                                self._rows(self.dG, t), s.grad(n("bias%d" % l)))
                if t > 0:
                    K.gemm(self._rows(self.dG, t), wh, out_f32=self.dh_rec[0], split_k=self.split_k,
                           split_stride=self.dh_rec.stride(0), block_n=self.bn_bwd, b_static=True)
            # weight gradients of this layer in two large GEMMs over all (t, b)
            K.transpose_bf16(self.dG, self.dGT)
            # [h_{t-1}]^T: column block t holds h_{t-1}; block 0 is the zero state
            K.zero_(self.HprevT.t[:, :B])
            if T > 1:
                K.transpose_bf16(self.Hall[l].rows_view(0, (T - 1) * B), self.HprevT.cols_view(B, TB))
            K.gemm(self.HprevT, self.dGT, out_f32=gk[I:])
            if l == 0:
                K.gemm(self.X0T, self.dGT, out_f32=gk[:I])
            else:
                K.transpose_bf16(self.Hall[l - 1], self.XinT)
                K.gemm(self.XinT, self.dGT, out_f32=gk[:I])
                K.gemm(self.dG, wx, out_f32=self.dY, b_static=True)          # gradient to the layer below, all timesteps

    def evaluate(self) -> None:
        self.forward(train=False)
        K.softmax_xent_grad(self.logits, self.inputs["y"], self.losses["loss"], None, None, None, self.correct)

    def kernels_per_step(self) -> int:
        return self.model.L * self.model.T * 4 + 20


def make_plan(model, store: NvlStore, batch: Optional[int] = None, **kw) -> DevicePlan:
    cls = {"encoder": EncoderPlan, "gan": GanPlan, "lstm": LstmPlan, "scaled_lstm": ScaledLstmPlan}[model.name]
    return cls(model, store, batch or model.batch_size, **kw)
