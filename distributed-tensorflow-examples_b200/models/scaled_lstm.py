"""Scaled stacked LSTM classifier — BASELINE.json config #4: "scaled 2-layer LSTM hidden 4096
seq 512 bf16 (comm-bound shard push/pull)".

Same family as ``lstm/distributed_lstm.py`` (BasicLSTMCell, zero initial state, linear head on
the last output, softmax cross-entropy, SGD) with ``MultiRNNCell``-style stacking; variable
names follow TF's scope naming (``rnn/multi_rnn_cell/cell_<l>/basic_lstm_cell/{kernel,bias}``).
With I = H = 4096 and two layers the cell kernels hold 2 x (8192 x 16384) = 268 M parameters.
"""
from __future__ import annotations

from typing import Dict, List

import torch

from ..ops import reference as R
from ..optim import GradientDescentOptimizer
from ..placement import VariableRegistry, glorot_uniform, random_normal, zeros
from .base import Model, StepOutput

forget_bias = 1.0


class ScaledLSTM(Model):
    name = "scaled_lstm"
    batch_size = 64
    num_steps = 1000

    def __init__(self, num_input=4096, timesteps=512, num_hidden=4096, num_layers=2, num_classes=10, lr=0.001):
        super().__init__()
        self.I, self.T, self.H, self.L, self.C, self.lr = num_input, timesteps, num_hidden, num_layers, num_classes, lr

    def build(self, registry: VariableRegistry) -> "ScaledLSTM":
        self.registry = registry
        H, C = self.H, self.C
        self.var["out_w"] = registry.variable(random_normal((H, C), stddev=0.05), (H, C))
        self.var["out_b"] = registry.variable(zeros((C,)), (C,))
        for l in range(self.L):
            I = self.I if l == 0 else H
            scope = "rnn/multi_rnn_cell/cell_%d/basic_lstm_cell" % l
            self.var["kernel%d" % l] = registry.variable(glorot_uniform((I + H, 4 * H)), (I + H, 4 * H),
                                                         name=scope + "/kernel")
            self.var["bias%d" % l] = registry.variable(zeros((4 * H,)), (4 * H,), name=scope + "/bias")
        self.global_step = registry.global_step()
        self._minimize("train", GradientDescentOptimizer(self.lr), list(self.var.values()), "loss")
        return self

    def in_dim(self, layer: int) -> int:
        return self.I if layer == 0 else self.H

    def step_reference(self, params: Dict[str, torch.Tensor], batch: Dict[str, torch.Tensor]) -> StepOutput:
        """``batch['x']`` is time-major ``[T, B, I]``."""
        p = {k: params[s.name] for k, s in self.var.items()}
        x = batch["x"].transpose(0, 1)          # oracle ops are batch-major
        y = batch["y"]
        saves, inp = [], x
        for l in range(self.L):
            _, sv = R.lstm_fwd(inp, p["kernel%d" % l], p["bias%d" % l], forget_bias)
            saves.append(sv)
            inp = sv.h[1:].transpose(0, 1)       # [B, T, H]
        h_last = saves[-1].h[-1]
        logits = R.dense_fwd(h_last, p["out_w"], p["out_b"], R.ACT_NONE)
        loss, dlogits = R.softmax_xent(logits, y)
        dh, dw, db = R.dense_bwd(h_last, p["out_w"], dlogits)
        grads = {self.var["out_w"].name: dw, self.var["out_b"].name: db}
        dh_seq = None
        for l in range(self.L - 1, -1, -1):
            res = R.lstm_bwd(dh if l == self.L - 1 else None, saves[l], p["kernel%d" % l], dh_seq=dh_seq,
                             need_dx=l > 0)
            grads[self.var["kernel%d" % l].name], grads[self.var["bias%d" % l].name] = res[0], res[1]
            if l > 0:
                dh_seq = res[2].transpose(0, 1)  # [T, B, H] for the layer below
        return StepOutput({"loss": loss}, {"train": grads}, {"logits": logits})
