"""M3 — the LSTM classifier of ``lstm/distributed_lstm.py``: ``BasicLSTMCell(128,
forget_bias=1.0)`` unrolled over the 28 rows of an MNIST image by ``static_rnn``
(``:49-53``), a linear-10 head on the last output (``:53,83-88``), mean softmax
cross-entropy (``:93-94``), ``GradientDescentOptimizer(0.001)`` (``:96-97``) and an
accuracy op (``:99-100``)."""
from __future__ import annotations

from typing import Dict

import torch

from ..ops import reference as R
from ..optim import GradientDescentOptimizer
from ..placement import VariableRegistry, glorot_uniform, random_normal, zeros
from .base import Model, StepOutput

learning_rate = 0.001
training_steps = 10000
batch_size = 128
num_input = 28
timesteps = 28
num_hidden = 128
num_classes = 10
forget_bias = 1.0


class LSTMClassifier(Model):
    name = "lstm"
    batch_size = batch_size
    num_steps = training_steps

    def __init__(self, num_input=num_input, timesteps=timesteps, num_hidden=num_hidden,
                 num_classes=num_classes, lr=learning_rate):
        super().__init__()
        self.I, self.T, self.H, self.C, self.lr = num_input, timesteps, num_hidden, num_classes, lr

    def build(self, registry: VariableRegistry) -> "LSTMClassifier":
        self.registry = registry
        I, H, C = self.I, self.H, self.C
        # creation order: weights['out'], biases['out'] (lstm:83-88), then the cell's
        # kernel and bias when RNN() runs (lstm:91 -> :51-52), then global_step (:95)
        self.var["out_w"] = registry.variable(random_normal((H, C)), (H, C))
        self.var["out_b"] = registry.variable(random_normal((C,)), (C,))
        self.var["kernel"] = registry.variable(glorot_uniform((I + H, 4 * H)), (I + H, 4 * H),
                                               name="rnn/basic_lstm_cell/kernel")
        self.var["bias"] = registry.variable(zeros((4 * H,)), (4 * H,),
                                             name="rnn/basic_lstm_cell/bias")
        self.global_step = registry.global_step()
        self._minimize("train", GradientDescentOptimizer(self.lr), list(self.var.values()), "loss")
        return self

    def logits_reference(self, params, x):
        p = {k: params[s.name] for k, s in self.var.items()}
        h_last, saved = R.lstm_fwd(x, p["kernel"], p["bias"], forget_bias)
        return R.dense_fwd(h_last, p["out_w"], p["out_b"], R.ACT_NONE), h_last, saved

    def step_reference(self, params: Dict[str, torch.Tensor], batch: Dict[str, torch.Tensor]) -> StepOutput:
        p = {k: params[s.name] for k, s in self.var.items()}
        x, y = batch["x"], batch["y"]
        logits, h_last, saved = self.logits_reference(params, x)
        loss, dlogits = R.softmax_xent(logits, y)
        dh, dw, db = R.dense_bwd(h_last, p["out_w"], dlogits)
        dk, dbias = R.lstm_bwd(dh, saved, p["kernel"])
        grads = {self.var["out_w"].name: dw, self.var["out_b"].name: db,
                 self.var["kernel"].name: dk, self.var["bias"].name: dbias}
        return StepOutput({"loss": loss}, {"train": grads}, {"logits": logits})

    def accuracy_reference(self, params, x, y) -> float:
        logits, _, _ = self.logits_reference(params, x)
        return float(R.accuracy(logits, y))
