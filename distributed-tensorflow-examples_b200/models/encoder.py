"""M1 — the 784-256-128-256-784 sigmoid auto-encoder of
``encoder/distributed_encoder.py`` (variables ``:114-125``, forward ``:67-85``, MSE
``:131-135``, ``RMSPropOptimizer(0.01).minimize(loss, global_step)`` ``:136-137``)."""
from __future__ import annotations

from typing import Dict

import torch

from ..ops import reference as R
from ..optim import RMSPropOptimizer
from ..placement import VariableRegistry, random_normal
from .base import Model, StepOutput

# module-level hyper-parameters, as in the reference (encoder:45-51)
learning_rate = 0.01
num_steps = 100000
batch_size = 256
num_hidden_1 = 256
num_hidden_2 = 128
num_input = 784

LAYERS = (("encoder_h1", "encoder_b1"), ("encoder_h2", "encoder_b2"),
          ("decoder_h1", "decoder_b1"), ("decoder_h2", "decoder_b2"))


class AutoEncoder(Model):
    name = "encoder"
    batch_size = batch_size
    num_steps = num_steps

    def build(self, registry: VariableRegistry) -> "AutoEncoder":
        self.registry = registry
        shapes = {
            "encoder_h1": (num_input, num_hidden_1), "encoder_h2": (num_hidden_1, num_hidden_2),
            "decoder_h1": (num_hidden_2, num_hidden_1), "decoder_h2": (num_hidden_1, num_input),
            "encoder_b1": (num_hidden_1,), "encoder_b2": (num_hidden_2,),
            "decoder_b1": (num_hidden_1,), "decoder_b2": (num_input,),
        }
        # weights dict first, then biases dict — dict literals evaluate in source order
        for key, shape in shapes.items():
            self.var[key] = registry.variable(random_normal(shape), shape)
        self.global_step = registry.global_step()
        self._minimize("train", RMSPropOptimizer(learning_rate), list(self.var.values()), "loss")
        return self

    def step_reference(self, params: Dict[str, torch.Tensor], batch: Dict[str, torch.Tensor]) -> StepOutput:
        p = {k: params[v.name] for k, v in self.var.items()}
        x = batch["x"]
        acts = [x]
        for wk, bk in LAYERS:
            acts.append(R.dense_fwd(acts[-1], p[wk], p[bk], R.ACT_SIGMOID))
        loss, dy = R.mse_loss(acts[-1], x)
        grads: Dict[str, torch.Tensor] = {}
        for li in range(3, -1, -1):
            wk, bk = LAYERS[li]
            dz = R.act_grad(acts[li + 1], dy, R.ACT_SIGMOID)
            dy, dw, db = R.dense_bwd(acts[li], p[wk], dz, need_dx=li > 0)
            grads[self.var[wk].name] = dw
            grads[self.var[bk].name] = db
        return StepOutput({"loss": loss}, {"train": grads}, {"y_pred": acts[-1]})
