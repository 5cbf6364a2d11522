"""Model protocol shared by the three reference models.

A model knows (1) which variables it creates, in the reference's creation order, and
which optimizers train which of them — :meth:`Model.build`; (2) how to turn parameters and
a batch into losses and gradients with the fp32 oracle ops — :meth:`Model.step_reference`;
(3) how to do the same with the sm_100a kernels on preallocated buffers — the
``*DevicePlan`` classes next to each model.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from ..optim import Optimizer
from ..placement import VariableRegistry, VariableSpec


@dataclass
class TrainOp:
    """One ``optimizer.minimize(loss, var_list=..., global_step=global_step)``."""
    name: str
    optimizer: Optimizer
    var_list: List[str]
    loss: str                         # key into StepOutput.losses
    increments_global_step: bool = True


@dataclass
class StepOutput:
    losses: Dict[str, torch.Tensor]
    grads: Dict[str, Dict[str, torch.Tensor]]     # train-op name -> {var name -> grad}
    extras: Dict[str, torch.Tensor] = field(default_factory=dict)


class Model:
    name = "model"
    batch_size = 0
    num_steps = 0

    def __init__(self):
        self.registry: Optional[VariableRegistry] = None
        self.train_ops: List[TrainOp] = []
        self.global_step: Optional[VariableSpec] = None
        self.var: Dict[str, VariableSpec] = {}     # friendly name -> spec

    # -- graph construction -------------------------------------------------------
    def build(self, registry: VariableRegistry) -> "Model":
        raise NotImplementedError

    def _minimize(self, name: str, opt: Optimizer, var_list: List[VariableSpec], loss: str) -> None:
        opt.create_slots(self.registry, var_list)
        self.train_ops.append(TrainOp(name, opt, [v.name for v in var_list], loss))

    # -- oracle step --------------------------------------------------------------
    def step_reference(self, params: Dict[str, torch.Tensor], batch: Dict[str, torch.Tensor]) -> StepOutput:
        raise NotImplementedError

    def trainable_names(self) -> List[str]:
        out: List[str] = []
        for op in self.train_ops:
            out.extend(op.var_list)
        return out

    def steps_per_run(self) -> int:
        """How much ``global_step`` advances per training call (GAN: 2, ``gan:158-159``)."""
        return sum(1 for op in self.train_ops if op.increments_global_step)
