"""Optimizer specifications (slot schema + update math).

Three optimizers appear in the reference: ``RMSPropOptimizer(0.01)``
(``encoder/distributed_encoder.py:137``), two ``AdamOptimizer(2e-4)``
(``gan/distributed_gan.py:146-147``) and ``GradientDescentOptimizer(0.001)``
(``lstm/distributed_lstm.py:96``).  Hyper-parameter defaults are TF 1.11's.

An optimizer here is *data*: a kind id, hyper-parameters, and the names of the slot /
non-slot variables it owns in the registry.  The arithmetic runs where the variable
lives — in the native ps server (``csrc/ps_server.cpp``), in the fused push kernel
(``csrc/push.cu``), or in :func:`apply_reference` (the fp32 torch oracle used by
tests and by nothing on a hot path).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch

from ..placement import VariableRegistry, VariableSpec, constant, ones, zeros

KIND_SGD = 0
KIND_RMSPROP = 1
KIND_ADAM = 2


@dataclass
class Optimizer:
    kind: int
    name: str                      # TF op name: "GradientDescent" | "RMSProp" | "Adam"
    lr: float
    hyper: Dict[str, float] = field(default_factory=dict)
    # filled by create_slots
    slots: Dict[str, List[str]] = field(default_factory=dict)     # var -> [slot names]
    nonslots: List[str] = field(default_factory=list)             # e.g. [beta1_power, beta2_power]
    var_list: List[str] = field(default_factory=list)

    # ---- graph construction ------------------------------------------------------
    def create_slots(self, registry: VariableRegistry, var_list: Sequence[VariableSpec]) -> None:
        """What ``Optimizer.minimize`` does to the variable set (names per SURVEY §2.4)."""
        self.var_list = [v.name for v in var_list]
        if self.kind == KIND_ADAM:
            # TF: non-slot variables are colocated with min(var_list, key=name) and are
            # created before the per-variable slots.
            first = min(var_list, key=lambda v: v.name + ":0")
            b1 = registry.nonslot("beta1_power", first, constant(self.hyper["beta1"]))
            b2 = registry.nonslot("beta2_power", first, constant(self.hyper["beta2"]))
            self.nonslots = [b1.name, b2.name]
            for v in var_list:
                m = registry.slot(v, self.name, zeros(v.shape))
                s = registry.slot(v, self.name, zeros(v.shape))
                self.slots[v.name] = [m.name, s.name]
        elif self.kind == KIND_RMSPROP:
            for v in var_list:
                rms = registry.slot(v, self.name, ones(v.shape))      # rms slot starts at ONE
                mom = registry.slot(v, self.name, zeros(v.shape))
                self.slots[v.name] = [rms.name, mom.name]
        else:
            for v in var_list:
                self.slots[v.name] = []

    # packed hyper-parameters in the order the native code expects
    def hyper_vector(self) -> List[float]:
        if self.kind == KIND_SGD:
            return [self.lr, 0.0, 0.0, 0.0]
        if self.kind == KIND_RMSPROP:
            return [self.lr, self.hyper["decay"], self.hyper["momentum"], self.hyper["epsilon"]]
        return [self.lr, self.hyper["beta1"], self.hyper["beta2"], self.hyper["epsilon"]]


def GradientDescentOptimizer(learning_rate: float) -> Optimizer:
    return Optimizer(KIND_SGD, "GradientDescent", float(learning_rate))


def RMSPropOptimizer(learning_rate: float, decay: float = 0.9, momentum: float = 0.0,
                     epsilon: float = 1e-10) -> Optimizer:
    return Optimizer(KIND_RMSPROP, "RMSProp", float(learning_rate),
                     dict(decay=decay, momentum=momentum, epsilon=epsilon))


def AdamOptimizer(learning_rate: float = 0.001, beta1: float = 0.9, beta2: float = 0.999,
                  epsilon: float = 1e-8) -> Optimizer:
    return Optimizer(KIND_ADAM, "Adam", float(learning_rate),
                     dict(beta1=beta1, beta2=beta2, epsilon=epsilon))


# ---- fp32 oracle -----------------------------------------------------------------------
def apply_reference(opt: Optimizer, var: torch.Tensor, grad: torch.Tensor,
                    slots: Sequence[torch.Tensor],
                    beta_powers: Optional[Sequence[float]] = None) -> None:
    """In-place dense update, TF 1.11 kernel arithmetic.

    * SGD:      ``w -= lr*g``
    * RMSProp:  ``ms = rho*ms + (1-rho)*g*g; mom = mu*mom + lr*g/sqrt(ms+eps); w -= mom``
    * Adam:     ``lr_t = lr*sqrt(1-b2p)/(1-b1p); m += (1-b1)(g-m); v += (1-b2)(g*g-v);
                 w -= lr_t*m/(sqrt(v)+eps)``   (``b1p``/``b2p`` are the *current* powers)
    """
    g = grad.to(torch.float32)
    if opt.kind == KIND_SGD:
        var.sub_(opt.lr * g)
    elif opt.kind == KIND_RMSPROP:
        ms, mom = slots
        rho, mu, eps = opt.hyper["decay"], opt.hyper["momentum"], opt.hyper["epsilon"]
        ms.mul_(rho).add_((1.0 - rho) * g * g)
        mom.mul_(mu).add_(opt.lr * g / torch.sqrt(ms + eps))
        var.sub_(mom)
    elif opt.kind == KIND_ADAM:
        m, v = slots
        b1, b2, eps = opt.hyper["beta1"], opt.hyper["beta2"], opt.hyper["epsilon"]
        b1p, b2p = beta_powers
        lr_t = opt.lr * (1.0 - b2p) ** 0.5 / (1.0 - b1p)
        m.add_((1.0 - b1) * (g - m))
        v.add_((1.0 - b2) * (g * g - v))
        var.sub_(lr_t * m / (torch.sqrt(v) + eps))
    else:
        raise ValueError("unknown optimizer kind %r" % opt.kind)


def adam_lr_t(opt: Optimizer, b1p: float, b2p: float) -> float:
    return opt.lr * (1.0 - b2p) ** 0.5 / (1.0 - b1p)
