"""Variable registry, TF-style auto-naming and round-robin ps placement.

What this replaces in the reference: ``tf.Variable(...)`` created under
``tf.device(tf.train.replica_device_setter(worker_device=..., cluster=cluster))``
(``encoder/distributed_encoder.py:111-125,136``; ``gan/distributed_gan.py:119-133,144``;
``lstm/distributed_lstm.py:79-88,95``).

Rules reproduced:

* an unnamed variable is called ``Variable``, ``Variable_1``, ... in creation order;
  an explicit name that collides gets the same ``_k`` suffix treatment;
* the i-th *primary* variable created goes to ``/job:ps/task:(i mod P)``;
* optimizer slots (``<var>/RMSProp``, ``<var>/RMSProp_1``, ``<var>/Adam``, ...) and
  non-slot optimizer variables (``beta1_power`` ...) are colocated with a primary
  and do **not** advance the round-robin counter.

The registry is pure metadata + initialisers; the bytes live in a parameter store.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import torch

Initializer = Callable[[torch.Generator], torch.Tensor]


@dataclass
class VariableSpec:
    name: str
    shape: Tuple[int, ...]
    dtype: torch.dtype
    ps_task: int
    init: Initializer
    trainable: bool = True
    kind: str = "primary"          # primary | slot | nonslot | step
    primary: Optional[str] = None  # for slots/nonslots: the variable they follow
    index: int = 0                 # creation index among *all* registry entries

    @property
    def numel(self) -> int:
        n = 1
        for d in self.shape:
            n *= d
        return n

    @property
    def device(self) -> str:
        return "/job:ps/task:%d" % self.ps_task


class RoundRobin:
    """``_RoundRobinStrategy(num_ps)``: hands out 0,1,..,P-1,0,1,.. per call."""

    def __init__(self, num_tasks: int):
        if num_tasks < 1:
            raise ValueError("need at least one ps task")
        self.num_tasks = num_tasks
        self._next = 0

    def __call__(self) -> int:
        t = self._next
        self._next = (self._next + 1) % self.num_tasks
        return t


class ReplicaDeviceSetter:
    """Device function: Variable ops -> ps round-robin, everything else -> worker."""

    def __init__(self, ps_tasks: int, worker_device: str = "/job:worker/task:0"):
        self.ps_tasks = ps_tasks
        self.worker_device = worker_device
        self.strategy = RoundRobin(ps_tasks)

    def place_variable(self) -> int:
        return self.strategy()

    def place_op(self) -> str:
        return self.worker_device


def replica_device_setter(worker_device: str = "/job:worker/task:0", cluster=None,
                          ps_tasks: int = 0) -> ReplicaDeviceSetter:
    """Same call shape as ``tf.train.replica_device_setter``; ``ps_tasks`` is taken
    from ``cluster`` when given (``encoder/distributed_encoder.py:111-113``)."""
    if cluster is not None:
        ps_tasks = cluster.num_tasks("ps")
    if ps_tasks < 1:
        raise ValueError("replica_device_setter needs a cluster with ps tasks or ps_tasks>=1")
    return ReplicaDeviceSetter(ps_tasks, worker_device)


class VariableRegistry:
    """Ordered collection of variable specs == ``tf.global_variables()`` of a worker graph."""

    def __init__(self, setter: ReplicaDeviceSetter):
        self.setter = setter
        self._specs: Dict[str, VariableSpec] = {}
        self._order: List[str] = []
        self._name_counts: Dict[str, int] = {}

    # ---- naming ------------------------------------------------------------------
    def _unique(self, base: str) -> str:
        n = self._name_counts.get(base, 0)
        self._name_counts[base] = n + 1
        if n == 0:
            return base
        name = "%s_%d" % (base, n)
        while name in self._name_counts:      # e.g. someone explicitly named "Variable_1"
            n += 1
            name = "%s_%d" % (base, n)
        self._name_counts[name] = 1           # a produced name is itself reserved
        return name

    def _add(self, spec: VariableSpec) -> VariableSpec:
        spec.index = len(self._order)
        self._specs[spec.name] = spec
        self._order.append(spec.name)
        return spec

    # ---- creation ----------------------------------------------------------------
    def variable(self, init: Initializer, shape: Sequence[int], dtype=torch.float32,
                 name: Optional[str] = None, trainable: bool = True,
                 kind: str = "primary") -> VariableSpec:
        """``tf.Variable(initial_value)`` under the device setter."""
        return self._add(VariableSpec(
            name=self._unique(name or "Variable"), shape=tuple(int(s) for s in shape),
            dtype=dtype, ps_task=self.setter.place_variable(), init=init,
            trainable=trainable, kind=kind))

    def global_step(self) -> VariableSpec:
        """``tf.Variable(0)`` — int32 scalar, unnamed, non-trainable in effect."""
        return self.variable(lambda g: torch.zeros((), dtype=torch.int32), (), torch.int32,
                             trainable=False, kind="step")

    def slot(self, primary: VariableSpec, op_name: str, init: Initializer) -> VariableSpec:
        """Slot variable ``<primary>/<op_name>[_k]`` colocated with ``primary``."""
        return self._add(VariableSpec(
            name=self._unique("%s/%s" % (primary.name, op_name)), shape=primary.shape,
            dtype=primary.dtype, ps_task=primary.ps_task, init=init, trainable=False,
            kind="slot", primary=primary.name))

    def nonslot(self, name: str, colocate_with: VariableSpec, init: Initializer,
                shape: Sequence[int] = (), dtype=torch.float32) -> VariableSpec:
        """Non-slot optimizer variable (Adam's ``beta1_power``) colocated with a primary."""
        return self._add(VariableSpec(
            name=self._unique(name), shape=tuple(shape), dtype=dtype,
            ps_task=colocate_with.ps_task, init=init, trainable=False, kind="nonslot",
            primary=colocate_with.name))

    # ---- views -------------------------------------------------------------------
    def __getitem__(self, name: str) -> VariableSpec:
        return self._specs[name]

    def __contains__(self, name: str) -> bool:
        return name in self._specs

    def __iter__(self) -> Iterable[VariableSpec]:
        return (self._specs[n] for n in self._order)

    def __len__(self) -> int:
        return len(self._order)

    def names(self) -> List[str]:
        return list(self._order)

    def on_task(self, ps_task: int) -> List[VariableSpec]:
        return [s for s in self if s.ps_task == ps_task]

    def trainable(self) -> List[VariableSpec]:
        return [s for s in self if s.trainable and s.kind == "primary"]

    def initial_values(self, seed: int = 0) -> Dict[str, torch.Tensor]:
        """Run every initialiser once (what ``init_op`` does on the chief)."""
        g = torch.Generator(device="cpu")
        g.manual_seed(seed)
        out = {}
        for s in self:
            v = s.init(g)
            if tuple(v.shape) != s.shape or v.dtype != s.dtype:
                raise ValueError("initialiser of %s produced %s %s, expected %s %s"
                                 % (s.name, tuple(v.shape), v.dtype, s.shape, s.dtype))
            out[s.name] = v.contiguous()
        return out

    def placement_table(self) -> List[Tuple[str, Tuple[int, ...], str]]:
        return [(s.name, s.shape, s.device) for s in self]


# ---- initialisers (K10) ------------------------------------------------------------
def random_normal(shape: Sequence[int], stddev: float = 1.0, mean: float = 0.0) -> Initializer:
    """``tf.random_normal(shape)`` (``encoder/distributed_encoder.py:115-124``)."""
    shape = tuple(shape)
    return lambda g: torch.randn(shape, generator=g, dtype=torch.float32) * stddev + mean


def zeros(shape: Sequence[int], dtype=torch.float32) -> Initializer:
    shape = tuple(shape)
    return lambda g: torch.zeros(shape, dtype=dtype)


def ones(shape: Sequence[int], dtype=torch.float32) -> Initializer:
    shape = tuple(shape)
    return lambda g: torch.ones(shape, dtype=dtype)


def constant(value: float, shape: Sequence[int] = (), dtype=torch.float32) -> Initializer:
    shape = tuple(shape)
    return lambda g: torch.full(shape, value, dtype=dtype)


def glorot_normal_ref(shape: Sequence[int]) -> Initializer:
    """The GAN script's ``glorot_init``: N(0, 1/sqrt(shape[0]/2)) (``gan/distributed_gan.py:72-73``)."""
    shape = tuple(shape)
    std = 1.0 / (shape[0] / 2.0) ** 0.5
    return random_normal(shape, stddev=std)


def glorot_uniform(shape: Sequence[int]) -> Initializer:
    """Variable-scope default initialiser used for ``rnn/basic_lstm_cell/kernel``:
    U(-l, l) with l = sqrt(6/(fan_in+fan_out))."""
    shape = tuple(shape)
    limit = (6.0 / (shape[0] + shape[1])) ** 0.5
    return lambda g: (torch.rand(shape, generator=g, dtype=torch.float32) * 2 - 1) * limit
