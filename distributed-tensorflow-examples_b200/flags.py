"""Command-line surface.

The eight reference flags keep their names, types and defaults
(``encoder/distributed_encoder.py:27-40``, ``gan/distributed_gan.py:32-45``,
``lstm/distributed_lstm.py:9-22``).  Everything after them is an opt-in extension
whose default reproduces the reference behaviour (async updates, MNIST from
``data_dir``, checkpoint every 60 s).

Like the reference, ``--ps`` is accepted and never consulted: the number of ps tasks
is ``len(ps_hosts.split(","))`` (``encoder/distributed_encoder.py:63``), while
``--workers`` sizes the done-queues and the ps wait loop (``:59,105``).
"""
from __future__ import annotations

import argparse
from dataclasses import dataclass, field, fields
from typing import List, Optional, Sequence


def _bool(v: str) -> bool:
    if isinstance(v, bool):
        return v
    s = str(v).strip().lower()
    if s in ("1", "true", "t", "yes", "y", "on"):
        return True
    if s in ("0", "false", "f", "no", "n", "off", ""):
        return False
    raise argparse.ArgumentTypeError("expected a boolean, got %r" % v)


@dataclass
class Flags:
    # ---- reference flags (names/defaults are public surface) -------------------
    ps_hosts: str = ""
    worker_hosts: str = ""
    job_name: str = ""
    task_index: int = 0
    data_dir: str = "/data_dir"
    model_dir: str = "/tmp/checkpoints"
    workers: int = 3
    ps: int = 1
    # ---- extensions (defaults = reference behaviour) ---------------------------
    sync: bool = False                 # SyncReplicasOptimizer-style aggregation
    replicas_to_aggregate: int = 0     # 0 -> number of workers
    backend: str = "cpu"               # "cpu" (TCP ps, torch-CPU math) | "nvl" (B200 peer memory)
    synthetic: bool = False            # MNIST-shaped synthetic data instead of data_dir
    num_steps: int = 0                 # 0 -> the script's module constant
    save_model_secs: int = 60          # Supervisor checkpoint period
    reinit_on_join: bool = False       # reproduce `sess.run(init)` on every worker (encoder:160, gan:181)
    gan_update: str = "simultaneous"   # "simultaneous" (gan:193) | "alternating"
    dtype: str = "fp32"                # cpu backend: weight precision seen by the worker math, fp32 | bf16 (nvl: always bf16 operands, fp32 accumulate)
    seed: int = 0
    log_every: int = 1                 # print the per-step line every n local steps
    ps_wait_timeout: float = 0.0       # 0 -> ps waits forever for done tokens (reference)
    recovery_wait_secs: float = 1.0    # non-chief polling period (TF default is 30 s)
    batch_size: int = 0                # 0 -> the script's module constant
    checkpoint_format: str = "psb200"  # bundle written by the chief: "psb200" (JSON index) | "tf" (TF V2 bundle)
    async_apply: str = "worker"        # nvl async: "worker" (remote read-modify-write on the owner's HBM) | "owner" (ps service thread applies)
    peer_timeout_secs: float = 30.0    # nvl: bound on every cross-GPU wait inside the push kernel
    extra: List[str] = field(default_factory=list)

    def ps_list(self) -> List[str]:
        return self.ps_hosts.split(",")

    def worker_list(self) -> List[str]:
        return self.worker_hosts.split(",")


_HELP = {
    "ps_hosts": "Comma-separated list of hostname:port pairs",
    "worker_hosts": "Comma-separated list of hostname:port pairs",
    "job_name": "One of 'ps', 'worker'",
    "task_index": "Index of task within the job",
    "data_dir": "Directory for storing mnist data",
    "model_dir": "Directory for storing the checkpoints",
    "workers": "Number of workers",
    "ps": "Number of ps",
}


def build_parser(prog: Optional[str] = None) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog=prog, allow_abbrev=False)
    for f in fields(Flags):
        if f.name == "extra":
            continue
        default = f.default
        kind = type(default)
        kw = dict(default=default, help=_HELP.get(f.name, "extension flag"))
        if kind is bool:
            # absl style: --sync, --sync=true, --nosync
            p.add_argument("--" + f.name, nargs="?", const=True, type=_bool, **kw)
            p.add_argument("--no" + f.name, dest=f.name, action="store_false",
                           help=argparse.SUPPRESS)
        else:
            p.add_argument("--" + f.name, type=kind, **kw)
    return p


def parse_flags(argv: Optional[Sequence[str]] = None, prog: Optional[str] = None) -> Flags:
    """Parse ``argv`` (default ``sys.argv[1:]``); unknown arguments are kept in
    ``Flags.extra`` the way ``tf.app.run`` hands leftovers to ``main``."""
    ns, extra = build_parser(prog).parse_known_args(argv)
    fl = Flags(**{k: v for k, v in vars(ns).items()})
    fl.extra = list(extra)
    if fl.backend not in ("cpu", "nvl"):
        raise SystemExit("--backend must be 'cpu' or 'nvl'")
    if fl.dtype not in ("fp32", "bf16"):
        raise SystemExit("--dtype must be 'fp32' or 'bf16'")
    if fl.checkpoint_format not in ("psb200", "tf"):
        raise SystemExit("--checkpoint_format must be 'psb200' or 'tf'")
    if fl.async_apply not in ("worker", "owner"):
        raise SystemExit("--async_apply must be 'worker' or 'owner'")
    if fl.gan_update not in ("simultaneous", "alternating"):
        raise SystemExit("--gan_update must be 'simultaneous' or 'alternating'")
    return fl
