"""Per-phase device timing and NVTX ranges (the reference only has wall-clock
``time.time()`` around ``sess.run``, ``encoder:162-167``).

    with PhaseTimer() as pt:
        with pt.phase("forward"): ...
        with pt.phase("push"): ...
    pt.report()   # {'forward': ms, 'push': ms}

Phases are bracketed with CUDA events on the current stream (and an NVTX range, visible in
Nsight tools); nothing synchronises until :meth:`report`.
"""
from __future__ import annotations

import contextlib
from typing import Dict, List, Tuple

import torch


class PhaseTimer:
    def __init__(self, enabled: bool = True):
        self.enabled = enabled and torch.cuda.is_available()
        self._events: List[Tuple[str, torch.cuda.Event, torch.cuda.Event]] = []

    def __enter__(self) -> "PhaseTimer":
        return self

    def __exit__(self, *exc) -> None:
        return None

    @contextlib.contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.nvtx.range_push(name)
        s.record()
        try:
            yield
        finally:
            e.record()
            torch.cuda.nvtx.range_pop()
            self._events.append((name, s, e))

    def report(self) -> Dict[str, float]:
        if not self.enabled:
            return {}
        torch.cuda.synchronize()
        out: Dict[str, float] = {}
        for name, s, e in self._events:
            out[name] = out.get(name, 0.0) + s.elapsed_time(e)
        return out
