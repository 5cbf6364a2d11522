"""The NVLink transport seen through the Supervisor's eyes.

``Supervisor`` (``supervisor.py``) drives any store that can declare / initialise variables,
report readiness, read everything back and tell the global step.  On the tcp transport that is
``TcpStore``; here it is a thin adapter over :class:`~psb200.parallel.trainer.DeviceTrainer`, so
``run_worker_nvl`` goes through the same ``managed_session`` / ``should_stop`` / checkpoint-timer
/ ``graph.json`` / ``events.jsonl`` machinery as the reference's
``Supervisor(...).managed_session(...)`` (``encoder:147-161,176``):

* chief init or restore, then "non-chief waits": initialisation is one collective -- the chief
  broadcasts the values it restored or drew, every rank writes its shards and the first parameter
  pull happens (``DeviceTrainer.initialize``).  A non-chief's ``uninitialized()`` poll joins it.
* the checkpoint timer thread reads the variables with copies on its own stream into pinned host
  memory (``NvlStore.snapshot_async``): the training stream is never synchronised or blocked.
* ``global_step()`` for the step-counter thread is the value the last finished step returned --
  a host integer, no device access from a foreign thread.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch


class NvlSession:
    def __init__(self, trainer, group=None, is_chief: bool = False):
        self.tr, self.group, self.is_chief = trainer, group, is_chief
        self.num_workers = trainer.world
        self._initialized = False
        self._last_step = 0
        self._snap_stream: Optional[torch.cuda.Stream] = None
        self._device = trainer.store.device

    # ---- what Supervisor.prepare_or_wait_for_session calls -----------------------------------
    def declare_all(self) -> None:            # storage was laid out when the trainer was built
        pass

    def sync_setup(self, is_chief: bool) -> None:   # aggregation happens on the device
        pass

    def _collective_init(self, values) -> None:
        if self.tr.world > 1:
            import torch.distributed as dist
            box = [values]
            dist.broadcast_object_list(box, src=0, group=self.group)
            values = box[0]
        self.tr.initialize(values)
        self._initialized = True
        self._last_step = self.tr.global_step()

    def initialize(self, values: Dict[str, torch.Tensor]) -> None:
        """Chief: ``init_op`` / restored checkpoint -> every shard (collective)."""
        self._collective_init(values)

    def uninitialized(self) -> int:
        """Non-chief ``ready_op``: joins the chief's initialisation, then reports 0."""
        if not self._initialized:
            self._collective_init(None)
        return 0

    # ---- services -------------------------------------------------------------------------------
    def note_step(self, global_step: int) -> None:
        self._last_step = int(global_step)

    def global_step(self) -> int:
        return self._last_step

    def read_all(self) -> Dict[str, torch.Tensor]:
        """Snapshot for the checkpoint thread: side-stream copies into pinned memory, then a wait on
        that stream only."""
        with torch.cuda.device(self._device):
            if self._snap_stream is None:
                self._snap_stream = torch.cuda.Stream(device=self._device)
            snap = self.tr.store.snapshot_async(self._snap_stream)
            self._snap_stream.synchronize()
        return {k: v.clone() for k, v in snap.items()}
