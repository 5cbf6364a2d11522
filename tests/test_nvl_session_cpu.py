"""The Supervisor life cycle of the NVLink transport (`runner.run_worker_nvl` = `Supervisor` +
`parallel/session.py`) with two gloo processes and a stand-in for the device trainer: chief init,
non-chief wait joining the collective initialisation, checkpoint timer thread, `graph.json`,
chief restart -> both ranks resume from the checkpoint (``encoder:147-161,176``)."""
import contextlib
import os
import socket
import time

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


class _Stream:
    def synchronize(self):
        pass


class _Store:
    device = torch.device("cpu")

    def __init__(self):
        self.values = {}

    def snapshot_async(self, stream):
        return {k: v.clone() for k, v in self.values.items()}


class _Trainer:
    """What NvlSession needs of a DeviceTrainer."""

    def __init__(self, world, gs_name):
        self.world, self.store, self.gs_name, self.inits = world, _Store(), gs_name, 0

    def initialize(self, values):
        assert values is not None, "initialisation values must have been broadcast by the chief"
        self.inits += 1
        self.store.values = {k: v.clone() for k, v in values.items()}

    def global_step(self):
        return int(self.store.values[self.gs_name])


def _worker(rank, port, logdir, phase, out):
    torch.cuda.device = lambda d: contextlib.nullcontext()
    torch.cuda.Stream = lambda device=None: _Stream()
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=2)
    from psb200.checkpoint import Saver
    from psb200.models import make_model
    from psb200.parallel.session import NvlSession
    from psb200.placement import VariableRegistry, replica_device_setter
    from psb200.supervisor import Supervisor
    registry = VariableRegistry(replica_device_setter(worker_device="/job:worker/task:%d" % rank, ps_tasks=2))
    model = make_model("lstm").build(registry)
    gs = model.global_step.name
    tr = _Trainer(2, gs)
    sess = NvlSession(tr, dist.group.WORLD, is_chief=(rank == 0))
    sv = Supervisor(is_chief=(rank == 0), logdir=logdir, store=sess, registry=registry, model=model, saver=Saver(),
                    save_model_secs=0.2, recovery_wait_secs=0.05, seed=3, sync=False)
    with sv.managed_session():
        assert tr.inits == 1 and set(tr.store.values) == {s.name for s in registry}
        start = sess.global_step()
        digest = float(sum(v.double().sum() for v in tr.store.values.values()))
        both = [None, None]
        dist.all_gather_object(both, (start, digest))
        assert both[0] == both[1], both                   # every rank initialised from the chief's values
        for i in range(5):                                # "training": the step advances, the timer thread snapshots
            tr.store.values[gs] = tr.store.values[gs] + 1
            sess.note_step(int(tr.store.values[gs]))
            time.sleep(0.1)
        assert not sv.should_stop()
        dist.barrier()
    sv.stop()
    if rank == 0:
        torch.save({"start": start, "restored": sv.restored_from, "last_saved": sv.last_saved,
                    "files": sorted(os.listdir(logdir))}, out)
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_supervisor_drives_the_nvlink_session_across_two_ranks(tmp_path):
    logdir = str(tmp_path / "model")
    r1, r2 = str(tmp_path / "r1.pt"), str(tmp_path / "r2.pt")
    mp.spawn(_worker, args=(_free_port(), logdir, 1, r1), nprocs=2, join=True)
    a = torch.load(r1)
    assert a["start"] == 0 and a["restored"] is None
    assert "graph.json" in a["files"] and "checkpoint" in a["files"]
    assert any(f.startswith("model.ckpt-5.") for f in a["files"]), a["files"]      # final checkpoint at step 5
    # the chief comes back: both ranks resume from the newest checkpoint instead of initialising
    mp.spawn(_worker, args=(_free_port(), logdir, 2, r2), nprocs=2, join=True)
    b = torch.load(r2)
    assert b["start"] == 5 and b["restored"] is not None and b["restored"].endswith("model.ckpt-5")
    assert any(f.startswith("model.ckpt-10.") for f in b["files"]), b["files"]
