"""T1 (in-process): Supervisor semantics on the tcp transport — chief init vs non-chief wait,
restore instead of init, reinit_on_join quirk, ps done-queue timeout, GAN alternating updates."""
import threading
import time

import pytest
import torch

from psb200.checkpoint import Saver, latest_checkpoint
from psb200.cluster import ClusterSpec, Server
from psb200.flags import parse_flags
from psb200.models import make_model
from psb200.placement import VariableRegistry, replica_device_setter
from psb200.runner import _train_call_cpu, run_ps
from psb200.store.tcp import TcpStore
from psb200.supervisor import Supervisor


@pytest.fixture
def cluster():
    ps = Server(ClusterSpec({"ps": ["127.0.0.1:0"], "worker": ["x:0"]}), "ps", 0)
    spec = ClusterSpec({"ps": ["127.0.0.1:%d" % ps.port], "worker": ["127.0.0.1:1", "127.0.0.1:2"]})
    yield spec
    ps.stop()


def _worker(spec, name, task, tmp, **kw):
    reg = VariableRegistry(replica_device_setter(worker_device="/job:worker/task:%d" % task, cluster=spec))
    model = make_model(name).build(reg)
    store = TcpStore(spec, reg, model, num_workers=2)
    sv = Supervisor(is_chief=(task == 0), logdir=str(tmp), store=store, registry=reg, model=model,
                    save_model_secs=0, recovery_wait_secs=0.05, seed=1, **kw)
    return reg, model, store, sv


def test_non_chief_waits_for_chief_init(cluster, tmp_path):
    _, _, store1, sv1 = _worker(cluster, "lstm", 1, tmp_path)
    ready = threading.Event()
    t = threading.Thread(target=lambda: (sv1.prepare_or_wait_for_session(), ready.set()), daemon=True)
    t.start()
    time.sleep(0.3)
    assert not ready.is_set()                     # variables declared but uninitialised: keeps polling
    _, _, store0, sv0 = _worker(cluster, "lstm", 0, tmp_path)
    sv0.prepare_or_wait_for_session()
    assert ready.wait(5.0)
    assert store1.uninitialized() == 0 and store1.global_step() == 0
    sv0.stop(); sv1.stop()


def test_chief_restores_latest_checkpoint_instead_of_init(cluster, tmp_path):
    reg, model, store, sv = _worker(cluster, "lstm", 0, tmp_path)
    vals = reg.initial_values(7)
    vals["Variable_2"] = torch.tensor(123, dtype=torch.int32)
    Saver().save(vals, str(tmp_path), 123)
    sv.prepare_or_wait_for_session()
    assert sv.restored_from == latest_checkpoint(str(tmp_path))
    assert store.global_step() == 123 and torch.equal(store.read("Variable"), vals["Variable"])
    sv.stop()


def test_reinit_on_join_reproduces_the_reference_quirk(cluster, tmp_path):
    reg, model, store, sv = _worker(cluster, "encoder", 0, tmp_path)
    sv.prepare_or_wait_for_session()
    store._of(model.global_step).var_add_i32(model.global_step.name, 5)
    assert store.global_step() == 5
    # a worker that joins with --reinit_on_join re-runs init: global_step goes back to zero (encoder:160)
    _, _, store1, sv1 = _worker(cluster, "encoder", 1, tmp_path, reinit_on_join=True)
    sv1.prepare_or_wait_for_session()
    assert store.global_step() == 0
    sv.stop(); sv1.stop()


def test_ps_wait_timeout(capsys):
    spec = ClusterSpec({"ps": ["127.0.0.1:0"], "worker": ["x:0"]})
    server = Server(spec, "ps", 0)
    flags = parse_flags(["--workers=2", "--ps_wait_timeout=0.3"])
    t0 = time.time()
    run_ps(flags, spec, server)
    server.stop()
    out = capsys.readouterr().out
    assert "timed out waiting for done 0" in out and "ps 0: quitting" in out and time.time() - t0 < 5


def test_gan_alternating_updates_advance_step_twice(cluster, tmp_path):
    reg, model, store, sv = _worker(cluster, "gan", 0, tmp_path)
    sv.prepare_or_wait_for_session()
    flags = parse_flags(["--gan_update=alternating"])
    batch = {"x": torch.rand(8, 784), "z": torch.rand(8, 100) * 2 - 1}
    before = store.read("Variable_2").clone(), store.read("Variable").clone()
    losses, step = _train_call_cpu(model, store, flags, batch, 0, sv)
    assert step == 2 and set(losses) == {"gen_loss", "disc_loss"}
    assert not torch.equal(store.read("Variable_2"), before[0]) and not torch.equal(store.read("Variable"), before[1])
    # Adam's beta powers advanced once per optimizer
    assert abs(float(store.read("beta1_power")) - 0.81) < 1e-6 and abs(float(store.read("beta1_power_1")) - 0.81) < 1e-6
    sv.stop()
