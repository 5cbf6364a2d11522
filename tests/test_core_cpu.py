"""T0: flags parity, ClusterSpec, checkpoint round-trip / naming / keep-5, data pipeline, the
native ps runtime's variable, queue and accumulator semantics."""
import os
import threading
import time

import numpy as np
import pytest
import torch

from psb200.checkpoint import Saver, latest_checkpoint, read_state
from psb200.cluster import ClusterSpec, Server, device_filters
from psb200.control.client import PsClient, PsError, QueueTimeout
from psb200.data import DataSet, read_data_sets, synthetic_mnist
from psb200.flags import Flags, parse_flags
from psb200.models import GAN, AutoEncoder, LSTMClassifier
from psb200.placement import VariableRegistry, replica_device_setter


def test_reference_flag_names_and_defaults():
    f = parse_flags([])
    assert (f.ps_hosts, f.worker_hosts, f.job_name, f.task_index) == ("", "", "", 0)
    assert (f.data_dir, f.model_dir, f.workers, f.ps) == ("/data_dir", "/tmp/checkpoints", 3, 1)
    assert f.sync is False and f.backend == "cpu" and f.save_model_secs == 60
    f = parse_flags(["--ps_hosts=a:1,b:2", "--worker_hosts=c:3", "--job_name=worker", "--task_index=1",
                     "--workers=2", "--ps=2", "--sync", "--nosynthetic", "--unknown=1"])
    assert f.ps_list() == ["a:1", "b:2"] and f.worker_list() == ["c:3"]
    assert f.job_name == "worker" and f.task_index == 1 and f.workers == 2 and f.sync is True
    assert f.extra == ["--unknown=1"]
    assert parse_flags(["--sync=false"]).sync is False


def test_cluster_spec_and_device_filters():
    c = ClusterSpec({"ps": ["h1:2222", "h2:2222"], "worker": {1: "w1:1", 0: "w0:1"}})
    assert c.num_tasks("ps") == 2 and c.task_address("worker", 0) == "w0:1" and c.jobs == ["ps", "worker"]
    assert c.as_dict()["worker"] == ["w0:1", "w1:1"]
    assert device_filters(3) == ["/job:ps", "/job:worker/task:3"]
    with pytest.raises(ValueError):
        Server(c, "chief", 0, start=False)


@pytest.mark.parametrize("cls,count", [(AutoEncoder, 25), (GAN, 29), (LSTMClassifier, 5)])
def test_checkpoint_roundtrip_and_key_set(tmp_path, cls, count):
    reg = VariableRegistry(replica_device_setter(ps_tasks=2))
    cls().build(reg)
    vals = reg.initial_values(1)
    saver = Saver()
    prefix = saver.save(vals, str(tmp_path), 7)
    assert os.path.basename(prefix) == "model.ckpt-7"
    for ext in (".index", ".data-00000-of-00001", ".meta"):
        assert os.path.exists(prefix + ext)
    assert open(os.path.join(tmp_path, "checkpoint")).readline().strip() == 'model_checkpoint_path: "model.ckpt-7"'
    assert latest_checkpoint(str(tmp_path)) == prefix
    back = Saver.restore(prefix)
    assert set(back) == set(reg.names()) and len(back) == count
    for k in vals:
        assert back[k].dtype == vals[k].dtype and torch.equal(back[k], vals[k])
    shapes = Saver.list_variables(prefix)
    assert all(tuple(shapes[s.name]) == s.shape for s in reg)


def test_checkpoint_keeps_five(tmp_path):
    saver = Saver()
    v = {"Variable": torch.zeros(3), "Variable_1": torch.zeros((), dtype=torch.int32)}
    for step in range(8):
        saver.save(v, str(tmp_path), step)
    kept = read_state(str(tmp_path))
    assert kept == ["model.ckpt-%d" % s for s in range(3, 8)]
    assert not os.path.exists(os.path.join(tmp_path, "model.ckpt-2.index"))
    assert latest_checkpoint(str(tmp_path)).endswith("model.ckpt-7")


def test_synthetic_and_idx_data(tmp_path):
    d = synthetic_mnist(n_train=256, n_test=64)
    x, y = d.train.next_batch(100)
    assert x.shape == (100, 784) and y.shape == (100, 10) and x.dtype == np.float32
    assert x.min() >= 0 and x.max() <= 1 and (y.sum(1) == 1).all()
    seen = 0
    for _ in range(5):
        seen += d.train.next_batch(100)[0].shape[0]
    assert seen == 500 and d.train.epochs_completed >= 1
    # idx files written by hand, read back through the reference-shaped API
    import gzip
    import struct
    img = (np.random.rand(6000, 28, 28) * 255).astype(np.uint8)
    lab = np.random.randint(0, 10, 6000).astype(np.uint8)
    for name, arr, magic in (("train-images-idx3-ubyte", img, 2051), ("t10k-images-idx3-ubyte", img[:100], 2051)):
        with gzip.open(os.path.join(tmp_path, name + ".gz"), "wb") as f:
            f.write(struct.pack(">IIII", magic, arr.shape[0], 28, 28) + arr.tobytes())
    for name, arr in (("train-labels-idx1-ubyte", lab), ("t10k-labels-idx1-ubyte", lab[:100])):
        with gzip.open(os.path.join(tmp_path, name + ".gz"), "wb") as f:
            f.write(struct.pack(">II", 2049, arr.shape[0]) + arr.tobytes())
    m = read_data_sets(str(tmp_path), one_hot=True)
    assert m.train.num_examples == 1000 and m.validation.num_examples == 5000 and m.test.num_examples == 100
    assert m.test.labels.shape == (100, 10) and abs(m.test.images[0, 0] - img[0, 0, 0] / 255.0) < 1e-6
    with pytest.raises(FileNotFoundError):
        read_data_sets(str(tmp_path / "nope"))


@pytest.fixture
def ps():
    c = ClusterSpec({"ps": ["127.0.0.1:0"], "worker": ["127.0.0.1:0"]})
    s = Server(c, "ps", 0)
    yield s
    s.stop()


def test_ps_runtime_variables_and_ready_op(ps):
    a, b = PsClient(ps.target), PsClient(ps.target)
    assert a.var_declare("Variable", (2, 3), torch.float32) is False
    assert b.uninitialized_count() == 1
    with pytest.raises(PsError):
        b.var_read("Variable", (2, 3), torch.float32)
    a.var_assign("Variable", torch.ones(2, 3))
    assert b.var_declare("Variable", (2, 3), torch.float32) is True      # idempotent, already initialised
    assert b.uninitialized_count() == 0
    with pytest.raises(PsError):
        b.var_declare("Variable", (3, 2), torch.float32)                   # shape clash
    a.var_declare("Variable_8", (), torch.int32)
    a.var_assign("Variable_8", torch.zeros((), dtype=torch.int32))
    assert [a.var_add_i32("Variable_8", 1), b.var_add_i32("Variable_8", 1)] == [1, 2]


def test_ps_runtime_optimizers_match_oracle(ps):
    from psb200.optim import AdamOptimizer, RMSPropOptimizer, apply_reference
    c = PsClient(ps.target)
    g = torch.randn(5, 4)
    w0 = torch.randn(5, 4)
    for name, opt, slots, bp in (("r", RMSPropOptimizer(0.01), [torch.ones(5, 4), torch.zeros(5, 4)], (0, 0)),
                                 ("a", AdamOptimizer(2e-4), [torch.zeros(5, 4), torch.zeros(5, 4)], (0.9, 0.999))):
        names = [name, name + "/s0", name + "/s1"]
        for n, v in zip(names, [w0] + slots):
            c.var_declare(n, (5, 4), torch.float32)
            c.var_assign(n, v)
        c.apply(name, g, opt.kind, opt.hyper_vector(), names[1:], bp)
        w, s0, s1 = w0.clone(), slots[0].clone(), slots[1].clone()
        apply_reference(opt, w, g, [s0, s1], bp)
        assert torch.allclose(c.var_read(name, (5, 4), torch.float32), w, rtol=2e-4, atol=1e-7)
        assert torch.allclose(c.var_read(names[1], (5, 4), torch.float32), s0, rtol=2e-4, atol=1e-8)
        assert torch.allclose(c.var_read(names[2], (5, 4), torch.float32), s1, rtol=2e-4, atol=1e-8)


def test_ps_runtime_queue_blocks_and_times_out(ps):
    c = PsClient(ps.target)
    c.queue_create("done_queue0", 2)
    with pytest.raises(QueueTimeout):
        c.queue_dequeue("done_queue0", timeout=0.2)
    got = []
    t = threading.Thread(target=lambda: got.append(PsClient(ps.target).queue_dequeue("done_queue0")))
    t.start()
    time.sleep(0.1)
    assert got == []
    c.queue_enqueue("done_queue0", 5)
    t.join(5)
    assert got == [5]


def test_ps_runtime_accumulator_mean_and_stale_drop(ps):
    c = PsClient(ps.target)
    c.var_declare("Variable", (4,), torch.float32)
    c.var_assign("Variable", torch.zeros(4))
    c.acc_declare("Variable")
    done = []
    t = threading.Thread(target=lambda: done.append(
        PsClient(ps.target).acc_take_apply("Variable", 2, 0, [1.0, 0, 0, 0], [])))
    t.start()
    assert c.acc_apply("Variable", torch.full((4,), 2.0), 0)
    time.sleep(0.1)
    assert done == []                                   # still waiting for the second replica
    assert c.acc_apply("Variable", torch.full((4,), 4.0), 0)
    t.join(5)
    assert done == [2]
    assert torch.allclose(c.var_read("Variable", (4,), torch.float32), torch.full((4,), -3.0))   # w -= mean
    # the take itself advanced the accumulator's step (no window in which a backup worker's old
    # gradient could slip into the next aggregate before acc_set_step arrives)
    assert c.acc_apply("Variable", torch.ones(4), 0) is False     # stale: dropped
    c.acc_set_step("Variable", 1)
    assert c.acc_apply("Variable", torch.ones(4), 0) is False
    assert c.acc_apply("Variable", torch.ones(4), 1) is True
    assert c.stats()["stale_dropped"] == 2


def test_ps_runtime_survives_malformed_requests(ps):
    """Requests a buggy or hostile client could send must be answered with an error, not crash the
    ps process: slot counts that do not match the optimizer kind, missing slots in the sync path,
    applies to uninitialised variables, absurd variable sizes."""
    import struct
    from psb200.control.client import OP_ACC_TAKE_APPLY, OP_APPLY, OP_VAR_DECLARE
    c = PsClient(ps.target)
    c.var_declare("w", (4,), torch.float32)
    c.var_assign("w", torch.zeros(4))
    c.var_declare("s", (4,), torch.float32)
    c.var_assign("s", torch.zeros(4))
    g = torch.ones(4)
    hdr = lambda kind, slots: PsClient._apply_header(kind, [0.1, 0.9, 0.9, 1e-8], (0.9, 0.999), slots)   # noqa: E731
    for kind, slots in ((0, ["s"]), (1, []), (2, ["s"]), (1, ["s", "nope"]), (7, [])):
        with pytest.raises(PsError):
            c._call(OP_APPLY, "w", hdr(kind, slots), "apply", tail=memoryview(g.numpy()).cast("B"))
    c.acc_declare("w")
    c.acc_apply("w", g, 0)
    for kind, slots in ((2, []), (1, ["s"])):
        with pytest.raises(PsError):
            c._call(OP_ACC_TAKE_APPLY, "w", struct.pack("<I", 1) + hdr(kind, slots), "take")
    c.var_declare("u", (4,), torch.float32)                     # never assigned
    c.acc_declare("u")
    with pytest.raises(PsError):
        c._call(OP_ACC_TAKE_APPLY, "u", struct.pack("<I", 0) + hdr(0, []), "take")
    with pytest.raises(PsError):                                # 2^40 x 2^40 elements
        c._call(OP_VAR_DECLARE, "huge", struct.pack("<IIQQ", 0, 2, 1 << 40, 1 << 40), "declare")
    with pytest.raises(PsError):                                # 2^31 elements x 4 B: bad_alloc or refused
        c._call(OP_VAR_DECLARE, "huge2", struct.pack("<IIQ", 0, 1, (1 << 32) + 1), "declare")
    # the server is still alive and consistent
    c.ping()
    assert torch.equal(c.var_read("w", (4,), torch.float32), torch.zeros(4))
    # a closed connection's descriptor is forgotten: stopping the server later cannot hit a reused fd
    PsClient(ps.target).close()


def test_dtype_flag_rounds_weight_matrices_only():
    from psb200.flags import parse_flags
    from psb200.runner import compute_params
    p = {"w": torch.tensor([[1.00390625, 2.0]]), "b": torch.tensor([1.00390625]), "step": torch.tensor(3, dtype=torch.int32)}
    assert compute_params(p, "fp32") is p
    q = compute_params(p, "bf16")
    assert q["w"][0, 0].item() == 1.0 and q["w"].dtype == torch.float32        # 1 + 2^-8 is not a bf16 value
    assert q["b"][0].item() == 1.00390625 and q["step"].dtype == torch.int32
    assert parse_flags(["--dtype=bf16"]).dtype == "bf16"
    with pytest.raises(SystemExit):
        parse_flags(["--dtype=fp8"])
