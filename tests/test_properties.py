"""T0 property tests (hypothesis): naming/placement invariants for arbitrary variable sequences,
checkpoint round-trips for arbitrary tensors, the ps apply header packing."""
import struct

import torch
from hypothesis import given, settings
from hypothesis import strategies as st

from psb200.checkpoint import Saver
from psb200.control.client import PsClient
from psb200.placement import VariableRegistry, replica_device_setter, zeros

names = st.one_of(st.none(), st.sampled_from(["Variable", "Variable_1", "w", "rnn/cell/kernel", "beta1_power"]))


@settings(max_examples=60, deadline=None)
@given(st.integers(min_value=1, max_value=9), st.lists(names, min_size=1, max_size=24))
def test_names_unique_and_round_robin(ps, seq):
    reg = VariableRegistry(replica_device_setter(ps_tasks=ps))
    specs = [reg.variable(zeros((2,)), (2,), name=n) for n in seq]
    assert len({s.name for s in specs}) == len(specs)                      # never a duplicate name
    assert [s.ps_task for s in specs] == [i % ps for i in range(len(specs))]  # creation-order round robin
    for s, n in zip(specs, seq):
        base = n or "Variable"
        assert s.name == base or s.name.startswith(base + "_")
    # slots follow their primary and do not disturb the round robin
    slot = reg.slot(specs[0], "Adam", zeros((2,)))
    nxt = reg.variable(zeros((2,)), (2,))
    assert slot.ps_task == specs[0].ps_task and nxt.ps_task == len(specs) % ps


@settings(max_examples=25, deadline=None)
@given(st.lists(st.tuples(st.sampled_from(["float32", "int32"]),
                          st.lists(st.integers(min_value=1, max_value=5), min_size=0, max_size=3)),
                min_size=1, max_size=6),
       st.integers(min_value=0, max_value=10 ** 6))
def test_checkpoint_roundtrip_any_tensors(tmp_path_factory, items, step):
    d = str(tmp_path_factory.mktemp("ck"))
    vals = {}
    for i, (dt, shape) in enumerate(items):
        if dt == "float32":
            vals["v%d" % i] = torch.randn(*shape) if shape else torch.randn(())
        else:
            vals["v%d" % i] = torch.randint(-5, 5, tuple(shape), dtype=torch.int32)
    prefix = Saver().save(vals, d, step)
    back = Saver.restore(prefix)
    assert prefix.endswith("model.ckpt-%d" % step) and set(back) == set(vals)
    for k in vals:
        assert back[k].dtype == vals[k].dtype and back[k].shape == vals[k].shape and torch.equal(back[k], vals[k])


@given(st.integers(min_value=0, max_value=2), st.lists(st.floats(-10, 10, width=32), min_size=4, max_size=4),
       st.lists(st.text(alphabet="abcXYZ_/0123", min_size=1, max_size=20), min_size=0, max_size=2))
def test_apply_header_layout(kind, hyper, slots):
    raw = PsClient._apply_header(kind, hyper, (0.5, 0.25), slots)
    k, lr, h1, h2, eps, b1p, b2p, n = struct.unpack("<IffffffI", raw[:32])
    assert (k, n) == (kind, len(slots)) and (b1p, b2p) == (0.5, 0.25)
    assert struct.pack("<ffff", *hyper) == struct.pack("<ffff", lr, h1, h2, eps)
    off = 32
    for s in slots:
        (ln,) = struct.unpack("<I", raw[off:off + 4])
        assert raw[off + 4:off + 4 + ln].decode() == s
        off += 4 + ln
    assert off == len(raw)
