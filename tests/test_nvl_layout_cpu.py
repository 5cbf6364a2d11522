"""Host-side logic of the NVLink store that needs no GPU: row-partitioned ownership, the block
budget of a push launch, and the Supervisor adapter (`parallel/session.py`) with a stand-in trainer."""
import threading

import torch
from hypothesis import given, settings, strategies as st

from psb200.store.nvl import row_blocks, share_blocks


@settings(max_examples=200, deadline=None)
@given(rows=st.integers(1, 20000), cols=st.integers(1, 4096), home=st.integers(0, 15), world=st.integers(1, 16))
def test_row_blocks_tile_the_variable(rows, cols, home, world):
    home %= world
    blocks = row_blocks(rows, cols, home, world)
    # contiguous cover of [0, rows), no empty block
    assert blocks[0][0] == 0 and blocks[-1][1] == rows
    for (a0, a1, _), (b0, _, _) in zip(blocks, blocks[1:]):
        assert a1 == b0 and a1 > a0
    assert all(0 <= o < world for _, _, o in blocks)
    if len(blocks) == 1:
        # small variables keep replica_device_setter's placement: whole, on their home rank
        assert blocks[0][2] == home
    else:
        assert rows * cols >= 8192 and rows >= 4 * world
        assert len(blocks) <= world
        # boundaries on multiples of four rows (8-byte stores of both bf16 copies), equal blocks but the last
        sizes = [b[1] - b[0] for b in blocks]
        assert all(b[0] % 4 == 0 for b in blocks)
        assert len(set(sizes[:-1])) <= 1 and sizes[-1] <= sizes[0]
        # owners rotate from the home rank: every rank owns at most one block
        assert [o for _, _, o in blocks] == [(home + i) % world for i in range(len(blocks))]


def test_row_blocks_reference_models_at_eight_ranks():
    # the LSTM cell kernel [156, 512] is shared by all eight ranks; the 10-column head stays whole
    assert len(row_blocks(156, 512, 2, 8)) == 8
    assert row_blocks(128, 10, 0, 8) == [(0, 128, 0)]
    assert row_blocks(784, 256, 0, 8)[0] == (0, 100, 0)
    assert row_blocks(784, 256, 0, 8, partition=False) == [(0, 784, 0)]
    # the scaled model: 1024 rows of [8192, 16384] per rank
    assert row_blocks(8192, 16384, 3, 8)[1] == (1024, 2048, 4)


@settings(max_examples=200, deadline=None)
@given(rounds=st.lists(st.integers(1, 200000), min_size=1, max_size=64))
def test_share_blocks_fits_one_wave_and_wastes_nothing(rounds):
    nb = share_blocks(rounds)
    assert len(nb) == len(rounds) and all(1 <= b <= r for b, r in zip(nb, rounds))
    assert sum(nb) <= max(296, len(rounds))
    if sum(rounds) <= 296:
        assert nb == rounds                     # small models: one iteration per block
    # every segment's blocks do at most R iterations, and no segment has more blocks than that needs
    R = max(-(-r // b) for r, b in zip(rounds, nb))
    assert all(-(-r // R) == b for r, b in zip(rounds, nb))


class _FakeStore:
    device = torch.device("cpu")

    def __init__(self):
        self.values = None

    def snapshot_async(self, stream):
        return {k: v.clone() for k, v in self.values.items()}


class _FakeTrainer:
    world = 1

    def __init__(self):
        self.store = _FakeStore()
        self.inits = 0

    def initialize(self, values):
        self.inits += 1
        self.store.values = {k: v.clone() for k, v in values.items()}

    def global_step(self):
        return int(self.store.values.get("Variable_2", torch.zeros(())))


def test_nvl_session_adapter_semantics(monkeypatch, tmp_path):
    """The Supervisor's view of the NVLink transport: chief initialises (collectively), a non-chief's
    ready poll joins that initialisation exactly once, the step counter reads a host integer."""
    import contextlib
    from psb200.parallel import session as sess_mod
    monkeypatch.setattr(torch.cuda, "device", lambda d: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "Stream", lambda device=None: threading.Event())     # has no .synchronize
    tr = _FakeTrainer()
    s = sess_mod.NvlSession(tr, None, is_chief=True)
    assert s.num_workers == 1
    s.declare_all()
    s.sync_setup(True)
    s.initialize({"Variable": torch.ones(3), "Variable_2": torch.tensor(7, dtype=torch.int32)})
    assert tr.inits == 1 and s.global_step() == 7 and s.uninitialized() == 0 and tr.inits == 1
    s.note_step(9)
    assert s.global_step() == 9
    # a non-chief's first poll performs the (collective) initialisation, later polls do not
    tr2 = _FakeTrainer()
    s2 = sess_mod.NvlSession(tr2, None, is_chief=False)
    monkeypatch.setattr(tr2, "initialize", lambda values: (setattr(tr2, "inits", tr2.inits + 1),
                                                           setattr(tr2.store, "values", {"Variable_2": torch.tensor(3)})))
    assert s2.uninitialized() == 0 and s2.uninitialized() == 0 and tr2.inits == 1 and s2.global_step() == 3
