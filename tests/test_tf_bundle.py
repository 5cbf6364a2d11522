"""TF V2 checkpoint bundle (csrc/tf_bundle.cpp): CRC32C known vectors, table layout bytes,
round trips through Saver, corruption detection."""
import os
import struct

import pytest
import torch

from psb200 import tf_bundle
from psb200.checkpoint import Saver, latest_checkpoint
from psb200.models import make_model
from psb200.placement import VariableRegistry, replica_device_setter

MAGIC = 0xdb4775248b80fb57


def test_crc32c_known_vectors():
    # RFC 3720 B.4 test patterns + the classic check value
    assert tf_bundle.crc32c(b"123456789") == 0xE3069283
    assert tf_bundle.crc32c(bytes(32)) == 0x8A9136AA
    assert tf_bundle.crc32c(b"\xff" * 32) == 0x62A8AB43
    assert tf_bundle.crc32c(bytes(range(32))) == 0x46DD794E
    assert tf_bundle.crc32c(bytes(range(31, -1, -1))) == 0x113FDB5C
    assert tf_bundle.crc32c(b"") == 0
    for n in (1, 7, 8, 9, 63, 1000):
        data = bytes((i * 131 + 7) & 0xff for i in range(n))
        assert tf_bundle.crc32c(data) == tf_bundle.crc32c(data, software=True)


def _varint(buf, pos):
    out, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7f) << shift
        shift += 7
        if not b & 0x80:
            return out, pos


def test_index_file_layout(tmp_path):
    """Walk the table by hand (independent of the native reader): footer, index block, data block,
    header entry first, prefix-compressed keys in sorted order, masked block CRCs."""
    prefix = str(tmp_path / "model.ckpt-7")
    vals = {"Variable": torch.arange(6, dtype=torch.float32).reshape(2, 3),
            "Variable_1": torch.tensor([1.5, -2.0]), "Variable_2": torch.tensor(7, dtype=torch.int32)}
    tf_bundle.write_bundle(prefix, vals)
    raw = open(prefix + ".index", "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == MAGIC and tf_bundle.is_tf_bundle(prefix)
    footer = raw[-48:]
    m_off, p = _varint(footer, 0)
    m_size, p = _varint(footer, p)
    i_off, p = _varint(footer, p)
    i_size, p = _varint(footer, p)
    assert i_off + i_size + 5 + 48 == len(raw)          # index block is the last block before the footer

    def block(off, size):
        body, typ, crc = raw[off:off + size], raw[off + size], struct.unpack("<I", raw[off + size + 1:off + size + 5])[0]
        assert typ == 0 and crc == tf_bundle.masked_crc32c(body + b"\0")
        n_restart = struct.unpack("<I", body[-4:])[0]
        end = len(body) - 4 - 4 * n_restart
        assert struct.unpack("<I", body[end:end + 4])[0] == 0       # first restart point is offset 0
        entries, key, q = [], b"", 0
        while q < end:
            shared, q = _varint(body, q)
            non_shared, q = _varint(body, q)
            vlen, q = _varint(body, q)
            key = key[:shared] + body[q:q + non_shared]
            q += non_shared
            entries.append((key, body[q:q + vlen], shared))
            q += vlen
        return entries

    assert block(m_off, m_size) == []                    # empty meta-index block
    index = block(i_off, i_size)
    assert len(index) == 1
    d_off, q = _varint(index[0][1], 0)
    d_size, _ = _varint(index[0][1], q)
    data = block(d_off, d_size)
    assert [k for k, _, _ in data] == [b"", b"Variable", b"Variable_1", b"Variable_2"]
    assert [s for _, _, s in data] == [0, 0, 8, 9]        # prefix compression against the previous key
    assert index[0][0] >= data[-1][0]
    # header: num_shards=1, version{producer=1};  first entry: DT_FLOAT, shape [2,3], size 24, crc of the bytes
    assert data[0][1] == bytes([0x08, 0x01, 0x1a, 0x02, 0x08, 0x01])
    payload = vals["Variable"].numpy().tobytes()
    want = bytes([0x08, 0x01, 0x12, 0x08, 0x12, 0x02, 0x08, 0x02, 0x12, 0x02, 0x08, 0x03, 0x28, 24, 0x35]) + \
        struct.pack("<I", tf_bundle.masked_crc32c(payload))
    assert data[1][1] == want
    # data shard: tensors back to back in key order
    shard = open(prefix + ".data-00000-of-00001", "rb").read()
    assert shard == payload + vals["Variable_1"].numpy().tobytes() + vals["Variable_2"].numpy().tobytes()


@pytest.mark.parametrize("model", ["encoder", "gan", "lstm"])
def test_saver_tf_format_round_trip(model, tmp_path):
    reg = VariableRegistry(replica_device_setter(ps_tasks=2))
    make_model(model).build(reg)
    vals = reg.initial_values(3)
    saver = Saver(bundle_format="tf")
    prefix = saver.save(vals, str(tmp_path), 42, {"model": model})
    assert os.path.basename(prefix) == "model.ckpt-42" and latest_checkpoint(str(tmp_path)) == prefix
    assert sorted(os.listdir(tmp_path)) == ["checkpoint", "model.ckpt-42.data-00000-of-00001",
                                            "model.ckpt-42.index", "model.ckpt-42.meta"]
    back = Saver.restore(prefix)
    assert set(back) == set(vals)
    for k in vals:
        assert back[k].dtype == vals[k].dtype and back[k].shape == vals[k].shape and torch.equal(back[k], vals[k]), k
    assert Saver.list_variables(prefix) == {k: list(v.shape) for k, v in vals.items()}
    # the default format still round-trips and both are recognised by the same restore()
    p2 = Saver().save(vals, str(tmp_path), 43)
    assert not tf_bundle.is_tf_bundle(p2) and torch.equal(Saver.restore(p2)["Variable"], vals["Variable"])


def test_many_keys_span_several_blocks_and_restart_points(tmp_path):
    prefix = str(tmp_path / "big")
    vals = {"layer_%05d/w" % i: torch.full((3,), float(i)) for i in range(15000)}
    vals["bf"] = torch.randn(4, 5).to(torch.bfloat16)
    vals["empty"] = torch.zeros(0, 4)
    tf_bundle.write_bundle(prefix, vals)
    assert os.path.getsize(prefix + ".index") > 262144          # more than one data block
    back = tf_bundle.read_bundle(prefix)
    assert len(back) == len(vals) and list(back) == sorted(vals)
    assert all(torch.equal(back[k], vals[k]) for k in vals)


def test_corruption_is_detected(tmp_path):
    prefix = str(tmp_path / "c")
    tf_bundle.write_bundle(prefix, {"a": torch.arange(100, dtype=torch.float32), "b": torch.ones(3)})
    data_path = prefix + ".data-00000-of-00001"
    blob = bytearray(open(data_path, "rb").read())
    blob[17] ^= 0x40
    open(data_path, "wb").write(bytes(blob))
    with pytest.raises(OSError, match="checksum mismatch for a"):
        tf_bundle.read_bundle(prefix)
    idx = bytearray(open(prefix + ".index", "rb").read())
    idx[5] ^= 0x01
    open(prefix + ".index", "wb").write(bytes(idx))
    with pytest.raises(OSError, match="block checksum"):
        tf_bundle.list_variables(prefix)
    open(prefix + ".index", "wb").write(b"{}")
    assert not tf_bundle.is_tf_bundle(prefix)


def test_random_bundles_round_trip():
    """Property: any set of distinct names / small tensors survives write -> read bit-exactly."""
    from hypothesis import given, settings, strategies as st
    import tempfile

    names = st.text(alphabet="abcXYZ_/0123456789", min_size=1, max_size=24)
    shapes = st.lists(st.integers(0, 5), min_size=0, max_size=3)
    dtypes = st.sampled_from([torch.float32, torch.int32, torch.float64, torch.int64, torch.float16, torch.bfloat16])

    @settings(max_examples=40, deadline=None)
    @given(st.dictionaries(names, st.tuples(shapes, dtypes, st.integers(0, 2 ** 31 - 1)), min_size=1, max_size=12))
    def check(spec):
        vals = {}
        for name, (shape, dt, seed) in spec.items():
            g = torch.Generator().manual_seed(seed)
            t = torch.randint(-1000, 1000, shape, generator=g)
            vals[name] = t.to(dt) if not dt.is_floating_point else (t.float() / 7).to(dt)
        with tempfile.TemporaryDirectory() as d:
            prefix = os.path.join(d, "m")
            tf_bundle.write_bundle(prefix, vals)
            back = tf_bundle.read_bundle(prefix)
            assert list(back) == sorted(vals)
            for k, v in vals.items():
                assert back[k].dtype == v.dtype and list(back[k].shape) == list(v.shape)
                assert back[k].reshape(-1).contiguous().numpy().tobytes() == v.reshape(-1).contiguous().numpy().tobytes() \
                    if v.dtype != torch.bfloat16 else torch.equal(back[k].view(torch.int16), v.view(torch.int16)), k

    check()


def test_inspect_checkpoint_tool(tmp_path, capsys):
    from psb200.tools import inspect_checkpoint as ic
    d = str(tmp_path)
    Saver(bundle_format="tf").save({"Variable": torch.ones(2, 3), "Variable_2": torch.tensor(5, dtype=torch.int32)}, d, 5)
    assert ic.main(["--file_name", d]) == 0
    out = capsys.readouterr().out
    assert "model.ckpt-5 (tf bundle, 2 variables)" in out and "Variable (DT_FLOAT) [2, 3]" in out and "Variable_2 (DT_INT32) []" in out
    assert ic.main(["--file_name", os.path.join(d, "model.ckpt-5"), "--tensor_name", "Variable_2"]) == 0
    assert capsys.readouterr().out.strip().endswith("5")
    assert ic.main(["--file_name", d, "--tensor_name", "nope"]) == 1
    assert ic.main(["--file_name", os.path.join(d, "missing")]) == 1
