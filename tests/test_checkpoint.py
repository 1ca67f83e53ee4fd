"""TF V2 bundle reader/writer + Estimator-style directory management."""
import os
import struct

import numpy as np
import pytest

from deepfm_b200.utils import checkpoint as ck


def _tensors():
    rng = np.random.default_rng(0)
    t = {"fm_bias": np.zeros(1, np.float32), "fm_w": rng.random(1000).astype(np.float32),
         "fm_v": rng.random((1000, 32)).astype(np.float32),
         "Deep-part/mlp0/weights": rng.random((1248, 256)).astype(np.float32),
         "Deep-part/mlp0/biases": np.zeros(256, np.float32),
         "Deep-part/mlp0/weights/Adam": np.zeros((1248, 256), np.float32),
         "beta1_power": np.asarray(0.9, np.float32)}
    for i in range(300):                       # enough keys to need several restart points
        t[f"zz/var{i:04d}"] = np.arange(i + 1, dtype=np.float32)
    return t


def test_bundle_roundtrip_names_shapes_dtypes(tmp_path):
    t = _tensors()
    t["global_step"] = np.asarray(1234, np.int64)
    prefix = str(tmp_path / "model.ckpt-1234")
    ck.write_bundle(prefix, t)
    assert os.path.exists(prefix + ".index") and os.path.exists(prefix + ".data-00000-of-00001")
    listing = ck.list_bundle(prefix)
    assert list(listing.keys()) == sorted(t.keys(), key=lambda s: s.encode())
    assert listing["fm_v"][:2] == (np.dtype("float32"), (1000, 32))
    assert listing["global_step"][:2] == (np.dtype("int64"), ())
    r = ck.read_bundle(prefix)
    assert all(np.array_equal(r[k], t[k]) and r[k].shape == np.asarray(t[k]).shape for k in t)
    one = ck.read_bundle(prefix, names=["fm_w"])
    assert list(one) == ["fm_w"]


def test_index_is_a_leveldb_table(tmp_path):
    prefix = str(tmp_path / "m")
    ck.write_bundle(prefix, {"a": np.ones(3, np.float32)})
    raw = open(prefix + ".index", "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == 0xDB4775248B80FB57      # table magic
    assert os.path.getsize(prefix + ".data-00000-of-00001") == 12


def test_corruption_is_detected(tmp_path):
    prefix = str(tmp_path / "m")
    ck.write_bundle(prefix, _tensors())
    p = prefix + ".data-00000-of-00001"
    raw = bytearray(open(p, "rb").read())
    raw[100] ^= 0x55
    open(p, "wb").write(raw)
    with pytest.raises(ValueError, match="CRC"):
        ck.read_bundle(prefix)


def test_manager_keeps_last_n_and_resumes(tmp_path):
    d = str(tmp_path / "ckpt")
    m = ck.CheckpointManager(d, keep_max=2, save_secs=0)
    t = _tensors()
    for step in (10, 20, 30):
        m.save(step, t)
    files = sorted(os.listdir(d))
    assert files == ["checkpoint", "model.ckpt-20.data-00000-of-00001", "model.ckpt-20.index",
                     "model.ckpt-30.data-00000-of-00001", "model.ckpt-30.index"]
    assert ck.latest_checkpoint(d).endswith("model.ckpt-30")
    assert 'model_checkpoint_path: "model.ckpt-30"' in open(os.path.join(d, "checkpoint")).read()
    step, r = ck.CheckpointManager(d).restore_latest()
    assert step == 30 and np.array_equal(r["fm_v"], t["fm_v"]) and "global_step" not in r
    assert ck.CheckpointManager(str(tmp_path / "empty")).restore_latest() is None


from hypothesis import given, settings, strategies as st  # noqa: E402


@settings(max_examples=40, deadline=None)
@given(seed=st.integers(0, 2 ** 31), n_vars=st.integers(1, 60))
def test_bundle_roundtrip_property(tmp_path_factory, seed, n_vars):
    """Any set of names (shared prefixes stress the SSTable's key compression), ranks 0-3, f32 / i64 / i32: identity."""
    rng = np.random.default_rng(seed)
    stems = ["a", "a/b", "a/b/c", "Deep-part/mlp", "Deep-part/mlp0/weights", "fm_v", "z" * 40]
    t = {}
    for i in range(n_vars):
        name = f"{stems[int(rng.integers(len(stems)))]}{int(rng.integers(0, 10 ** int(rng.integers(1, 6))))}"
        shape = tuple(int(x) for x in rng.integers(1, 7, int(rng.integers(0, 4))))
        dt = [np.float32, np.int64, np.int32][int(rng.integers(3))]
        t[name] = (rng.standard_normal(shape) * 100).astype(dt)
    prefix = str(tmp_path_factory.mktemp("b") / "model.ckpt-7")
    ck.write_bundle(prefix, t)
    r = ck.read_bundle(prefix)
    assert set(r) == set(t)
    for k in t:
        assert r[k].dtype == t[k].dtype and r[k].shape == t[k].shape and np.array_equal(r[k], t[k]), k


# ---------------------------------------------------------------------------------------------------
# An index file assembled BY HAND from the published formats (LevelDB table_format.md; tensor_bundle.proto), with
# none of the module's own encoding helpers: if the module reads it, and this file's own little parser reads what the
# module writes, reader and writer agree with the spec and not merely with each other.  (No TensorFlow exists for this
# Python, so a TF-written golden file cannot be produced here: DESIGN.md section 6 says "unverified against TF".)
# ---------------------------------------------------------------------------------------------------
def _crc32c(data: bytes) -> int:
    crc = 0xFFFFFFFF
    for b in data:
        crc ^= b
        for _ in range(8):
            crc = (crc >> 1) ^ (0x82F63B78 & -(crc & 1))
    return crc ^ 0xFFFFFFFF


def _masked(data: bytes) -> int:
    c = _crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def _vi(v: int) -> bytes:
    out = b""
    while v >= 0x80:
        out += bytes([(v & 0x7F) | 0x80])
        v >>= 7
    return out + bytes([v])


def _block(entries) -> bytes:
    """One LevelDB block with a restart point at every entry (shared = 0), then the restart array and its length."""
    body, restarts = b"", []
    for k, v in entries:
        restarts.append(len(body))
        body += _vi(0) + _vi(len(k)) + _vi(len(v)) + k + v
    return body + b"".join(struct.pack("<I", r) for r in restarts) + struct.pack("<I", len(restarts))


def _with_trailer(block: bytes) -> bytes:
    return block + b"\x00" + struct.pack("<I", _masked(block + b"\x00"))      # type 0 = uncompressed, masked CRC32C


def _hand_made_index(value_bytes: bytes) -> bytes:
    header = bytes([0x08, 0x01, 0x1A, 0x02, 0x08, 0x01])          # BundleHeaderProto{num_shards: 1, version{producer: 1}}
    shape = bytes([0x12, 0x02, 0x08, 0x03])                        # TensorShapeProto{dim{size: 3}}
    entry = bytes([0x08, 0x01]) + bytes([0x12, len(shape)]) + shape + bytes([0x28, len(value_bytes)]) + \
        bytes([0x35]) + struct.pack("<I", _masked(value_bytes))    # dtype DT_FLOAT, shape, size, crc32c (fixed32)
    data = _with_trailer(_block([(b"", header), (b"fm_w", entry)]))
    meta_off = len(data)
    meta = _with_trailer(_block([]) if False else struct.pack("<II", 0, 1))   # empty block: restarts [0], count 1
    index_off = meta_off + len(meta)
    index = _with_trailer(_block([(b"fm_x", _vi(0) + _vi(len(data) - 5))]))    # separator >= last key -> handle of block 0
    footer = _vi(meta_off) + _vi(len(meta) - 5) + _vi(index_off) + _vi(len(index) - 5)
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", 0xDB4775248B80FB57)
    return data + meta + index + footer


def _spec_parse_index(raw: bytes):
    """Minimal reader straight from table_format.md: footer -> index block -> data blocks -> (key, value) pairs."""
    def vi(buf, pos):
        v = shift = 0
        while True:
            b = buf[pos]; pos += 1
            v |= (b & 0x7F) << shift
            if b < 0x80:
                return v, pos
            shift += 7

    def block(off, size):
        blk = raw[off:off + size]
        assert struct.unpack("<I", raw[off + size + 1:off + size + 5])[0] == _masked(raw[off:off + size + 1])
        n_restarts = struct.unpack("<I", blk[-4:])[0]
        end = len(blk) - 4 - 4 * n_restarts
        pos, key, out = 0, b"", []
        while pos < end:
            shared, pos = vi(blk, pos); non_shared, pos = vi(blk, pos); vlen, pos = vi(blk, pos)
            key = key[:shared] + blk[pos:pos + non_shared]; pos += non_shared
            out.append((key, blk[pos:pos + vlen])); pos += vlen
        return out

    assert struct.unpack("<Q", raw[-8:])[0] == 0xDB4775248B80FB57
    f = raw[-48:-8]
    _, p = vi(f, 0); _, p = vi(f, p)
    ioff, p = vi(f, p); isz, p = vi(f, p)
    items = []
    for _, handle in block(ioff, isz):
        o, q = vi(handle, 0); s, _ = vi(handle, q)
        items += block(o, s)
    return items


def test_reader_accepts_an_index_assembled_from_the_format_spec(tmp_path):
    vals = np.asarray([1.5, -2.0, 3.25], np.float32)
    prefix = str(tmp_path / "model.ckpt-1")
    open(prefix + ".index", "wb").write(_hand_made_index(vals.tobytes()))
    open(prefix + ".data-00000-of-00001", "wb").write(vals.tobytes())
    r = ck.read_bundle(prefix)
    assert list(r) == ["fm_w"] and r["fm_w"].dtype == np.float32 and np.array_equal(r["fm_w"], vals)


def test_writer_output_parses_with_a_reader_built_from_the_format_spec(tmp_path):
    prefix = str(tmp_path / "m")
    t = _tensors()
    ck.write_bundle(prefix, t)
    items = _spec_parse_index(open(prefix + ".index", "rb").read())
    keys = [k for k, _ in items]
    assert keys[0] == b"" and keys[1:] == sorted(k.encode() for k in t)          # header entry first, then sorted names
    assert items[0][1][:2] == bytes([0x08, 0x01])                                # BundleHeaderProto.num_shards = 1
    entry = dict(items)[b"fm_v"]
    assert entry[:2] == bytes([0x08, 0x01]) and bytes([0x12]) in entry           # DT_FLOAT, then the shape message
