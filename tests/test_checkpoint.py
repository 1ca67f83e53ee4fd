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
