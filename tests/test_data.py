"""Native IO layer: CRC32C, framing, Example codec, sharding, loader, converter, generator.
Golden facts about the reference's bundled data/val.tfrecords are from SURVEY 2.8."""
import os
import struct
import subprocess
import threading

import numpy as np
import pytest
import torch

from deepfm_b200.data import tfrecord as T
from deepfm_b200.data.loader import BatchLoader
from deepfm_b200.data.sharding import hvd_shard, ps_shard, shard_batches_per_epoch


def test_crc32c_known_vectors():
    assert T.crc32c(b"123456789") == 0xE3069283
    assert T.crc32c(b"") == 0
    assert T.crc32c(bytes(32)) == 0x8A9136AA
    c = T.crc32c(b"hello")
    assert T.masked_crc32c(b"hello") == (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def test_val_tfrecords_golden_facts(val_tfrecords):
    if "reference" not in val_tfrecords:
        pytest.skip("golden numbers are for the reference's bundled file")
    assert T.count_records(val_tfrecords) == 10000
    ids, vals, labels = T.read_all(val_tfrecords, 39)
    assert ids.shape == (10000, 39) and ids.min() == 1 and ids.max() == 117565
    assert len(np.unique(ids)) == 28116
    assert abs(labels.mean() - 0.2553) < 1e-4
    assert (ids[:, :13] == np.arange(1, 14)).all()              # numeric fields: constant ids 1..13
    assert (vals[:, 13:] == 1.0).all()                          # categorical fields: value 1.0
    lens = {}
    order = {"label_first": 0, "values_first": 0}
    for p in T.iter_records(val_tfrecords):
        lens[len(p)] = lens.get(len(p), 0) + 1
        order["label_first" if p.index(b"label") < p.index(b"values") else "values_first"] += 1
    assert lens == {297: 9038, 298: 952, 299: 10}
    assert order == {"label_first": 5066, "values_first": 4934}   # key order is not stable


@pytest.mark.parametrize("packed", [True, False])
def test_example_roundtrip_packed_and_unpacked(packed):
    rng = np.random.default_rng(0)
    ids = rng.integers(0, 2 ** 40, 39)
    vals = rng.random(39).astype(np.float32)
    payload = T.encode_example(1.0, ids, vals, packed=packed)
    label, i2, v2 = T.parse_example(payload, 39)
    assert label == 1.0 and (i2 == ids).all() and (v2 == vals).all()
    _, i32, _ = T.parse_example(T.encode_example(0.0, ids % 100000, vals, packed=packed), 39, ids_dtype=np.int32)
    assert i32.dtype == np.int32 and (i32 == ids % 100000).all()


def test_example_key_order_independent_and_errors():
    ids, vals = np.arange(5), np.ones(5, np.float32)
    p = T.encode_example(1.0, ids, vals)
    # rebuild with the map entries reversed: Example{features{entry*}}
    body = p[2:] if p[1] < 0x80 else p[3:]
    entries, pos = [], 0
    while pos < len(body):
        assert body[pos] == 0x0A
        n = body[pos + 1]
        entries.append(body[pos:pos + 2 + n])
        pos += 2 + n
    rev = b"".join(reversed(entries))
    q = bytes([0x0A, len(rev)]) + rev
    label, i2, v2 = T.parse_example(q, 5)
    assert label == 1.0 and (i2 == ids).all()
    with pytest.raises(T.DataError, match="field_size"):
        T.parse_example(p, 6)
    with pytest.raises(T.DataError):
        T.parse_example(b"\x0a\x05abc", 5)


def test_writer_reader_roundtrip_and_corruption(tmp_path):
    f = str(tmp_path / "a.tfrecords")
    with T.TFRecordWriter(f) as w:
        for i in range(100):
            w.write_example(i % 2, np.arange(39) + i, np.full(39, i, np.float32))
    assert T.count_records(f) == 100
    ids, vals, labels = T.read_all(f, 39)
    assert (ids[7] == np.arange(39) + 7).all() and vals[9, 0] == 9 and labels[3] == 1
    raw = bytearray(open(f, "rb").read())
    n = struct.unpack("<Q", raw[:8])[0]
    assert struct.unpack("<I", raw[8:12])[0] == T.masked_crc32c(bytes(raw[:8]))
    assert struct.unpack("<I", raw[12 + n:16 + n])[0] == T.masked_crc32c(bytes(raw[12:12 + n]))
    raw[40] ^= 0xFF
    open(f, "wb").write(raw)
    with pytest.raises(T.DataError, match="CRC"):
        T.count_records(f, verify_crc=True)


def test_shard_policies_match_reference_matrix():
    # (pipe, s3shard, multipath, worker_per_host, hosts, rank, local_rank, world)
    assert hvd_shard(0, True, False, 4, 1, 2, 2, 4) == (4, 2)       # HVD.py:130-131
    assert hvd_shard(0, False, False, 4, 1, 2, 2, 4) == (4, 2)      # HVD.py:132-133 (size, rank)
    assert hvd_shard(0, False, False, 4, 2, 5, 1, 8) == (8, 5)
    assert hvd_shard(1, True, True, 4, 2, 5, 1, 8) == (1, 0)        # HVD.py:140-141
    assert hvd_shard(1, False, True, 4, 2, 5, 1, 8) == (2, 1)       # HVD.py:142-144
    assert hvd_shard(1, False, True, 4, 1, 3, 3, 4) == (1, 0)
    assert hvd_shard(1, True, False, 4, 2, 5, 1, 8) == (4, 1)       # HVD.py:146-147
    assert hvd_shard(1, False, False, 4, 2, 5, 1, 8) == (8, 5)      # HVD.py:148-149
    assert ps_shard(False, ["algo-1", "algo-2"], "algo-2") == (2, 1)   # PS.py:153-156
    assert ps_shard(True, ["algo-1", "algo-2"], "algo-2") == (1, 0)
    assert shard_batches_per_epoch(10000, 4, 3, 1024) == 2
    assert shard_batches_per_epoch(10, 4, 1, 2, drop_remainder=False) == 2


def _write(path, n, start=0):
    with T.TFRecordWriter(path) as w:
        for i in range(start, start + n):
            w.write_example(i % 2, np.full(4, i), np.full(4, i, np.float32))


def test_loader_shard_batch_repeat_order(tmp_path):
    a, b = str(tmp_path / "tr_a.tfrecords"), str(tmp_path / "tr_b.tfrecords")
    _write(a, 50)
    _write(b, 37, start=50)
    # shard (3,1) over the concatenated stream, batch 8, drop remainder, 2 epochs
    got = [hb.feat_ids[:, 0].clone() for hb in BatchLoader([a, b], 4, 8, num_epochs=2, shard_n=3, shard_i=1,
                                                            ids_dtype=torch.int64, num_workers=3, num_slots=3)]
    mine = [i for i in range(87) if i % 3 == 1]
    per_epoch = len(mine) // 8
    assert len(got) == 2 * per_epoch
    for e in range(2):
        for j in range(per_epoch):
            assert got[e * per_epoch + j].tolist() == mine[j * 8:(j + 1) * 8]
    # keep the remainder (eval semantics of this framework)
    tail = [hb.count for hb in BatchLoader([a, b], 4, 8, drop_remainder=False, ids_dtype=torch.int32)]
    assert sum(tail) == 87 and tail[-1] == 87 % 8


def test_loader_reports_bad_records(tmp_path):
    f = str(tmp_path / "bad.tfrecords")
    with T.TFRecordWriter(f) as w:
        w.write_example(1, np.arange(4), np.ones(4, np.float32))
        w.write_example(1, np.arange(5), np.ones(5, np.float32))     # wrong field count
    with pytest.raises(T.DataError, match="field_size"):
        list(BatchLoader([f], 4, 1))


def test_pipe_mode_streams_from_fifo(tmp_path):
    """SageMaker pipe mode (HVD.py:136): records arrive through a FIFO, not a seekable file."""
    src, fifo = str(tmp_path / "src.tfrecords"), str(tmp_path / "training_0")
    _write(src, 40)
    os.mkfifo(fifo)
    t = threading.Thread(target=lambda: open(fifo, "wb").write(open(src, "rb").read()))
    t.start()
    got = [hb.feat_ids[:, 0].tolist() for hb in BatchLoader([fifo], 4, 8, shard_n=2, shard_i=0, streaming=True,
                                                            ids_dtype=torch.int64)]
    t.join()
    assert got == [list(range(0, 16, 2)), list(range(16, 32, 2))]


def test_libsvm_converter(tmp_path):
    txt = tmp_path / "tr.libsvm"
    txt.write_text("1 1:0.5 2:0.03519 3:1\n0 4:1 5:2.5 6:1\n\n")
    out = str(tmp_path / "tr.tfrecords")
    assert T.libsvm_to_tfrecord(str(txt), out, expect_fields=3) == 2
    ids, vals, labels = T.read_all(out, 3)
    assert ids.tolist() == [[1, 2, 3], [4, 5, 6]] and labels.tolist() == [1.0, 0.0]
    assert np.allclose(vals[0], [0.5, 0.03519, 1.0])
    txt.write_text("1 1:0.5 2:1\n")
    with pytest.raises(T.DataError, match="expected 3"):
        T.libsvm_to_tfrecord(str(txt), out, expect_fields=3)


def test_synthetic_generator_is_criteo_shaped(tmp_path):
    f = str(tmp_path / "tr_syn.tfrecords")
    assert T.generate_synthetic(f, 4096, 39, 117581, n_dense=13, seed=7) == 4096
    ids, vals, labels = T.read_all(f, 39)
    assert (ids[:, :13] == np.arange(1, 14)).all() and (vals[:, 13:] == 1.0).all()
    assert ids.max() < 117581 and ids[:, 13:].min() >= 14
    assert 0.15 < labels.mean() < 0.40
    uniq = len(np.unique(ids[:2048]))
    assert 2000 < uniq < 40000       # skewed: far fewer distinct ids than 2048*39 lookups
    ids2, _, _ = T.read_all(f, 39)
    assert (ids2 == ids).all()


# ---------------------------------------------------------------- property tests (hypothesis)
from hypothesis import given, settings, strategies as st  # noqa: E402


@settings(max_examples=150, deadline=None)
@given(n=st.integers(1, 64), packed=st.booleans(), seed=st.integers(0, 2 ** 31), label=st.sampled_from([0.0, 1.0, 0.5]))
def test_example_codec_roundtrip_property(n, packed, seed, label):
    """encode -> parse is the identity for any field count, id magnitude and encoding."""
    rng = np.random.default_rng(seed)
    ids = rng.integers(0, 2 ** 62, n)
    vals = rng.standard_normal(n).astype(np.float32)
    lab, i2, v2 = T.parse_example(T.encode_example(label, ids, vals, packed=packed), n)
    assert lab == np.float32(label) and (i2 == ids).all() and (v2 == vals).all()


@settings(max_examples=300, deadline=None)
@given(seed=st.integers(0, 2 ** 31), n_flips=st.integers(1, 8), cut=st.integers(0, 40))
def test_example_parser_survives_corruption(seed, n_flips, cut):
    """A damaged payload is either rejected with DataError or parsed into well-formed arrays -- never a crash."""
    rng = np.random.default_rng(seed)
    payload = bytearray(T.encode_example(1.0, rng.integers(0, 10 ** 6, 39), rng.random(39).astype(np.float32),
                                         packed=bool(seed & 1)))
    for _ in range(n_flips):
        payload[int(rng.integers(0, len(payload)))] = int(rng.integers(0, 256))
    data = bytes(payload[: len(payload) - cut]) if cut else bytes(payload)
    try:
        lab, ids, vals = T.parse_example(data, 39)
    except T.DataError:
        return
    assert ids.shape == (39,) and vals.shape == (39,) and np.isfinite(lab) | np.isnan(lab) | np.isinf(lab)


@settings(max_examples=40, deadline=None)
@given(seed=st.integers(0, 2 ** 31), n=st.integers(1, 300), batch=st.integers(1, 64), workers=st.integers(1, 4),
       shard_n=st.integers(1, 3), n_files=st.integers(1, 4), streaming=st.booleans())
def test_loader_delivers_every_record_once_in_order(tmp_path_factory, seed, n, batch, workers, shard_n, n_files, streaming):
    """Whatever the batch size, worker count, shard, number of files (some of them empty, batches span them) and
    transport (mapped / block stream): the loader yields exactly the shard's records, in file order -- the record index
    of the shard runs on across files."""
    d = tmp_path_factory.mktemp("prop")
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, 2, n).astype(np.float32)
    first_ids = rng.integers(1, 10 ** 6, n)
    cuts = sorted(int(c) for c in rng.integers(0, n + 1, n_files - 1))
    bounds = [0] + cuts + [n]
    paths = []
    for f in range(n_files):
        paths.append(str(d / f"x{f}.tfrecords"))
        with T.TFRecordWriter(paths[-1]) as w:
            for i in range(bounds[f], bounds[f + 1]):
                ids = np.full(5, first_ids[i]); ids[1:] = np.arange(4)
                w.write(T.encode_example(float(labels[i]), ids, np.ones(5, np.float32), packed=bool(i & 1)))
    from deepfm_b200.data.loader import BatchLoader
    for shard_i in range(shard_n):
        got = []
        loader = BatchLoader(paths, 5, batch, num_epochs=1, drop_remainder=False, shard_n=shard_n, shard_i=shard_i,
                             num_workers=workers, num_slots=3, ids_dtype=torch.int64, pin_memory=False,
                             streaming=streaming)
        for hb in loader:
            got.extend(zip(hb.feat_ids[:, 0].tolist(), hb.labels.tolist()))
        loader.close()
        want = [(int(first_ids[i]), float(labels[i])) for i in range(n) if i % shard_n == shard_i]
        assert got == want


def test_loader_rejects_out_of_range_ids(tmp_path):
    """a too-small --feature_size (or a corrupt record) must fail the batch with file and offset, not reach
    the embedding tables."""
    import torch
    from deepfm_b200.data.loader import BatchLoader
    from deepfm_b200.data.tfrecord import DataError, generate_synthetic
    path = str(tmp_path / "tr.tfrecords")
    generate_synthetic(path, 2000, 39, 5000, n_dense=13, seed=3, pos_rate=0.25, zipf_s=1.05)
    ok = BatchLoader([path], 39, 100, feature_size=5000, ids_dtype=torch.int32, num_workers=2)
    assert sum(hb.count for hb in ok) == 2000
    ok.close()
    bad = BatchLoader([path], 39, 100, feature_size=1000, ids_dtype=torch.int32, num_workers=2)
    with pytest.raises(DataError) as ei:
        for _ in bad:
            pass
    assert "outside [0, feature_size = 1000)" in str(ei.value) and "tr.tfrecords" in str(ei.value) and "offset" in str(ei.value)
    bad.close()


def test_parallel_framing_index_matches_sequential_read(tmp_path):
    """Files larger than one index range (1 MiB) are walked by several threads; the batches must be the records in file
    order, whatever the shard, batch size or worker count."""
    import numpy as np
    import torch
    from deepfm_b200.data.loader import BatchLoader
    from deepfm_b200.data import tfrecord as T
    path = str(tmp_path / "big.tfrecords")
    n = 12000                                                   # ~3.7 MB: four index ranges
    T.generate_synthetic(path, n, 39, 117581, n_dense=13, seed=11, pos_rate=0.25, zipf_s=1.05)
    ref_ids = np.stack([T.parse_example(r, 39)[1] for r in T.iter_records(path)])
    for workers, batch, shard in ((1, 256, (1, 0)), (7, 100, (1, 0)), (4, 64, (3, 1))):
        ld = BatchLoader([path], 39, batch, num_epochs=2, drop_remainder=False, shard_n=shard[0], shard_i=shard[1],
                         num_workers=workers, ids_dtype=torch.int64)
        got = np.concatenate([hb.feat_ids.numpy().copy() for hb in ld])
        ld.close()
        want = ref_ids[shard[1]::shard[0]]
        assert got.shape[0] == 2 * want.shape[0]
        assert np.array_equal(got[:want.shape[0]], want) and np.array_equal(got[want.shape[0]:], want)


def test_loader_can_be_closed_mid_stream(tmp_path):
    """Closing a loader that still has framing-index ranges in flight must not touch an unmapped file (regression:
    the batcher dropped the mapping while a worker was still walking it)."""
    import torch
    from deepfm_b200.data.loader import BatchLoader
    from deepfm_b200.data.tfrecord import generate_synthetic
    path = str(tmp_path / "big.tfrecords")
    generate_synthetic(path, 40000, 39, 117581, n_dense=13, seed=5)      # ~12 MB: a dozen index ranges
    for it in range(40):
        ld = BatchLoader([path], 39, 512, num_epochs=1000, num_workers=6, num_slots=4, ids_dtype=torch.int32)
        for i, _ in enumerate(ld):
            if i == it % 5:
                break
        ld.close()


# ---------------------------------------------------------------- fast paths of the codec (canonical layout, SIMD id runs)
def _crc32c_bitwise(data: bytes) -> int:
    c = 0xFFFFFFFF
    for b in data:
        c ^= b
        for _ in range(8):
            c = (c >> 1) ^ (0x82F63B78 & -(c & 1))
    return c ^ 0xFFFFFFFF


def test_crc32c_three_lane_joins_match_bitwise_definition():
    """Every length class of the interleaved CRC (3 x 1024 / 256 / 96 / 32 byte lanes, 8-byte words, byte tail), at
    every alignment of the first byte."""
    rng = np.random.default_rng(5)
    blob = rng.integers(0, 256, 8200, dtype=np.uint8).tobytes()
    lengths = [0, 1, 7, 8, 9, 95, 96, 97, 287, 288, 289, 297, 298, 299, 767, 768, 769, 1000, 3071, 3072, 3073, 3500,
               4096, 7000, 8190]
    for n in lengths:
        for off in (0, 1, 3, 5, 7):
            assert T.crc32c(blob[off:off + n]) == _crc32c_bitwise(blob[off:off + n]), (n, off)


_ID_CLASSES = [(0, 128), (128, 16384), (16384, 2 ** 21), (2 ** 21, 2 ** 28), (2 ** 40, 2 ** 62)]


@settings(max_examples=400, deadline=None)
@given(n=st.integers(1, 80), seed=st.integers(0, 2 ** 31), wide=st.booleans(), i32=st.booleans(),
       run=st.sampled_from([0, 1, 2, 3]))
def test_example_fast_path_matches_input_for_mixed_varint_widths(n, seed, wide, i32, run):
    """The canonical-layout decoder takes ids in steps of four when it sees a run of one- or three-byte varints and one
    at a time otherwise; ids of four bytes and more send the record to the generic walker.  Any mix must round-trip."""
    rng = np.random.default_rng(seed)
    top = len(_ID_CLASSES) if (wide and not i32) else (4 if wide else 3)
    if run == 0:       # every id its own width
        cls = rng.integers(0, top, n)
    elif run == 1:     # Criteo shape: a prefix of small ids, then large ones
        k = int(rng.integers(0, n + 1))
        cls = np.concatenate([np.zeros(k, np.int64), np.full(n - k, 2)])
    elif run == 2:     # long runs with one odd id somewhere
        cls = np.full(n, int(rng.integers(0, 3)))
        cls[int(rng.integers(0, n))] = int(rng.integers(0, top))
    else:              # all three-byte
        cls = np.full(n, 2)
    ids = np.array([int(rng.integers(*_ID_CLASSES[c])) for c in cls], dtype=np.int64)
    vals = rng.standard_normal(n).astype(np.float32)
    payload = T.encode_example(0.25, ids, vals, packed=True)
    lab, i2, v2 = T.parse_example(payload, n, ids_dtype=np.int32 if i32 else np.int64)
    assert lab == np.float32(0.25) and (i2 == ids).all() and (v2 == vals).all()
    assert i2.dtype == (np.int32 if i32 else np.int64)


def test_canonical_and_generic_decoders_agree_on_reordered_and_padded_records():
    """A record the fast path declines (unknown extra key, unpacked lists) must decode to the same arrays."""
    rng = np.random.default_rng(11)
    ids = np.concatenate([np.arange(1, 14), rng.integers(16384, 117581, 26)])
    vals = rng.random(39).astype(np.float32)
    canon = T.encode_example(1.0, ids, vals, packed=True)
    lab, i1, v1 = T.parse_example(canon, 39, ids_dtype=np.int32)
    for other in (T.encode_example(1.0, ids, vals, packed=False),
                  _with_extra_entry(canon)):
        lab2, i2, v2 = T.parse_example(other, 39, ids_dtype=np.int32)
        assert lab2 == lab and (i2 == i1).all() and (v2 == v1).all()
    assert (i1 == ids).all()


def _with_extra_entry(payload: bytes) -> bytes:
    """Example{features{... + entry{key 'zz', value Feature{bytes_list{'x'}}}}}"""
    assert payload[0] == 0x0A
    body = payload[2:] if payload[1] < 0x80 else payload[3:]
    extra = bytes([0x0A, 0x0B, 0x0A, 0x02]) + b"zz" + bytes([0x12, 0x05, 0x0A, 0x03, 0x0A, 0x01]) + b"x"
    body = body + extra
    n = len(body)
    return bytes([0x0A, (n & 0x7F) | 0x80, n >> 7]) + body if n >= 128 else bytes([0x0A, n]) + body


def test_loader_range_check_inside_a_vector_step(tmp_path):
    """One id of a three-byte run is out of range: the loader names it exactly (the vector step declines, the generic
    walker reports)."""
    from deepfm_b200.data.tfrecord import DataError
    path = str(tmp_path / "tr.tfrecords")
    rng = np.random.default_rng(2)
    with T.TFRecordWriter(path) as w:
        for r in range(64):
            ids = np.concatenate([np.arange(1, 14), rng.integers(20000, 90000, 26)])
            if r == 40:
                ids[22] = 99999
            w.write_example(float(r & 1), ids, np.ones(39, np.float32))
    ok = BatchLoader([path], 39, 16, feature_size=100000, ids_dtype=torch.int32, num_workers=2)
    assert sum(hb.count for hb in ok) == 64
    ok.close()
    bad = BatchLoader([path], 39, 16, feature_size=99999, ids_dtype=torch.int32, num_workers=2)
    with pytest.raises(DataError, match=r"feature id 99999 is outside \[0, feature_size = 99999\)"):
        for _ in bad:
            pass
    bad.close()


# ---------------------------------------------------------------- streams: blocks cut into records in place
def _all_batches(loader):
    out = []
    for hb in loader:
        out.append((hb.feat_ids.clone(), hb.feat_vals.clone(), hb.labels.clone()))
    loader.close()
    return out


def test_stream_blocks_deliver_what_the_mapped_reader_delivers(tmp_path):
    """A stream is read in 4 MiB blocks; records that straddle a block boundary are carried over.  10 MB of records,
    fed through a FIFO in writes of odd sizes, must come out exactly like the mapped file: every shard, two epochs
    of the file-backed stream."""
    from deepfm_b200.data.tfrecord import generate_synthetic
    path, fifo = str(tmp_path / "tr.tfrecords"), str(tmp_path / "training_0")
    generate_synthetic(path, 33000, 39, 117581, n_dense=13, seed=9, pos_rate=0.25, zipf_s=1.05)
    assert os.path.getsize(path) > 2 * (4 << 20)
    for shard_n, shard_i in ((1, 0), (3, 2)):
        kw = dict(shard_n=shard_n, shard_i=shard_i, ids_dtype=torch.int32, num_workers=3, drop_remainder=False,
                  feature_size=117581)
        want = _all_batches(BatchLoader([path], 39, 500, **kw))
        os.mkfifo(fifo)

        def feed():
            data, rng, pos = open(path, "rb").read(), np.random.default_rng(1), 0
            with open(fifo, "wb", buffering=0) as f:
                while pos < len(data):
                    n = int(rng.integers(1, 300000))
                    f.write(data[pos:pos + n])
                    pos += n

        t = threading.Thread(target=feed)
        t.start()
        got = _all_batches(BatchLoader([fifo], 39, 500, streaming=True, **kw))
        t.join()
        os.unlink(fifo)
        assert len(got) == len(want)
        for g, w in zip(got, want):
            assert all(torch.equal(a, b) for a, b in zip(g, w))
    # a regular file forced into stream mode, two epochs
    want2 = _all_batches(BatchLoader([path], 39, 512, num_epochs=2, ids_dtype=torch.int32))
    got2 = _all_batches(BatchLoader([path], 39, 512, num_epochs=2, ids_dtype=torch.int32, streaming=True))
    assert len(got2) == len(want2) == 2 * (33000 // 512)
    assert all(torch.equal(g[0], w[0]) and torch.equal(g[2], w[2]) for g, w in zip(got2, want2))


def test_stream_record_larger_than_a_block(tmp_path):
    """One Example of 700k fields is ~6 MB: larger than a stream block, so it gets a block of its own size."""
    F = 700_000
    path = str(tmp_path / "big.tfrecords")
    rng = np.random.default_rng(4)
    ids = rng.integers(0, 100000, F)
    vals = rng.random(F).astype(np.float32)
    with T.TFRecordWriter(path) as w:
        w.write_example(1.0, np.arange(F) % 7, np.ones(F, np.float32))
        w.write_example(0.0, ids, vals)
        w.write_example(1.0, np.arange(F) % 5, np.ones(F, np.float32))
    assert os.path.getsize(path) > 2 * (4 << 20) and len(T.encode_example(0.0, ids, vals)) > (4 << 20)
    got = _all_batches(BatchLoader([path], F, 1, streaming=True, ids_dtype=torch.int64, num_workers=2, num_slots=2))
    assert len(got) == 3
    assert (got[1][0][0].numpy() == ids).all() and (got[1][1][0].numpy() == vals).all() and got[1][2].item() == 0.0
    assert (got[2][0][0].numpy() == np.arange(F) % 5).all()


def test_stream_errors_name_the_offset(tmp_path):
    src = str(tmp_path / "src.tfrecords")
    with T.TFRecordWriter(src) as w:
        for i in range(50):
            w.write_example(float(i & 1), np.arange(4) + i, np.ones(4, np.float32))
    data = open(src, "rb").read()
    rec = len(data) // 50
    cases = {
        "payload": (bytearray(data), 20 * rec + 30, r"CRC mismatch.*byte offset %d\b" % (20 * rec + 12)),
        "length": (bytearray(data), 10 * rec + 2, r"corrupt length CRC at stream offset %d\b" % (10 * rec)),
    }
    for name, (buf, at, pattern) in cases.items():
        buf[at] ^= 0x40
        f = str(tmp_path / f"{name}.tfrecords")
        open(f, "wb").write(bytes(buf))
        with pytest.raises(T.DataError, match=pattern):
            list(BatchLoader([f], 4, 5, streaming=True))
    f = str(tmp_path / "cut.tfrecords")
    open(f, "wb").write(data[: 30 * rec + 7])
    with pytest.raises(T.DataError, match=r"stream ends inside a record \(stream offset %d\)" % (30 * rec)):
        list(BatchLoader([f], 4, 5, streaming=True))
