"""Shard record file — the portable spec of src/test/test_shard.cc:12-57, plus crash recovery and the native
batch decoder."""
import os
import struct

import numpy as np

from singa_b200.io.shard import Shard, read_image_batch
from singa_b200.proto import schema as S

KEYS = [f"firstkey{i}" for i in range(5)]
TUPLES = [f"tuple number {i} with some payload" * (i + 1) for i in range(5)]


def test_create_append_count_read(tmp_path):
    folder = str(tmp_path / "shard_test")
    with Shard(folder, Shard.kCreate, 50) as sh:
        for i in range(3):
            assert sh.insert(KEYS[i], TUPLES[i].encode())
        sh.flush()
    with Shard(folder, Shard.kAppend, 50) as sh:
        assert sh.insert(KEYS[3], TUPLES[3].encode())
        assert sh.insert(KEYS[4], TUPLES[4].encode())
        assert not sh.insert(KEYS[0], b"duplicate key is ignored")
        sh.flush()
    with Shard(folder, Shard.kRead, 50) as sh:
        assert sh.count() == 5
        for i in range(5):
            k, v = sh.next()
            assert k.decode() == KEYS[i] and v.decode() == TUPLES[i]
        assert sh.next() is None
        sh.seek_to_first()
        k, v = sh.next()
        assert k.decode() == KEYS[0] and v.decode() == TUPLES[0]


def test_layout_is_u64_lengths(tmp_path):
    folder = str(tmp_path / "s")
    with Shard(folder, Shard.kCreate) as sh:
        sh.insert("ab", b"xyz")
        sh.flush()
    raw = open(os.path.join(folder, "shard.dat"), "rb").read()
    assert raw == struct.pack("<Q", 2) + b"ab" + struct.pack("<Q", 3) + b"xyz"


def test_append_recovers_from_torn_tail(tmp_path):
    folder = str(tmp_path / "s")
    with Shard(folder, Shard.kCreate) as sh:
        sh.insert("k0", b"v0")
        sh.insert("k1", b"v1" * 100)
        sh.flush()
    path = os.path.join(folder, "shard.dat")
    size = os.path.getsize(path)
    with open(path, "r+b") as f:          # a crashed writer: the last tuple is incomplete
        f.truncate(size - 17)
    with Shard(folder, Shard.kAppend) as sh:
        assert sh.insert("k1", b"again")  # k1 was torn, so it is not a duplicate
        assert not sh.insert("k0", b"dup")
        sh.flush()
    with Shard(folder, Shard.kRead) as sh:
        assert [sh.next()[0] for _ in range(2)] == [b"k0", b"k1"]
        assert sh.next() is None


def test_native_image_batch_decoder_wraps_around(tmp_path):
    folder = str(tmp_path / "img")
    rng = np.random.default_rng(0)
    imgs = rng.integers(0, 256, size=(5, 28 * 28), dtype=np.uint8)
    with Shard(folder, Shard.kCreate) as sh:
        for i in range(5):
            r = S.Record()
            r.image.shape.extend([28, 28])
            r.image.label = i + 3
            if i % 2 == 0:
                r.image.pixel = imgs[i].tobytes()
            else:
                r.image.data.extend(imgs[i].astype(np.float32).tolist())
            sh.insert(f"img{i}", r)
        sh.flush()
    with Shard(folder, Shard.kRead) as sh:
        # byte- and float-valued records mixed: an fp32 batch takes both
        px = np.zeros((7, 784), dtype=np.float32)
        lab = np.zeros(7, dtype=np.int32)
        read_image_batch(sh, px, lab)
        assert lab.tolist() == [3, 4, 5, 6, 7, 3, 4]
        assert np.array_equal(px[:5], imgs.astype(np.float32)) and np.array_equal(px[5:], imgs[:2].astype(np.float32))


def test_float_records_keep_negative_and_fractional_values(tmp_path):
    """Mean-subtracted images (the loader tool's ImageNet path) are float records with negative and fractional values:
    they must reach the batch unchanged (src/worker/layer.cc:595-598 memcpy's them), never clamped into bytes."""
    import pytest
    folder = str(tmp_path / "flt")
    vals = np.array([[-3.5, 10.25, 300.0, 0.75], [1e-3, -255.0, 127.5, 2.0]], dtype=np.float32)
    with Shard(folder, Shard.kCreate) as sh:
        for i, v in enumerate(vals):
            r = S.Record()
            r.image.shape.extend([2, 2])
            r.image.label = i
            r.image.data.extend(v.tolist())
            sh.insert(f"f{i}", r)
        sh.flush()
    with Shard(folder, Shard.kRead) as sh:
        px = np.zeros((2, 4), dtype=np.float32)
        lab = np.zeros(2, dtype=np.int32)
        read_image_batch(sh, px, lab)
        assert np.array_equal(px, vals) and lab.tolist() == [0, 1]
        sh.seek_to_first()
        with pytest.raises(IOError, match="float pixels"):          # a uint8 batch refuses instead of clamping
            read_image_batch(sh, np.zeros((2, 4), dtype=np.uint8), lab)
    # the data layer stages such a shard as float32 and the parser passes the values through
    from singa_b200.layers.data import ShardSource
    src = ShardSource(folder)
    assert src.is_float and src.sample_shape == [2, 2]
