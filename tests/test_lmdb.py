"""Native LMDB reader (csrc/lmdb_reader.cc) against files produced by the in-tree writer: key order, leaf / branch /
overflow pages, Datum decoding into batch buffers, wrap-around, and the kLMDBData layer end to end (CPU)."""
import os

import numpy as np
import pytest

from singa_b200.io.lmdb_reader import LMDBEnv, LMDBSource, datum_bytes, write_lmdb


def _items(n, shape, seed=0):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        img = rng.integers(0, 256, size=shape, dtype=np.uint8)
        out.append((f"{i:08d}".encode(), img, int(rng.integers(0, 10))))
    return out


@pytest.mark.parametrize("n,shape", [(1, (1, 8, 8)), (37, (1, 28, 28)), (700, (1, 28, 28)), (25, (3, 64, 64)),
                                     (3000, (1, 4, 4))])
def test_roundtrip_key_order_and_values(tmp_path, n, shape):
    recs = _items(n, shape)
    shuffled = list(recs)
    np.random.default_rng(1).shuffle(shuffled)
    path = str(tmp_path / "db")
    assert write_lmdb(path, [(k, datum_bytes(img, lab)) for k, img, lab in shuffled]) == n
    env = LMDBEnv(path)
    assert env.count() == n
    assert env.first_shape() == list(shape)
    seen = []
    while True:
        kv = env.next()
        if kv is None:
            break
        seen.append(kv)
    assert [k for k, _ in seen] == [k for k, _, _ in recs]          # cursor order = key order
    for (k, v), (_, img, lab) in zip(seen, recs):
        assert v == datum_bytes(img, lab)
    assert env.next()[0] == recs[0][0]                               # restarts after the end
    env.close()


def test_batch_decode_wraps_around(tmp_path):
    n, shape = 10, (1, 28, 28)
    recs = _items(n, shape, seed=3)
    path = str(tmp_path / "db")
    write_lmdb(path, [(k, datum_bytes(img, lab)) for k, img, lab in recs])
    src = LMDBSource(path)
    assert src.sample_shape == [28, 28] and src.count() == n
    px = np.zeros((4, 28 * 28), dtype=np.uint8)
    lb = np.zeros(4, dtype=np.int32)
    pos = 0
    for _ in range(7):                                               # 28 records: wraps twice
        src.next_batch(px, lb)
        for i in range(4):
            _, img, lab = recs[(pos + i) % n]
            assert lb[i] == lab and np.array_equal(px[i], img.reshape(-1))
        pos += 4
    src.skip(3)
    src.next_batch(px, lb)
    assert lb[0] == recs[(pos + 3) % n][2]


def test_corrupt_file_is_rejected(tmp_path):
    p = tmp_path / "bad"
    os.makedirs(p)
    (p / "data.mdb").write_bytes(b"\0" * 8192)
    with pytest.raises(RuntimeError):
        LMDBEnv(str(p))


def test_lmdb_data_layer_trains(tmp_path):
    """kLMDBData -> kMnistImage/kLabel -> MLP: a few steps through the worker on the reference path."""
    from singa_b200.models import configs
    from singa_b200.worker import Worker
    recs = _items(64, (1, 28, 28), seed=5)
    path = str(tmp_path / "mnist_lmdb")
    write_lmdb(path, [(k, datum_bytes(img, lab)) for k, img, lab in recs])
    from singa_b200 import device
    device.set_context("cpu", "torch")
    model = configs.mnist_mlp(batch=16, train_steps=4, hidden=(32, 16))
    for l in model.neuralnet.layer:
        if l.type in ("kShardData", "kLMDBData"):
            l.type = "kLMDBData"
            l.data_param.path = path
            l.data_param.ClearField("source")
    w = Worker()
    w.build(model)
    d = w.train_net.datalayers[0]
    assert d.sample_shape == [28, 28] and d.source.count() == 64
    w.executor.setup()
    for step in range(4):
        w.executor.ticks += 1
        w.executor.train_one_batch(step)
    assert float(w.train_net.losslayers[0].metric[0]) > 0


def test_loader_tool_writes_lmdb(tmp_path):
    from singa_b200.tools import loader
    out = str(tmp_path / "syn_lmdb")
    assert loader.main(["--datasource", "synthetic", "--shape", "3,8,8", "--n", "50", "--shard_folder", out,
                        "--format", "lmdb"]) == 0
    src = LMDBSource(out)
    assert src.count() == 50 and src.sample_shape == [3, 8, 8]
    px = np.zeros((5, 192), dtype=np.uint8)
    lb = np.zeros(5, dtype=np.int32)
    src.next_batch(px, lb)
    assert px.any() and (lb >= 0).all() and (lb < 10).all()


def test_encoded_datums_are_decoded_on_the_host(tmp_path):
    """Datum.encoded = true (convert_imageset --encoded): the native batch path refuses, the source falls back to an
    image decoder; PNG is lossless so the pixels must come back exactly (BGR channel order)."""
    cv2 = pytest.importorskip("cv2")
    from singa_b200.proto import schema as S
    rng = np.random.default_rng(7)
    imgs = [rng.integers(0, 256, size=(3, 10, 12), dtype=np.uint8) for _ in range(5)]
    items = []
    for i, im in enumerate(imgs):
        ok, buf = cv2.imencode(".png", np.ascontiguousarray(im.transpose(1, 2, 0)))
        assert ok
        d = S.Datum()
        d.data, d.label, d.encoded = buf.tobytes(), i % 3, True
        items.append((f"{i:08d}".encode(), d.SerializeToString()))
    path = str(tmp_path / "enc")
    write_lmdb(path, items)
    src = LMDBSource(path)
    assert src.encoded and src.sample_shape == [3, 10, 12]
    px = np.zeros((7, 360), dtype=np.uint8)
    lb = np.zeros(7, dtype=np.int32)
    src.next_batch(px, lb)                                           # 7 > 5 records: wraps around
    for i in range(7):
        assert lb[i] == (i % 5) % 3 and np.array_equal(px[i], imgs[i % 5].reshape(-1))


def test_value_sizes_around_the_overflow_threshold_and_deep_trees(tmp_path):
    """Leaf nodes larger than nodemax (2038 bytes at 4 KB pages) move their value to overflow pages; 20 000 small
    records need a three-level tree.  Raw values (not Datums) are read back through the cursor."""
    rng = np.random.default_rng(11)
    items = []
    for i, n in enumerate(list(range(2000, 2060)) + [1, 4079, 4080, 4081, 8176, 20000]):
        items.append((f"a{i:06d}".encode(), rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()))
    for i in range(20000):
        items.append((f"b{i:07d}".encode(), bytes([i % 251]) * (1 + i % 40)))
    path = str(tmp_path / "db")
    write_lmdb(path, items)
    env = LMDBEnv(path)
    assert env.count() == len(items)
    expect = sorted(items)
    for k, v in expect:
        kv = env.next()
        assert kv is not None and kv[0] == k and kv[1] == v, k
    assert env.next() is None


def test_float_datums_keep_negative_and_fractional_values(tmp_path):
    """Datum.float_data (mean-subtracted images, the reference's own LMDB conversion) is staged as float32 unchanged;
    a uint8 batch refuses it instead of clamping."""
    vals = [np.array([[[-3.5, 10.25], [300.0, 0.75]]], dtype=np.float32),
            np.array([[[1e-3, -255.0], [127.5, 2.0]]], dtype=np.float32)]
    path = str(tmp_path / "fdb")
    write_lmdb(path, [(b"k%d" % i, datum_bytes(v, i + 5)) for i, v in enumerate(vals)])
    src = LMDBSource(path)
    assert src.is_float and src.sample_shape == [2, 2]
    px = np.zeros((3, 4), dtype=np.float32)
    lb = np.zeros(3, dtype=np.int32)
    src.next_batch(px, lb)
    assert lb.tolist() == [5, 6, 5]
    assert np.array_equal(px[0], vals[0].reshape(-1)) and np.array_equal(px[1], vals[1].reshape(-1))
    with pytest.raises(RuntimeError, match="float_data"):
        src.next_batch(np.zeros((2, 4), dtype=np.uint8), lb[:2].copy())
