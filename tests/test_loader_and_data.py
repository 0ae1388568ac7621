"""loader tool -> Shard -> ShardDataLayer -> training on CPU, and split / partition helpers."""
import gzip
import os
import struct

import numpy as np

from singa_b200.io.shard import Shard
from singa_b200.tools import loader


def _write_idx(tmp_path, n=40):
    rng = np.random.default_rng(1)
    imgs = rng.integers(0, 256, size=(n, 28, 28), dtype=np.uint8)
    labels = rng.integers(0, 10, size=n, dtype=np.uint8)
    ip, lp = tmp_path / "images.idx3", tmp_path / "labels.idx1"
    ip.write_bytes(struct.pack(">IIII", 2051, n, 28, 28) + imgs.tobytes())
    lp.write_bytes(struct.pack(">II", 2049, n) + labels.tobytes())
    return str(ip), str(lp), imgs, labels


def test_mnist_loader_restartable_and_split(tmp_path):
    ip, lp, imgs, labels = _write_idx(tmp_path)
    folder = str(tmp_path / "train")
    assert loader.build(loader.MnistSource(ip, lp), folder) == 40
    assert loader.build(loader.MnistSource(ip, lp), folder) == 0          # restart: every key already present
    with Shard(folder, Shard.kRead) as sh:
        assert sh.count() == 40
    counts = loader.split(folder, str(tmp_path / "part"), 3)
    assert sum(counts) == 40 and max(counts) - min(counts) <= 1


def test_partition_record_list(tmp_path):
    rid = tmp_path / "rid.txt"
    rid.write_text("".join(f"img{i}.jpg {i % 7}\n" for i in range(10)))
    parts = loader.partition_record_list(str(rid), 3)
    assert [len(p) for p in parts] == [4, 4, 2]
    parts2 = loader.partition_record_list(str(rid), 3, replicate=2)
    assert sum(len(p) for p in parts2) == 20


def test_training_from_a_real_shard_on_cpu(tmp_path):
    from singa_b200 import device
    from singa_b200.models import configs
    from singa_b200.worker import Worker
    ip, lp, imgs, labels = _write_idx(tmp_path, n=64)
    folder = str(tmp_path / "train")
    loader.build(loader.MnistSource(ip, lp), folder)
    device.set_context("cpu", "torch")
    m = configs.mnist_mlp(batch=16, train_steps=6, hidden=(32,))
    m.display_frequency = 0
    dl = m.neuralnet.layer[0]
    dl.data_param.ClearField("source")
    dl.data_param.path = folder
    w = Worker()
    w.build(m)
    assert w.train_net.datalayers[0].sample_shape == [28, 28]
    w.executor.setup()
    for step in range(6):
        w.executor.ticks += 1
        w.executor.train_one_batch(step)
    d = w.train_net.datalayers[0]
    assert float(w.train_net.losslayers[0].metric[0]) > 0
    # the first batch the data layer read are the first 16 records of the shard
    with Shard(folder, Shard.kRead) as sh:
        first = sh.next()
    assert first[0] == b"00000000"
    device.reset_context()


def test_mnist_distortion_is_applied_in_training_only():
    """MnistProto gamma / beta / elastic fields change training batches (the reference parses and ignores them) and
    leave test batches alone; identity parameters reproduce the plain normalisation."""
    import torch
    from singa_b200 import device
    from singa_b200.neuralnet import NeuralNet
    from singa_b200.ops import reference as R
    from singa_b200.proto import schema as S
    device.set_context("cpu", "torch")
    p = S.NetProto()
    l = p.layer.add(); l.name, l.type = "data", "kShardData"
    l.data_param.batchsize = 6; l.data_param.source = "synthetic"; l.data_param.synthetic_shape.extend([12, 12])
    l = p.layer.add(); l.name, l.type = "mnist", "kMnistImage"; l.srclayers.append("data")
    l.mnist_param.norm_a, l.mnist_param.norm_b = 127.5, 1.0
    l.mnist_param.gamma, l.mnist_param.beta = 15.0, 15.0
    l.mnist_param.kernel, l.mnist_param.sigma, l.mnist_param.alpha = 5, 2.0, 3.0
    net = NeuralNet(p)
    d, m = net.name2layer("data"), net.name2layer("mnist")
    d.compute_feature(True)
    plain = d.pixels.float().view(6, 12, 12) / 127.5 - 1.0
    m.compute_feature(False)
    assert torch.allclose(m.data, plain)                     # test phase: no distortion
    m.stage_ready = -1
    m.compute_feature(True)
    assert not torch.allclose(m.data, plain)                 # training phase: distorted
    assert float(m.data.min()) >= -1.0 - 1e-5 and float(m.data.max()) <= 1.0 + 1e-5
    A, disp = m._last_distortion
    assert A.shape == (6, 6) and disp.shape == (6, 2, 12, 12) and float(disp.abs().max()) < 3.0
    # identity map reproduces the normalisation
    eye = torch.tensor([[1., 0., 0., 0., 1., 0.]]).repeat(6, 1)
    out = R.mnist_distort(d.pixels.view(6, 12, 12), eye, None, 12, 127.5, 1.0)
    assert torch.allclose(out, plain, atol=1e-6)
    device.reset_context()
