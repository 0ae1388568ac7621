"""Whole-net finite-difference gradient check on the CPU oracle path (the reference has no numerical
gradient checks).  A small conv - pool - LRN - relu - inner product - tanh - inner product - softmax-loss net is run
through NeuralNet / the layers' compute_feature + compute_gradient, and dLoss/dw of randomly chosen entries of every
parameter is compared with a central difference of the loss."""
import numpy as np
import pytest
import torch

from singa_b200 import device
from singa_b200.models import configs as C
from singa_b200.proto import schema as S
from singa_b200.worker import Worker


def _tiny_model(pool_method):
    m = S.ModelProto()
    m.name = "gradcheck"
    m.train_steps = 1
    m.display_frequency = 0
    C._sgd(m, 0.01)
    net = m.neuralnet
    C._data(net, 6, [3, 12, 12], 5)
    l = net.layer.add(); l.name, l.type = "rgb", "kRGBImage"; l.srclayers.append("data")
    l.rgbimage_param.scale = 1.0 / 64
    l = net.layer.add(); l.name, l.type = "label", "kLabel"; l.srclayers.append("data")
    C._conv(net, "conv1", "rgb", 4, 3, stride=1, pad=1, std=0.2, bias=0.1)
    C._pool(net, "pool1", "conv1", 3, 2, method=pool_method)
    C._lrn(net, "norm1", "pool1", size=3, alpha=0.05, beta=0.75)
    C._act(net, "relu1", "norm1", "kReLU")
    C._ip(net, "fc1", "relu1", 7, low=-0.3, high=0.3)
    C._act(net, "tanh1", "fc1", "kTanh")
    C._ip(net, "fc2", "tanh1", 5, low=-0.3, high=0.3)
    C._loss(net, "fc2")
    return m


def _loss(net):
    """Mean softmax cross-entropy of one forward pass, evaluated in float64 from the last layer's logits (the loss
    layer's own metric blob is fp32 and accumulates over a display window)."""
    for l in net.layers:
        if not l.is_datalayer and not l.is_parserlayer:
            l.compute_feature(True)
    loss_layer = net.losslayers[0]
    logits = loss_layer.srclayers[0].data_for(loss_layer).double().reshape(loss_layer.srclayers[0].data.shape[0], -1)
    labels = loss_layer.srclayers[1].data_for(loss_layer).long().reshape(-1)
    return float(torch.nn.functional.cross_entropy(logits, labels))


@pytest.mark.parametrize("pool_method", ["MAX", "AVE"])
def test_backward_matches_central_differences(pool_method, monkeypatch):
    # the oracle path casts its inputs with .float(); for this test "float" means float64
    monkeypatch.setattr(torch.Tensor, "float", lambda t: t.double())
    device.set_context("cpu", "torch")
    torch.manual_seed(0)
    w = Worker(seed=1)
    w.build(_tiny_model(pool_method))
    net = w.train_net
    for d in net.datalayers:
        d.compute_feature(True)
    for p in net.parserlayers:
        p.compute_feature(True)
    # float64 everywhere below the parsers so the central difference is meaningful
    converted = {}                      # fused activations alias their producer's blobs: keep the aliasing
    for l in net.layers:
        for name in ("data", "grad"):
            t = getattr(l, name, None)
            if isinstance(t, torch.Tensor) and t.dtype == torch.float32:
                key = (t.data_ptr(), tuple(t.shape), tuple(t.stride()))
                if key not in converted:
                    converted[key] = t.double()
                setattr(l, name, converted[key])
    for p in net.params:
        p.data = p.data.double()
        p.grad = p.grad.double()
    base = _loss(net)
    assert np.isfinite(base)
    for l in reversed(net.layers):
        if not l.is_datalayer and not l.is_parserlayer:
            l.compute_gradient()
    rng = np.random.default_rng(0)
    eps = 1e-5
    checked = 0
    for p in net.params:
        flat, g = p.data.view(-1), p.grad.view(-1)
        for i in rng.choice(flat.numel(), size=min(6, flat.numel()), replace=False):
            old = float(flat[i])
            flat[i] = old + eps
            lp = _loss(net)
            flat[i] = old - eps
            lm = _loss(net)
            flat[i] = old
            num = (lp - lm) / (2 * eps)
            assert abs(num - float(g[i])) <= 1e-6 + 1e-4 * max(abs(num), abs(float(g[i]))), \
                (p.name, int(i), num, float(g[i]))
            checked += 1
    assert checked >= 20
    device.reset_context()
