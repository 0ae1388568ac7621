"""A partitioned net must compute what the unpartitioned net computes (the reference only ever BUILT partitioned
graphs: its Slice/Concate/Split bodies are empty and `singa` never passes group_size > 1)."""
import pytest
import torch

from singa_b200 import device
from singa_b200.neuralnet import NeuralNet
from singa_b200.proto import schema as S
from test_config_and_graph import convnet


def _run(net, steps=1):
    for d in net.datalayers:
        d.compute_feature(True)
    from singa_b200.parallel.bridge import BridgeTransport
    bridge = BridgeTransport(None)          # every location lives in this process: bridges are device copies
    for l in net.layers:
        if not l.is_datalayer:
            if l.is_bridgedstlayer:
                bridge.recv_data(l)
            l.compute_feature(True)
            if l.is_bridgesrclayer:
                bridge.send_data(l)
    net.zero_shared_grads()
    for l in reversed(net.layers):
        if not (l.is_datalayer or l.is_parserlayer):
            if l.is_bridgesrclayer:
                bridge.recv_grad(l)
            l.compute_gradient()
            if l.is_bridgedstlayer:
                bridge.send_grad(l)
    return net


def _init(net, seed=3):
    for p in net.params:
        if p.is_owner:
            p.init(seed)
    return net


@pytest.mark.parametrize("ptype,gsize", [(S.kDataPartition, 2), (S.kLayerPartition, 2)])
def test_partitioned_net_matches_unpartitioned(ptype, gsize):
    device.set_context("cpu", "torch")
    torch.manual_seed(0)
    base = _init(NeuralNet(convnet(), fuse=False))
    # give the parameters non-trivial values
    g = torch.Generator().manual_seed(5)
    for p in base.params:
        p.data.copy_(torch.randn(p.shape, generator=g) * 0.1)
    p2 = convnet()
    p2.partition_type = ptype
    part = NeuralNet(p2, gsize, fuse=False)
    # copy parameters: data partition shares them, layer partition slices dim 1 of conv / ip weights
    for name in ("conv1", "fc1"):
        bw, bb = base.name2layer(name).get_params()
        if ptype == S.kDataPartition:
            w, b = part.name2layer(f"{name}-00").get_params()
            w.data.copy_(bw.data); b.data.copy_(bb.data)
        else:
            off = 0
            for i in range(gsize):
                w, b = part.name2layer(f"{name}-{i:02d}").get_params()
                n = b.shape[0]
                if name == "conv1":
                    w.data.copy_(bw.data[off:off + n]); b.data.copy_(bb.data[off:off + n])
                else:
                    # fc1 splits its OUTPUT dim; its input is the concatenation of the conv partitions
                    w.data.copy_(bw.data[:, off:off + n]); b.data.copy_(bb.data[off:off + n])
                off += n
    _run(base)
    _run(part)
    loss_b = float(base.losslayers[0].metric[0])
    if ptype == S.kDataPartition:
        loss_p = sum(float(l.metric[0]) for l in part.losslayers) / len(part.losslayers)
        assert loss_p == pytest.approx(loss_b, rel=1e-4)
        # shared, accumulated gradient = gsize x the full-batch mean gradient
        wg = part.name2layer("fc1-00").get_params()[0].grad
        torch.testing.assert_close(wg / gsize, base.name2layer("fc1").get_params()[0].grad, rtol=1e-3, atol=1e-5)
        assert part.name2layer("fc1-01").get_params()[0].data.data_ptr() == \
            part.name2layer("fc1-00").get_params()[0].data.data_ptr()
    else:
        loss_p = float(part.losslayers[0].metric[0])
        assert loss_p == pytest.approx(loss_b, rel=1e-4)
        off = 0
        for i in range(gsize):
            w = part.name2layer(f"fc1-{i:02d}").get_params()[0]
            n = w.shape[1]
            torch.testing.assert_close(w.grad, base.name2layer("fc1").get_params()[0].grad[:, off:off + n],
                                       rtol=1e-3, atol=1e-5)
            off += n
    device.reset_context()
