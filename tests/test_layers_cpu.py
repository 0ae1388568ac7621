"""CPU oracle checks: every layer's backward against autograd / finite differences of its own forward, the
updaters against hand-written recurrences, LR schedules, init statistics (the kept parts of the reference's legacy
test_param.cc:25-126), sync rules and the checkpoint round trip."""
import math
import os

import numpy as np
import pytest
import torch

from singa_b200.ops import reference as R
from singa_b200.proto import schema as S

torch.manual_seed(0)


# ------------------------------------------------------------------------------------------------ layer math
def test_stanh_constants_and_gradient():
    x = torch.linspace(-3, 3, 41, dtype=torch.float64, requires_grad=True)
    y = R.stanh_fwd(x)
    assert abs(float(R.stanh_fwd(torch.tensor(1e9)))) == pytest.approx(1.7159047, rel=1e-6)
    y.sum().backward()
    torch.testing.assert_close(R.stanh_bwd(torch.ones_like(y), y.detach()), x.grad, rtol=1e-6, atol=1e-8)


@pytest.mark.parametrize("is_max", [True, False])
@pytest.mark.parametrize("h,k,s,pad", [(8, 2, 2, 0), (9, 4, 2, 0), (32, 3, 2, 0), (13, 3, 2, 1)])
def test_pool_backward_matches_autograd(h, k, s, pad, is_max):
    x = torch.randn(2, 3, h, h, dtype=torch.float64, requires_grad=True)
    y = R.pool_fwd(x, k, s, pad, is_max)
    assert y.shape[-1] == R.pool_out_size(h, k, s, pad)
    dy = torch.randn_like(y)
    (y * dy).sum().backward()
    dx = R.pool_bwd(x.detach(), y.detach(), dy, k, s, pad, is_max)
    torch.testing.assert_close(dx, x.grad, rtol=1e-9, atol=1e-9)    # no ties in random data


def test_pool_max_ties_route_to_every_maximum():
    x = torch.ones(1, 1, 2, 2)
    y = R.pool_fwd(x, 2, 2, 0, True)
    dx = R.pool_bwd(x, y, torch.full_like(y, 3.0), 2, 2, 0, True)
    assert torch.equal(dx, torch.full((1, 1, 2, 2), 3.0))            # mshadow unpool semantics


def test_pool_ceil_output_and_ave_divides_by_k2():
    assert R.pool_out_size(32, 3, 2) == 16 and R.pool_out_size(9, 4, 2) == 4 and R.pool_out_size(24, 2, 2) == 12
    x = torch.ones(1, 1, 4, 4)
    y = R.pool_fwd(x, 3, 2, 0, False)        # second window is clipped to 3x2... still divided by 9
    assert y.shape == (1, 1, 2, 2) and float(y[0, 0, 0, 0]) == 1.0 and float(y[0, 0, 1, 1]) == pytest.approx(4 / 9)


def test_conv_layout_and_gradients():
    n, c, h, k, f = 2, 3, 7, 3, 4
    x = torch.randn(n, c, h, h, dtype=torch.float64, requires_grad=True)
    w = torch.randn(f, c * k * k, dtype=torch.float64, requires_grad=True)
    b = torch.randn(f, dtype=torch.float64, requires_grad=True)
    y = R.conv_fwd(x, w, b, k, 2, 1)
    # column index (c*k + ky)*k + kx: compare one output element by hand
    xp = torch.nn.functional.pad(x.detach(), (1, 1, 1, 1))
    patch = xp[0, :, 2:2 + k, 4:4 + k].reshape(-1)
    assert float(y[0, 1, 1, 2]) == pytest.approx(float(patch @ w[1].detach() + b[1].detach()), rel=1e-9)
    dy = torch.randn_like(y)
    (y * dy).sum().backward()
    dx, dw, db = R.conv_bwd(x.detach(), w.detach(), dy, k, 2, 1)
    torch.testing.assert_close(dx, x.grad)
    torch.testing.assert_close(dw, w.grad)
    torch.testing.assert_close(db, b.grad)


def test_inner_product_gradients():
    x = torch.randn(5, 7, dtype=torch.float64, requires_grad=True)
    w = torch.randn(7, 3, dtype=torch.float64, requires_grad=True)
    b = torch.randn(3, dtype=torch.float64, requires_grad=True)
    y = R.ip_fwd(x, w, b)
    dy = torch.randn_like(y)
    (y * dy).sum().backward()
    dx, dw, db = R.ip_bwd(x.detach(), w.detach(), dy)
    torch.testing.assert_close(dx, x.grad); torch.testing.assert_close(dw, w.grad); torch.testing.assert_close(db, b.grad)


def test_softmax_loss_metrics_and_gradient():
    logits = torch.randn(6, 5, dtype=torch.float64, requires_grad=True)
    labels = torch.tensor([0, 1, 2, 3, 4, 0])
    m, g = R.softmax_loss(logits.detach().float(), labels, topk=2, scale=0.5)
    loss = torch.nn.functional.cross_entropy(logits, labels) * 0.5
    loss.backward()
    assert float(m[0]) == pytest.approx(float(loss), rel=1e-5)
    torch.testing.assert_close(g.double(), logits.grad, rtol=1e-5, atol=1e-7)
    top2 = logits.detach().topk(2, dim=1).indices
    acc = float((top2 == labels[:, None]).any(1).float().mean()) * 0.5
    assert float(m[1]) == pytest.approx(acc)
    # labels >= 10 are fine (the reference CHECKs label < 10)
    R.softmax_loss(torch.randn(3, 1000), torch.tensor([999, 500, 10]))


def test_lrn_window_is_clipped_at_the_channel_edges():
    x = torch.ones(1, 4, 1, 1)
    y, norm = R.lrn_fwd(x, 3, alpha=3.0, beta=1.0, knorm=1.0)
    assert norm.flatten().tolist() == [3.0, 4.0, 4.0, 3.0]           # 1 + (3/3) * #in-window


def test_philox_reference_is_uniform_and_reproducible():
    u = R.philox_uniform(12345, 100000)
    assert torch.equal(u, R.philox_uniform(12345, 100000)) and not torch.equal(u, R.philox_uniform(12346, 100000))
    assert float(u.min()) >= 0 and float(u.max()) < 1 and abs(float(u.mean()) - 0.5) < 0.01
    assert abs(float(R.dropout_mask(7, 100000, 0.8).mean()) - 1.0) < 0.02


# ------------------------------------------------------------------------------------------------ updaters
def _up(**kw):
    u = S.UpdaterProto()
    u.base_learning_rate = 0.1
    for k, v in kw.items():
        setattr(u, k, v if not isinstance(v, str) or k == "param_type" else
                S.enum_value("UpdaterProto.Type" if k == "type" else "UpdaterProto.ChangeProto", v))
    return u


def test_learning_rate_schedules():
    assert R.learning_rate(_up(), 7) == pytest.approx(0.1)
    u = _up(learning_rate_change_method="kLinear", learning_rate_change_frequency=10, final_learning_rate=0.0)
    assert R.learning_rate(u, 5) == pytest.approx(0.05)
    u = _up(learning_rate_change_method="kExponential", learning_rate_change_frequency=10, final_learning_rate=0.05)
    assert R.learning_rate(u, 20) == pytest.approx(0.025)
    u = _up(learning_rate_change_method="kInverse_t", final_learning_rate=0.05)
    assert R.learning_rate(u, 1) == pytest.approx(0.1 / (1 + 1 / 0.05))
    u = _up(learning_rate_change_method="kInverse", gamma=0.0001, pow=0.75)
    assert R.learning_rate(u, 100) == pytest.approx(0.1 * (1 + 0.0001 * 100) ** -0.75)
    u = _up(learning_rate_change_method="kStep", gamma=0.997, learning_rate_change_frequency=60)
    assert R.learning_rate(u, 119) == pytest.approx(0.1 * 0.997) and R.learning_rate(u, 120) == pytest.approx(0.1 * 0.997 ** 2)


@pytest.mark.parametrize("kind", ["kSGD", "kNesterov", "kAdaGrad", "kRMSProp", "kAdaDelta"])
def test_updaters_match_scalar_recurrences(kind):
    from singa_b200.param import Param
    from singa_b200.updater import create_updater
    u = _up(type=kind, momentum=0.9, weight_decay=0.01, delta=1e-7, rho=0.9)
    upd = create_updater(u)
    pp = S.ParamProto(); pp.name = "w"
    p = Param(); p.setup(pp, [3], 0)
    p.data.copy_(torch.tensor([1.0, -2.0, 0.5]))
    w, h, q = [1.0, -2.0, 0.5], [0.0] * 3, [0.0] * 3
    for step in range(3):
        g = [0.3 * (step + 1), -0.2, 0.1]
        p.grad.copy_(torch.tensor(g))
        upd.update(step, p)
        for i in range(3):
            lr, wd, mu, rho, d = 0.1, 0.01, 0.9, 0.9, 1e-7
            if kind == "kSGD":
                gg = g[i] + wd * w[i]; h[i] = mu * h[i] + lr * gg; w[i] -= h[i]
            elif kind == "kNesterov":
                gg = g[i] + wd * w[i]; old = h[i]; h[i] = mu * h[i] + lr * gg; w[i] -= (1 + mu) * h[i] - mu * old
            elif kind == "kAdaGrad":
                h[i] += g[i] ** 2; gg = g[i] + wd * w[i]; w[i] -= lr * gg / math.sqrt(h[i] + d)
            elif kind == "kRMSProp":
                h[i] = rho * h[i] + (1 - rho) * g[i] ** 2; gg = g[i] + wd * w[i]; w[i] -= lr * gg / math.sqrt(h[i] + d)
            else:
                gg = g[i] + wd * w[i]; h[i] = rho * h[i] + (1 - rho) * gg ** 2
                t = gg * math.sqrt(q[i] + d) / math.sqrt(h[i] + d); q[i] = rho * q[i] + (1 - rho) * t * t; w[i] -= t
    assert p.data.tolist() == pytest.approx(w, rel=1e-5)


# ------------------------------------------------------------------------------------------------ params
def _param(method, shape, fan_in=0, **kw):
    from singa_b200.param import Param
    pp = S.ParamProto(); pp.name = "p"
    pp.init_method = S.enum_value("ParamProto.InitMethod", method)
    for k, v in kw.items():
        setattr(pp, k, v)
    p = Param(); p.setup(pp, shape, fan_in); p.id = 3
    p.init(seed=1)
    return p


def test_param_init_distributions():
    assert torch.equal(_param("kConstant", [4, 5], value=0.25).data, torch.full((4, 5), 0.25))
    u = _param("kUniform", [100, 50], low=-0.05, high=0.05).data
    assert float(u.min()) >= -0.05 and float(u.max()) <= 0.05 and abs(float(u.mean())) < 2e-3
    u = _param("kUniform", [100, 50], low=-1, high=1, value=3.0).data
    assert 2.5 < float(u.max()) <= 3.0
    u = _param("kUniformSqrtFanInOut", [60, 40], low=-1, high=1, value=6.0).data
    assert float(u.abs().max()) <= 6.0 / math.sqrt(100) + 1e-6
    u = _param("kUniformSqrtFanIn", [60, 40], fan_in=300, low=-1, high=1).data
    assert float(u.abs().max()) <= 1.0 / math.sqrt(300 / 3.0) + 1e-6
    g = _param("kGaussain", [100, 50], mean=0.5, std=2.0).data
    assert abs(float(g.mean()) - 0.5) < 0.1 and abs(float(g.std()) - 2.0) < 0.1
    g = _param("kGaussainSqrtFanIn", [100, 50], mean=0.0, std=1.0).data
    assert abs(float(g.std()) - 1.0 / math.sqrt(100)) < 0.01
    assert torch.equal(_param("kGaussain", [10, 5]).data, _param("kGaussain", [10, 5]).data)   # seeded


def test_elastic_and_randomsync_message_rules():
    from singa_b200.param import ElasticParam, RandomSyncParam
    pp = S.ParamProto(); pp.name = "w"
    w, s = ElasticParam(), ElasticParam()
    for p in (w, s):
        p.setup(pp, [6], 0)
    w.data.copy_(torch.arange(6.0)); s.data.copy_(torch.ones(6))
    control, payload = w.gen_sync_msg_from_worker(0.25)
    control2, reply = s.handle_sync_msg(control, payload)
    w.parse_sync_msg_from_ps(control2, reply)
    d = (torch.arange(6.0) - 1.0) * 0.25
    torch.testing.assert_close(s.data, torch.ones(6) + d)
    torch.testing.assert_close(w.data, torch.arange(6.0) - d)
    rw, rs = RandomSyncParam(), RandomSyncParam()
    for p in (rw, rs):
        p.setup(pp, [1000], 0)
    rw.data.copy_(torch.randn(1000)); rw.snapshot.copy_(rw.data - 0.1)
    rs.data.copy_(torch.randn(1000))
    w0, s0, sn0 = rw.data.clone(), rs.data.clone(), rw.snapshot.clone()
    control, payload = rw.gen_sync_msg_from_worker(0.3)
    seed, ratio, count = control.split("-")
    assert 200 < int(count) < 400
    control2, reply = rs.handle_sync_msg(control, payload)
    rw.parse_sync_msg_from_ps(control2, reply)
    idx = RandomSyncParam.random_sample(int(seed), float(ratio), 1000)
    mask = torch.zeros(1000, dtype=torch.bool); mask[idx] = True
    torch.testing.assert_close(rs.data[mask], s0[mask] + 0.1)                     # server += w - snap
    torch.testing.assert_close(rw.data[mask], w0[mask] + s0[mask] - sn0[mask])    # w += old - snap
    assert torch.equal(rw.data[~mask], w0[~mask]) and torch.equal(rw.snapshot[mask], rw.data[mask])


def test_checkpoint_roundtrip_with_optimizer_state(tmp_path):
    from singa_b200 import checkpoint, device
    from singa_b200.models import configs
    from singa_b200.worker import Worker
    device.set_context("cpu", "torch")
    m = configs.mnist_mlp(batch=8, train_steps=4, hidden=(16, 8))
    m.display_frequency = 0
    m.updater.momentum = 0.9
    w = Worker(); w.build(m); w.executor.setup()
    for step in range(4):
        w.executor.ticks += 1
        w.executor.train_one_batch(step)
    for p in w.train_net.params:
        p.proto.split_threshold = 50          # force multi-segment blobs
    path = checkpoint.save(str(tmp_path), w.train_net, w.pm, 4, m.name)
    assert checkpoint.latest(str(tmp_path)) == path
    w2 = Worker(seed=99); w2.build(m)
    assert not torch.equal(w2.train_net.params[0].data, w.train_net.params[0].data)
    assert checkpoint.load(path, w2.train_net, w2.pm) == 4
    for a, b in zip(w.train_net.params, w2.train_net.params):
        assert torch.equal(a.data, b.data) and torch.equal(a.history, b.history)
    ck = S.ParamCheckpoint(); ck.ParseFromString(open(path, "rb").read())
    assert ck.step == 4 and ck.param[0].name == "fc1/weight" and len(ck.param[0].blob) > 1
    device.reset_context()


def test_train_one_batch_async_matches_sync():
    """The pipelined result handle returns the same per-step metrics as reading the loss layer after each step."""
    from singa_b200 import device
    from singa_b200.models import configs
    from singa_b200.worker import Worker
    device.set_context("cpu", "torch")
    m = configs.mnist_mlp(batch=8, train_steps=5, hidden=(16,))
    m.display_frequency = 0
    ref = []
    w = Worker(seed=3); w.build(m); w.executor.setup()
    for step in range(5):
        w.executor.ticks += 1
        w.executor.train_one_batch(step)
        ref.append(w.train_net.losslayers[0].metric.reshape(-1).clone())
    w2 = Worker(seed=3); w2.build(m); w2.executor.setup()
    got, pending = [], None
    for step in range(5):
        w2.executor.ticks += 1
        h = w2.executor.train_one_batch_async(step)
        if pending is not None:
            got.append(pending.result().clone())
        pending = h
    got.append(pending.result().clone())
    assert pending.ready() and pending.step == 4
    for a, b in zip(ref, got):
        torch.testing.assert_close(a, b[:a.numel()])


def test_sharded_checkpoint_written_per_owner_and_resharded_on_load(tmp_path):
    """Every shard owner writes only the element ranges it owns (offset-tagged segments, no gather); loading reads all
    shard files, so a job with a different number of ranks can resume.  Ownership is emulated on one process."""
    from singa_b200 import checkpoint, device
    from singa_b200.models import configs
    from singa_b200.worker import Worker
    device.set_context("cpu", "torch")
    m = configs.mnist_mlp(batch=8, train_steps=3, hidden=(40, 24))
    m.display_frequency = 0
    m.updater.momentum = 0.9
    w = Worker(); w.build(m); w.executor.setup()
    for step in range(3):
        w.executor.ticks += 1
        w.executor.train_one_batch(step)
    for p in w.train_net.params:
        p.proto.split_threshold = 700
    total = w.pm.arena.total
    cut = [0, 3 * 1024, 3 * 1024 + 512 * 33, total]      # three owners, boundaries inside parameters

    class Owner:
        world = 3

        def __init__(self, r):
            self.rank = r

        def owned_ranges(self):
            return [(cut[self.rank], cut[self.rank + 1])]

    paths = []
    for r in range(3):
        w.pm.backend = Owner(r)
        paths.append(checkpoint.save(str(tmp_path), w.train_net, w.pm, 3, m.name))
    w.pm.backend = None
    assert [os.path.basename(p) for p in paths] == [f"step3.shard{r}of3.bin" for r in range(3)]
    assert checkpoint.latest(str(tmp_path)) == paths[0]
    sizes = [os.path.getsize(p) for p in paths]
    assert max(sizes) < 8 * total * 0.9                  # no file holds the whole model (value + history = 8 B/param)
    w2 = Worker(seed=7); w2.build(m)
    assert checkpoint.load(paths[1], w2.train_net, w2.pm) == 3          # any shard file names the checkpoint
    for a, b in zip(w.train_net.params, w2.train_net.params):
        assert torch.equal(a.data, b.data) and torch.equal(a.history, b.history)
    os.remove(paths[2])                                                  # an incomplete checkpoint is not "latest"
    assert checkpoint.latest(str(tmp_path)) is None
    with pytest.raises(FileNotFoundError):
        checkpoint.load(paths[0], w2.train_net, w2.pm)
    device.reset_context()


@pytest.mark.parametrize("case", [(3, 11, 4, 0, 35), (3, 7, 2, 3, 31), (1, 5, 4, 1, 33)])
def test_space_to_depth_convolution_identity(case):
    """The re-statement behind ops/conv.py::S2dConvPlan: a k x k stride-s convolution equals a ceil(k/s)^2 stride-1
    convolution of the space-to-depth image with the filter re-laid out through s2d_filter_index (CPU, fp64)."""
    import torch.nn.functional as F_
    from singa_b200.ops.conv import s2d_filter_index, space_to_depth_ref
    C, k, s, pad, H = case
    W, F, N = H - 2, 6, 2
    torch.manual_seed(k)
    x = torch.randn(N, C, H, W, dtype=torch.float64)
    w = torch.randn(F, C * k * k, dtype=torch.float64)
    ref = F_.conv2d(x, w.view(F, C, k, k), None, stride=s, padding=pad)
    ks = (k + s - 1) // s
    OH, OW = ref.shape[2], ref.shape[3]
    Hs, Ws = OH + ks - 1, OW + ks - 1
    Cs = (s * s * C + 3) // 4 * 4
    idx = s2d_filter_index(C, k, s, Cs).long()
    w_in = torch.where(idx >= 0, w[:, idx.clamp(min=0)], torch.zeros((), dtype=w.dtype))
    xs = space_to_depth_ref(x, s, pad, Hs, Ws, Cs)
    got = F_.conv2d(xs, w_in.view(F, Cs, ks, ks), None, stride=1, padding=0)
    torch.testing.assert_close(got, ref)
    # and back: the weight gradient of the inner convolution scatters onto the reference layout
    dy = torch.randn_like(ref)
    dw_ref = torch.nn.grad.conv2d_weight(x, (F, C, k, k), dy, stride=s, padding=pad).reshape(F, -1)
    dw_in = torch.nn.grad.conv2d_weight(xs, (F, Cs, ks, ks), dy, stride=1, padding=0).reshape(F, -1)
    dw = torch.zeros_like(dw_ref)
    live = idx >= 0
    dw[:, idx[live]] = dw_in[:, live]
    torch.testing.assert_close(dw, dw_ref)
