"""Every native memory-bound kernel against the plain-PyTorch fp32 restatement of the same op."""
import pytest
import torch

from singa_b200.ops import reference as R

pytestmark = pytest.mark.gpu

DEV = "cuda"


def K():
    from singa_b200.ops import kernels
    return kernels


def nhwc(t):
    return t.contiguous(memory_format=torch.channels_last)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("act", [1, 2])
@pytest.mark.parametrize("n", [1000 * 2500, 12345])
def test_act_fwd_bwd(dtype, act, n):
    x = torch.randn(n, device=DEV).to(dtype)
    y = torch.empty_like(x)
    K().act_fwd(x, y, act)
    ref = R.relu_fwd(x.float()) if act == 1 else R.stanh_fwd(x.float())
    tol = dict(rtol=1e-2, atol=1e-2) if dtype == torch.bfloat16 else dict(rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(y.float(), ref, **tol)
    dy = torch.randn(n, device=DEV).to(dtype)
    dx = torch.empty_like(x)
    K().act_bwd(dy, y, dx, act)
    refg = R.relu_bwd(dy.float(), y.float()) if act == 1 else R.stanh_bwd(dy.float(), y.float())
    torch.testing.assert_close(dx.float(), refg, **tol)


def test_dropout_matches_philox_reference():
    n, pkeep, seed = 100003, 0.6, 987654321
    x = torch.randn(n, device=DEV)
    y = torch.empty_like(x)
    K().dropout(x, y, pkeep, seed)
    mask = R.dropout_mask(seed, n, pkeep, DEV)
    torch.testing.assert_close(y, x * mask)
    assert abs(float((y != 0).float().mean()) - pkeep) < 0.01
    # device-side seed changes the mask
    sd = torch.tensor([5], dtype=torch.int64, device=DEV)
    y2 = torch.empty_like(x)
    K().dropout(x, y2, pkeep, seed, sd)
    assert not torch.equal(y, y2)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_colsum(dtype):
    rows, cols, ld = 1000, 2500, 2504
    buf = torch.randn(rows, ld, device=DEV).to(dtype)
    out = torch.empty(cols, device=DEV)
    K().colsum(buf, rows, cols, ld, out)
    torch.testing.assert_close(out, buf[:, :cols].float().sum(0), rtol=1e-3, atol=2e-2)
    K().colsum(buf, rows, cols, ld, out, accumulate=True)
    torch.testing.assert_close(out, 2 * buf[:, :cols].float().sum(0), rtol=1e-3, atol=4e-2)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("rows,cols,ld", [(5000, 96, 96), (777, 2000, 2000), (3000, 256, 264), (100, 8, 8),
                                          (40000, 384, 384), (1000, 2500, 2504), (513, 10, 16), (64, 50, 56)])
def test_colsum_vec8(rows, cols, ld, dtype):
    """Row strides of a 16-byte multiple take the 128-bit kernel (tall-skinny conv bias gradients, ragged widths)."""
    buf = torch.randn(rows, ld, device=DEV).to(dtype)
    out = torch.empty(cols, device=DEV)
    K().colsum(buf, rows, cols, ld, out)
    ref = buf[:, :cols].float().sum(0)
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=5e-2)
    K().colsum(buf, rows, cols, ld, out, accumulate=True)
    torch.testing.assert_close(out, 2 * ref, rtol=1e-3, atol=1e-1)


def test_cast_pack_and_copy2d_and_convert():
    F_, C, kk = 20, 3, 25
    w = torch.randn(F_, C * kk, device=DEV)
    dst = torch.full((F_, 80), 7.0, device=DEV, dtype=torch.bfloat16)
    K().cast_pack(w, C * kk, dst, 80, F_, C * kk, perm=(C, kk))
    ref = w.view(F_, C, kk).permute(0, 2, 1).reshape(F_, C * kk)
    torch.testing.assert_close(dst[:, :C * kk].float(), ref.to(torch.bfloat16).float())
    assert float(dst[:, C * kk:].abs().sum()) == 0.0
    src = torch.randn(33, 20, device=DEV).to(torch.bfloat16)
    d2 = torch.full((33, 24), 3.0, device=DEV, dtype=torch.bfloat16)
    K().copy2d_bf16(src, 20, d2, 24, 33, 20)
    assert torch.equal(d2[:, :20], src) and float(d2[:, 20:].abs().sum()) == 0.0
    a = torch.randn(5000, device=DEV)
    b = torch.empty(5000, device=DEV, dtype=torch.bfloat16)
    K().convert(a, b)
    assert torch.equal(b, a.to(torch.bfloat16))


def test_transpose_planes_roundtrip():
    n, c, h, w = 7, 48, 6, 5
    x = nhwc(torch.randn(n, c, h, w, device=DEV).to(torch.bfloat16))
    flat = torch.empty(n, c * h * w, device=DEV, dtype=torch.bfloat16)
    K().transpose_planes(x, flat, n, h * w, c)
    assert torch.equal(flat, x.reshape(n, -1))           # logical (c,h,w) flatten
    back = nhwc(torch.empty(n, c, h, w, device=DEV, dtype=torch.bfloat16))
    K().transpose_planes(flat, back, n, c, h * w)
    assert torch.equal(back, x)


def test_asum_and_fill_random():
    x = torch.randn(1 << 20, device=DEV)
    out = torch.zeros(1, device=DEV)
    K().asum(x, out)
    torch.testing.assert_close(out[0], x.abs().sum(), rtol=1e-4, atol=1e-2)
    u = torch.empty(1 << 20, device=DEV)
    K().fill_random(u, 0, -0.05, 0.05, 1.0, 42)
    assert float(u.min()) >= -0.05 and float(u.max()) <= 0.05 and abs(float(u.mean())) < 1e-3
    assert abs(float(u.std()) - 0.1 / 12 ** 0.5) < 1e-3
    g = torch.empty(1 << 20, device=DEV)
    K().fill_random(g, 1, 1.0, 2.0, 0.5, 43)
    assert abs(float(g.mean()) - 0.5) < 0.01 and abs(float(g.std()) - 1.0) < 0.01
    u2 = torch.empty(1 << 20, device=DEV)
    K().fill_random(u2, 0, -0.05, 0.05, 1.0, 42)
    assert torch.equal(u, u2)                               # reproducible


def test_parsers():
    n = 64
    px = torch.randint(0, 256, (n, 784), device=DEV, dtype=torch.uint8)
    out = torch.empty(n, 28, 28, device=DEV, dtype=torch.bfloat16)
    K().mnist_parse(px, out, 127.5, 1.0)
    torch.testing.assert_close(out.float().view(n, -1), px.float() / 127.5 - 1.0, rtol=1e-2, atol=1e-2)
    C, H, W, crop = 3, 32, 32, 24
    img = torch.randint(0, 256, (n, C, H, W), device=DEV, dtype=torch.uint8)
    oh = torch.randint(0, H - crop + 1, (n,), device=DEV, dtype=torch.int32)
    ow = torch.randint(0, W - crop + 1, (n,), device=DEV, dtype=torch.int32)
    mir = torch.randint(0, 2, (n,), device=DEV, dtype=torch.int32)
    dst = nhwc(torch.empty(n, C, crop, crop, device=DEV, dtype=torch.float32))
    K().rgb_parse(img, None, dst, n, C, H, W, crop, oh, ow, mir, 1 / 128)
    for i in (0, 5, 63):
        ref = img[i, :, oh[i]:oh[i] + crop, ow[i]:ow[i] + crop].float() / 128
        if mir[i]:
            ref = ref.flip(-1)
        torch.testing.assert_close(dst[i], ref)


@pytest.mark.parametrize("rows,dim,topk", [(1000, 10, 1), (256, 1000, 5), (33, 7, 2)])
def test_softmax_loss(rows, dim, topk):
    ld = (dim + 3) // 4 * 4
    logits = (torch.randn(rows, ld, device=DEV) * 3)[:, :dim]
    labels = torch.randint(0, dim, (rows,), device=DEV, dtype=torch.int32)
    ldg = (dim + 7) // 8 * 8
    grad = torch.zeros(rows, ldg, device=DEV, dtype=torch.bfloat16)[:, :dim]
    prob = torch.zeros(rows, dim, device=DEV)
    metric = torch.zeros(2, device=DEV)
    K().softmax_loss(logits, ld, labels, rows, dim, topk, 1.0, prob, grad, ldg, metric)
    m, g = R.softmax_loss(logits, labels, topk, 1.0)
    torch.testing.assert_close(metric, m, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(prob, torch.softmax(logits, 1), rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(grad.float(), g, rtol=2e-2, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("is_max", [True, False])
@pytest.mark.parametrize("shape,k,s,pad", [((4, 32, 32, 32), 3, 2, 0), ((2, 20, 24, 24), 2, 2, 0),
                                           ((3, 8, 9, 9), 4, 2, 0), ((2, 16, 13, 13), 3, 2, 1),
                                           ((2, 24, 11, 14), 3, 1, 1), ((2, 6, 10, 10), 3, 2, 0),
                                           ((5, 96, 55, 55), 3, 2, 0)])
def test_pool(shape, k, s, pad, is_max, dtype):
    """Row kernels (16-byte channel vectors; windows of 2 / 3, one to three windows per pixel) and the scalar
    fallbacks (k = 4, channel counts that are not a vector multiple) against the fp32 reference."""
    n, c, h, w = shape
    x32 = torch.randn(shape, device=DEV)
    if is_max:   # exercise the tie rule
        x32 = (x32 * 2).round() / 2
    x = nhwc(x32.to(dtype))
    oh, ow = R.pool_out_size(h, k, s, pad), R.pool_out_size(w, k, s, pad)
    y = nhwc(torch.full((n, c, oh, ow), float("nan"), device=DEV, dtype=dtype))
    K().pool_fwd(x, y, n, h, w, c, oh, ow, k, s, pad, is_max)
    yr = R.pool_fwd(x.float(), k, s, pad, is_max)
    tol = dict(rtol=1e-2, atol=1e-2) if dtype == torch.bfloat16 else dict(rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(y.float(), yr, **tol)
    dy = nhwc(torch.randn(n, c, oh, ow, device=DEV).to(dtype))
    dx = nhwc(torch.full(shape, float("nan"), device=DEV, dtype=dtype))
    K().pool_bwd(x, y, dy, dx, n, h, w, c, oh, ow, k, s, pad, is_max)
    dxr = R.pool_bwd(x.float(), y.float(), dy.float(), k, s, pad, is_max)
    tol = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dx.float(), dxr, **tol)


def test_pool_reference_matches_torch_when_exact():
    x = torch.randn(2, 4, 8, 8)
    torch.testing.assert_close(R.pool_fwd(x, 2, 2, 0, True), torch.nn.functional.max_pool2d(x, 2, 2))
    torch.testing.assert_close(R.pool_fwd(x, 2, 2, 0, False), torch.nn.functional.avg_pool2d(x, 2, 2))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("lsize,C,hw", [(5, 96, 13), (3, 32, 13), (5, 256, 7), (3, 6, 5), (5, 2400, 2)])
def test_lrn(lsize, C, hw, dtype):
    """Tiled kernels (whole pixels staged in shared memory) and the scalar fallbacks (C = 6: not a vector
    multiple, C = 2400: wider than a tile)."""
    n, h, w = 4, hw, hw
    alpha, beta, knorm = 1e-2, 0.75, 1.0
    x = nhwc((torch.randn(n, C, h, w, device=DEV) * 3).to(dtype))
    y = nhwc(torch.full_like(x, float("nan")))
    norm = nhwc(torch.full((n, C, h, w), float("nan"), device=DEV))
    K().lrn_fwd(x, y, norm, n * h * w, C, lsize, alpha, beta, knorm)
    yr, nr = R.lrn_fwd(x.float(), lsize, alpha, beta, knorm)
    torch.testing.assert_close(norm, nr, rtol=1e-4, atol=1e-5)
    tol = dict(rtol=1e-2, atol=1e-2) if dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(y.float(), yr, **tol)
    dy = nhwc(torch.randn(n, C, h, w, device=DEV).to(dtype))
    dx = nhwc(torch.full_like(x, float("nan")))
    K().lrn_bwd(x, norm, dy, dx, n * h * w, C, lsize, alpha, beta)
    dxr = R.lrn_bwd(x.float(), nr, dy.float(), lsize, alpha, beta)
    tol = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(dx.float(), dxr, **tol)


def test_lrn_reference_matches_autograd():
    x = torch.randn(2, 8, 3, 3, dtype=torch.float64, requires_grad=True)
    y, nm = R.lrn_fwd(x, 5, 1e-1, 0.75, 1.0)
    dy = torch.randn_like(y)
    y.backward(dy)
    torch.testing.assert_close(R.lrn_bwd(x.detach(), nm.detach(), dy, 5, 1e-1, 0.75), x.grad)


@pytest.mark.parametrize("N,C,H,W,k,s,pad", [(4, 16, 12, 12, 5, 1, 2), (2, 3, 32, 32, 5, 1, 2),
                                              (2, 8, 27, 27, 11, 4, 0), (3, 1, 28, 28, 5, 1, 0),
                                              (2, 3, 35, 35, 11, 4, 2)])
def test_im2col_col2im(N, C, H, W, k, s, pad):
    OH, OW = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
    x = nhwc(torch.randn(N, C, H, W, device=DEV).to(torch.bfloat16))
    K_ = k * k * C
    ld = (K_ + 7) // 8 * 8
    col = torch.zeros(N * OH * OW, ld, device=DEV, dtype=torch.bfloat16)
    K().im2col_nhwc(x, col, ld, N, H, W, C, OH, OW, k, s, pad)
    unf = torch.nn.functional.unfold(x.float(), k, padding=pad, stride=s)          # [N, C*k*k, L] (c,ky,kx)
    ref = unf.view(N, C, k * k, OH * OW).permute(0, 3, 2, 1).reshape(N * OH * OW, K_)   # (ky,kx,c)
    torch.testing.assert_close(col[:, :K_].float(), ref)
    dcol = torch.zeros(N * OH * OW, ld, device=DEV, dtype=torch.bfloat16)
    dcol[:, :K_] = torch.randn(N * OH * OW, K_, device=DEV).to(torch.bfloat16)
    dx = nhwc(torch.empty(N, C, H, W, device=DEV, dtype=torch.bfloat16))
    K().col2im_nhwc(dcol, ld, dx, N, H, W, C, OH, OW, k, s, pad)
    d = dcol[:, :K_].float().view(N, OH * OW, k * k, C).permute(0, 3, 2, 1).reshape(N, C * k * k, OH * OW)
    refdx = torch.nn.functional.fold(d, (H, W), k, padding=pad, stride=s)
    torch.testing.assert_close(dx.float(), refdx, rtol=2e-2, atol=3e-2)
    aux = nhwc(torch.randn(N, C, H, W, device=DEV).to(torch.bfloat16))
    K().col2im_nhwc(dcol, ld, dx, N, H, W, C, OH, OW, k, s, pad, 1, aux)
    torch.testing.assert_close(dx.float(), refdx * (aux.float() > 0), rtol=2e-2, atol=3e-2)


@pytest.mark.parametrize("kind", ["kSGD", "kNesterov", "kAdaGrad", "kRMSProp", "kAdaDelta"])
def test_update_arena(kind):
    from singa_b200.updater import _KIND_ID
    nchunks, n = 37, 37 * 1024
    w = torch.randn(n, device=DEV)
    g = torch.randn(n, device=DEV)
    h = torch.rand(n, device=DEV)
    u = torch.rand(n, device=DEV)
    w16 = torch.zeros(n, device=DEV, dtype=torch.bfloat16)
    chunk_param = torch.zeros(nchunks, dtype=torch.int32, device=DEV)
    chunk_param[10:30] = 1
    chunk_param[30:] = -1
    lr_mult = torch.tensor([1.0, 2.0], device=DEV)
    wd_mult = torch.tensor([1.0, 0.0], device=DEV)
    wr, hr, ur = w.clone(), h.clone(), u.clone()
    hp = dict(momentum=0.9, delta=1e-7, rho=0.9, grad_scale=0.5)
    K().update_arena(w, g, h, u, w16, chunk_param, lr_mult, wd_mult, 0, nchunks, kind=_KIND_ID[kind], lr=0.01,
                     weight_decay=0.001, **hp)
    for pid, (lo, hi) in enumerate(((0, 10 * 1024), (10 * 1024, 30 * 1024))):
        R.update_ref(kind, wr[lo:hi], g[lo:hi], hr[lo:hi], ur[lo:hi], lr=0.01 * float(lr_mult[pid]),
                     wd=0.001 * float(wd_mult[pid]), **hp)
    torch.testing.assert_close(w, wr, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(h, hr, rtol=1e-5, atol=1e-6)
    if kind == "kAdaDelta":
        torch.testing.assert_close(u, ur, rtol=1e-5, atol=1e-6)
    assert torch.equal(w16[:30 * 1024], w[:30 * 1024].to(torch.bfloat16))
    # device-resident LR overrides the by-value one
    lr_dev = torch.tensor([0.0], device=DEV)
    w2 = w.clone()
    K().update_arena(w2, g, h.clone(), u.clone(), None, chunk_param, lr_mult, wd_mult, 0, nchunks, kind=0, lr=1.0,
                     lr_dev=lr_dev)
    assert torch.equal(w2, w)


def test_easgd_and_randomsync():
    n = 100000
    w, s = torch.randn(n, device=DEV), torch.randn(n, device=DEV)
    wr, sr = w.clone(), s.clone()
    K().easgd_sync(w, s, 0.3)
    R.easgd_ref(wr, sr, 0.3)
    torch.testing.assert_close(w, wr)
    torch.testing.assert_close(s, sr)
    w, snap, s = torch.randn(n, device=DEV), torch.randn(n, device=DEV), torch.randn(n, device=DEV)
    wr, snr, sr = w.clone(), snap.clone(), s.clone()
    nsel = torch.zeros(1, dtype=torch.int64, device=DEV)
    K().randomsync(w, snap, s, 0.25, 777, None, nsel)
    cnt = R.randomsync_ref(wr, snr, sr, 0.25, 777)
    assert int(nsel.item()) == cnt and abs(cnt / n - 0.25) < 0.01
    torch.testing.assert_close(w, wr)
    torch.testing.assert_close(snap, snr)
    torch.testing.assert_close(s, sr)


def test_sharded_exchange_and_group_average():
    """One launch over all server shards (pointer table) equals the per-shard launches; the neighbour-group averaging
    moves a shard to the mean of the group copies.  'Peers' are local buffers here (one GPU)."""
    n, nsh = 8 * 1024, 4
    per = n // nsh
    w = torch.randn(n, device=DEV)
    servers = [torch.randn(n, device=DEV) for _ in range(nsh)]       # full-size arenas; shard k uses [k*per, (k+1)*per)
    wr, sr = w.clone(), [s.clone() for s in servers]
    tbl = K().shard_table([s.data_ptr() for s in servers], per)
    K().easgd_sync_sharded(w, tbl, 0.3)
    for k in range(nsh):
        R.easgd_ref(wr[k * per:(k + 1) * per], sr[k][k * per:(k + 1) * per], 0.3)
    torch.testing.assert_close(w, wr)
    for k in range(nsh):
        torch.testing.assert_close(servers[k], sr[k])
    # RandomSync: same subset as the single-shard kernel with the same (device-mixed) seed
    w, snap = torch.randn(n, device=DEV), torch.randn(n, device=DEV)
    servers = [torch.randn(n, device=DEV) for _ in range(nsh)]
    w1, snap1, s1 = w.clone(), snap.clone(), [s.clone() for s in servers]
    step = torch.tensor([41], dtype=torch.int64, device=DEV)
    nsel = torch.zeros(1, dtype=torch.int64, device=DEV)
    tbl = K().shard_table([s.data_ptr() for s in servers], per)
    K().randomsync_sharded(w, snap, tbl, 0.25, 777, step, None, nsel)
    mixed = 777 ^ ((42 * 0x9E3779B97F4A7C15) & (2**64 - 1))
    joined = torch.cat([s1[k][k * per:(k + 1) * per] for k in range(nsh)])
    cnt = R.randomsync_ref(w1, snap1, joined, 0.25, mixed)
    assert int(nsel.item()) == cnt
    torch.testing.assert_close(w, w1)
    torch.testing.assert_close(snap, snap1)
    torch.testing.assert_close(torch.cat([servers[k][k * per:(k + 1) * per] for k in range(nsh)]), joined)
    # neighbour-group averaging of one shard
    mine = torch.randn(n, device=DEV)
    others = [torch.randn(n, device=DEV) for _ in range(3)]
    ref = mine.clone()
    lo, cnt = 2 * per, per
    ref[lo:lo + cnt] = (mine[lo:lo + cnt] + sum(o[lo:lo + cnt] for o in others)) / 4
    K().group_average(mine, K().shard_table([o.data_ptr() for o in others], per), lo, cnt)
    torch.testing.assert_close(mine, ref, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_mnist_distort_kernel_matches_reference(dtype):
    n, S_, out = 16, 28, 24
    src = torch.randint(0, 256, (n, S_, S_), device=DEV, dtype=torch.uint8)
    gen = torch.Generator(device=DEV)
    gen.manual_seed(5)
    labels = torch.randint(0, 10, (n,), device=DEV, dtype=torch.int32)
    A, disp = R.mnist_distortion_params(n, S_, out, 15.0, 20.0, 7, 2.5, 4.0, labels, gen, torch.device(DEV))
    dst = torch.empty(n, out, out, device=DEV, dtype=dtype)
    K().mnist_distort(src, dst, n, S_, out, A, disp, 127.5, 1.0)
    ref = R.mnist_distort(src, A, disp, out, 127.5, 1.0)
    tol = dict(rtol=1e-2, atol=1e-2) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(dst.float(), ref, **tol)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,Kd,N", [(1000, 500, 10), (37, 1024, 16), (64, 13, 1)])
def test_ip_small_kernels(M, Kd, N, dtype):
    """Classifier-sized inner product (<= 16 outputs) on CUDA cores: forward with bias + activation, data gradient
    with a fused activation gradient, weight + bias gradient (plain and accumulating) against fp32 torch."""
    from singa_b200.ops import gemm as G
    torch.manual_seed(M + N)
    q = 8 if dtype == torch.bfloat16 else 4
    ldn = (N + q - 1) // q * q
    x = (torch.randn(M, Kd, device=DEV) * 0.5).to(dtype)
    w = torch.zeros(Kd, ldn, device=DEV, dtype=dtype)
    w[:, :N] = (torch.randn(Kd, N, device=DEV) * 0.1).to(dtype)
    b = torch.randn(N, device=DEV)
    y = torch.full((M, ldn), float("nan"), device=DEV)                       # fp32 output (feeds the loss)
    K().ip_small_fwd(x, w[:, :N], y[:, :N], M, Kd, N, bias=b, act=G.ACT_STANH)
    ref = 1.7159047 * torch.tanh(0.66666667 * (x.float() @ w[:, :N].float() + b))
    torch.testing.assert_close(y[:, :N], ref, rtol=2e-3, atol=2e-3)
    dy = (torch.randn(M, ldn, device=DEV) * 0.5).to(dtype)
    aux = (torch.randn(M, Kd, device=DEV)).to(dtype)
    dx = torch.full((M, Kd), float("nan"), device=DEV, dtype=dtype)
    K().ip_small_dgrad(dy[:, :N], w[:, :N], dx, M, Kd, N, grad_act=G.ACT_RELU, aux=aux)
    dref = (dy[:, :N].float() @ w[:, :N].float().t()) * (aux.float() > 0)
    tol = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(dx.float(), dref, **tol)
    K().ip_small_dgrad(dy[:, :N], w[:, :N], dx, M, Kd, N)
    torch.testing.assert_close(dx.float(), dy[:, :N].float() @ w[:, :N].float().t(), **tol)
    dw = torch.full((Kd, ldn), float("nan"), device=DEV)
    db = torch.full((N,), float("nan"), device=DEV)
    ws = K().ip_small_wgrad_workspace(M, Kd, DEV)
    K().ip_small_wgrad(x, dy[:, :N], dw[:, :N], M, Kd, N, ws, db=db)
    wref = x.float().t() @ dy[:, :N].float()
    bref = dy[:, :N].float().sum(0)
    torch.testing.assert_close(dw[:, :N], wref, rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(db, bref, rtol=1e-4, atol=1e-3)
    first = dw.clone()
    K().ip_small_wgrad(x, dy[:, :N], dw[:, :N], M, Kd, N, ws, db=db, accumulate=True)
    torch.testing.assert_close(dw[:, :N], 2 * wref, rtol=1e-4, atol=2e-3)
    torch.testing.assert_close(db, 2 * bref, rtol=1e-4, atol=2e-3)
    again = torch.full_like(dw, float("nan"))
    K().ip_small_wgrad(x, dy[:, :N], again[:, :N], M, Kd, N, ws)       # same bits on every run (slice-ordered sums)
    assert torch.equal(again[:, :N], first[:, :N])
