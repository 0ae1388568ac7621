"""Plain-PyTorch fp32 restatement of every layer / updater / init rule of the reference.

This module is (a) the CPU execution path (BASELINE config 1, CI without a GPU), (b) the oracle that the
native kernels are tested against, and (c) on a GPU, the cuBLAS/cuDNN "library baseline" backend.
All functions take/return logical NCHW tensors and never touch the native library.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

STANH_A, STANH_B = 1.7159047, 0.66666667
FLT_MIN = 1.17549435e-38

# Library-baseline precision knob (bench.py --impl torch --dtype bf16): run the matmul / convolution sites under
# torch.autocast(bf16) — cuBLAS/cuDNN bf16 tensor-core kernels with fp32 accumulation, fp32 everywhere else — the
# library counterpart of the native bf16 path.  Off (fp32, optionally TF32 via torch.backends) by default.
AUTOCAST_BF16 = False


def _f(t):
    return t.float() if t.dtype == torch.bfloat16 else t


def _mm_ctx(t):
    return torch.autocast(device_type=t.device.type, dtype=torch.bfloat16, enabled=AUTOCAST_BF16 and t.is_cuda)


# ------------------------------------------------------------------------------------------- activations
def relu_fwd(x):
    return torch.clamp_min(x, 0)


def relu_bwd(dy, y):
    return dy * (y > 0).to(dy.dtype)


def stanh_fwd(x):
    """1.7159047 * tanh(0.66666667 x)  (include/mshadow/cxxnet_op.h:77-81)."""
    return STANH_A * torch.tanh(STANH_B * x)


def stanh_bwd(dy, y):
    """dy * (A*B - B/A * y^2)  (cxxnet_op.h:83-87)."""
    return dy * (STANH_A * STANH_B - (STANH_B / STANH_A) * y * y)


# ------------------------------------------------------------------------------------------- philox
def philox_uniform(seed: int, n: int, device="cpu") -> torch.Tensor:
    """n uniforms in [0,1): element e uses counter e//4, lane e%4 — bit-identical to csrc/philox.cuh."""
    n4 = (n + 3) // 4
    idx = torch.arange(n4, dtype=torch.int64, device=device)
    M32 = 0xFFFFFFFF
    c = [idx & M32, (idx >> 32) & M32, torch.zeros_like(idx), torch.zeros_like(idx)]
    k0, k1 = seed & M32, (seed >> 32) & M32
    M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85

    def mulhilo(a: int, b: torch.Tensor):
        # 32x32 -> 64 without overflowing int64: split b into 16-bit halves
        b_lo, b_hi = b & 0xFFFF, b >> 16
        p_lo = a * b_lo                      # < 2^48
        p_hi = a * b_hi                      # < 2^48
        lo = (p_lo + ((p_hi & 0xFFFF) << 16))
        carry = lo >> 32
        hi = (p_hi >> 16) + carry
        return hi & M32, lo & M32

    for _ in range(10):
        hi0, lo0 = mulhilo(M0, c[0])
        hi1, lo1 = mulhilo(M1, c[2])
        c = [(hi1 ^ c[1] ^ k0) & M32, lo1, (hi0 ^ c[3] ^ k1) & M32, lo0]
        k0 = (k0 + W0) & M32
        k1 = (k1 + W1) & M32
    u = torch.stack([(ci >> 8).to(torch.float32) * (1.0 / 16777216.0) for ci in c], dim=1).reshape(-1)
    return u[:n]


def dropout_mask(seed: int, n: int, pkeep: float, device="cpu") -> torch.Tensor:
    return (philox_uniform(seed, n, device) < pkeep).to(torch.float32) / pkeep


# ------------------------------------------------------------------------------------------- pooling
def pool_out_size(h: int, k: int, s: int, pad: int = 0) -> int:
    """ceil((h + 2p - k)/s) + 1 with the last window required to start inside the padded image
    (the reference's ceil form, src/worker/layer.cc:498-501, made consistent with the window clipping of
    include/mshadow/tensor_expr_ext.h:790-811 — the ceil/floor mismatch of the survey's §2.8 is not kept)."""
    o = int(math.ceil((h + 2 * pad - k) / s)) + 1
    if pad > 0 and (o - 1) * s >= h + pad:
        o -= 1
    return max(o, 1)


def _windows(x, k, s, pad, oh, ow, fill):
    """[N,C,oh,ow,k*k] view of clipped windows (out-of-image taps = fill)."""
    n, c, h, w = x.shape
    need_h = (oh - 1) * s + k
    need_w = (ow - 1) * s + k
    xp = F.pad(x, (pad, max(0, need_w - w - pad), pad, max(0, need_h - h - pad)), value=fill)
    cols = F.unfold(xp, kernel_size=k, stride=s)          # [N, C*k*k, L]
    L = cols.shape[-1]
    full_ow = (xp.shape[-1] - k) // s + 1
    cols = cols.view(n, c, k * k, L // full_ow, full_ow)[:, :, :, :oh, :ow]
    return cols.permute(0, 1, 3, 4, 2)


def pool_fwd(x, k, s, pad, is_max):
    n, c, h, w = x.shape
    oh, ow = pool_out_size(h, k, s, pad), pool_out_size(w, k, s, pad)
    if is_max:
        return _windows(x, k, s, pad, oh, ow, float("-inf")).amax(dim=-1)
    return _windows(x, k, s, pad, oh, ow, 0.0).sum(dim=-1) / (k * k)


def pool_bwd(x, y, dy, k, s, pad, is_max):
    """MAX: gradient goes to EVERY input equal to the window max (mshadow unpool, tensor_expr_ext.h:819-848);
    AVE: dy / k^2 to every tap."""
    n, c, h, w = x.shape
    oh, ow = y.shape[2], y.shape[3]
    dx = torch.zeros_like(x)
    for ky in range(k):
        for kx in range(k):
            # taps: input (oy*s - pad + ky, ox*s - pad + kx)
            iy = torch.arange(oh, device=x.device) * s - pad + ky
            ix = torch.arange(ow, device=x.device) * s - pad + kx
            vy = (iy >= 0) & (iy < h)
            vx = (ix >= 0) & (ix < w)
            if not vy.any() or not vx.any():
                continue
            iyv, ixv = iy[vy], ix[vx]
            g = dy[:, :, vy][:, :, :, vx]
            if is_max:
                xs = x[:, :, iyv][:, :, :, ixv]
                ys = y[:, :, vy][:, :, :, vx]
                g = g * (xs == ys).to(g.dtype)
            else:
                g = g / (k * k)
            dx[:, :, iyv[:, None], ixv[None, :]] += g
    return dx


# ------------------------------------------------------------------------------------------- LRN
def _chwindow_sum(t, lsize):
    half = lsize // 2
    c = t.shape[1]
    tp = F.pad(t, (0, 0, 0, 0, half, half))
    out = torch.zeros_like(t)
    for d in range(lsize):
        out = out + tp[:, d:d + c]
    return out


def lrn_fwd(x, lsize, alpha, beta, knorm):
    norm = knorm + (alpha / lsize) * _chwindow_sum(x * x, lsize)
    return x * norm.pow(-beta), norm


def lrn_bwd(x, norm, dy, lsize, alpha, beta):
    t = dy * x * norm.pow(-beta - 1.0)
    return dy * norm.pow(-beta) - (2.0 * beta * alpha / lsize) * x * _chwindow_sum(t, lsize)


# ------------------------------------------------------------------------------------------- MNIST distortion
def mnist_distortion_params(n, in_size, out_size, gamma, beta, kernel, sigma, alpha, labels, gen, device):
    """Per-image inverse affine maps [n, 6] (output pixel -> source pixel) and the elastic displacement field
    [n, 2, out, out] (or None), following the reference's disabled block (src/worker/layer.cc:406-438): resize by
    (1 + r * gamma / 100) per axis, then — with equal probability — a rotation by r * beta degrees about the centre of
    the resized image or a horizontal shear of r * beta / 90 (halved for the digits 1 and 7); the elastic field is
    alpha x a Gaussian blur (kernel, sigma) of uniform noise in [-1, 1]."""
    def rnd(*shape):
        return torch.rand(*shape, generator=gen, device=device) * 2 - 1
    sx = 1 + (rnd(n) * gamma / 100.0 if gamma else torch.zeros(n, device=device))
    sy = 1 + (rnd(n) * gamma / 100.0 if gamma else torch.zeros(n, device=device))
    M = torch.eye(3, device=device).repeat(n, 1, 1)
    if beta:
        r = rnd(n)
        rot = torch.rand(n, generator=gen, device=device) < 0.5
        th = torch.deg2rad(r * beta)
        c, s_ = torch.cos(th), torch.sin(th)
        cx, cy = in_size * sx / 2, in_size * sy / 2
        R = torch.eye(3, device=device).repeat(n, 1, 1)
        R[:, 0, 0], R[:, 0, 1], R[:, 0, 2] = c, s_, (1 - c) * cx - s_ * cy
        R[:, 1, 0], R[:, 1, 1], R[:, 1, 2] = -s_, c, s_ * cx + (1 - c) * cy
        sh = r * beta / 90.0
        if labels is not None:
            lab = labels.to(device)
            sh = torch.where((lab == 1) | (lab == 7), sh / 2, sh)
        Sh = torch.eye(3, device=device).repeat(n, 1, 1)
        Sh[:, 0, 1] = sh
        M = torch.where(rot[:, None, None], R, Sh)
    S_ = torch.eye(3, device=device).repeat(n, 1, 1)
    S_[:, 0, 0], S_[:, 1, 1] = sx, sy
    A = torch.linalg.inv(M @ S_)[:, :2, :].reshape(n, 6).contiguous().float()
    disp = None
    if kernel and sigma and alpha:
        k = int(kernel) | 1
        ax = torch.arange(k, device=device, dtype=torch.float32) - k // 2
        g1 = torch.exp(-ax * ax / (2 * sigma * sigma))
        g1 = g1 / g1.sum()
        noise = rnd(n * 2, 1, out_size, out_size)
        blur = F.conv2d(F.conv2d(noise, g1.view(1, 1, 1, k), padding=(0, k // 2)), g1.view(1, 1, k, 1),
                        padding=(k // 2, 0))
        disp = (blur * alpha).view(n, 2, out_size, out_size).contiguous()
    return A, disp


def mnist_distort(src, A, disp, out_size, norm_a, norm_b):
    """Bilinear sampling of src [n, S, S] at A (x, y, 1) + disp, zero outside, then x / norm_a - norm_b."""
    n, S_, _ = src.shape
    dev = src.device
    ys, xs = torch.meshgrid(torch.arange(out_size, device=dev, dtype=torch.float32),
                            torch.arange(out_size, device=dev, dtype=torch.float32), indexing="ij")
    a = A.view(n, 6, 1, 1)
    sx = a[:, 0] * xs + a[:, 1] * ys + a[:, 2]
    sy = a[:, 3] * xs + a[:, 4] * ys + a[:, 5]
    if disp is not None:
        sx, sy = sx + disp[:, 0], sy + disp[:, 1]
    x0, y0 = torch.floor(sx), torch.floor(sy)
    wx, wy = sx - x0, sy - y0
    img = src.float()

    def px(yy, xx):
        ok = (yy >= 0) & (yy < S_) & (xx >= 0) & (xx < S_)
        idx = (yy.clamp(0, S_ - 1) * S_ + xx.clamp(0, S_ - 1)).long().view(n, -1)
        return (img.view(n, -1).gather(1, idx).view(n, out_size, out_size)) * ok

    v = (1 - wy) * ((1 - wx) * px(y0, x0) + wx * px(y0, x0 + 1)) + wy * ((1 - wx) * px(y0 + 1, x0) + wx * px(y0 + 1, x0 + 1))
    return v / norm_a - norm_b


# ------------------------------------------------------------------------------------------- conv / inner product
def conv_fwd(x, weight, bias, k, s, pad):
    """weight {F, C*k*k} with column (c*k + ky)*k + kx  (SURVEY A.1)."""
    f = weight.shape[0]
    c = x.shape[1]
    with _mm_ctx(x):
        y = F.conv2d(x, weight.view(f, c, k, k), bias, stride=s, padding=pad)
    return _f(y)


def conv_bwd(x, weight, dy, k, s, pad, need_dx=True):
    f = weight.shape[0]
    c = x.shape[1]
    w4 = weight.view(f, c, k, k)
    db = dy.sum(dim=(0, 2, 3))
    if AUTOCAST_BF16 and x.is_cuda:   # conv2d_weight / conv2d_input are not autocast-aware: cast explicitly
        xb, wb, dyb = x.bfloat16(), w4.bfloat16(), dy.bfloat16()
        dw = torch.nn.grad.conv2d_weight(xb, w4.shape, dyb, stride=s, padding=pad).float().reshape(f, -1)
        dx = torch.nn.grad.conv2d_input(x.shape, wb, dyb, stride=s, padding=pad).float() if need_dx else None
        return dx, dw, db
    dw = torch.nn.grad.conv2d_weight(x, w4.shape, dy, stride=s, padding=pad).reshape(f, -1)
    dx = torch.nn.grad.conv2d_input(x.shape, w4, dy, stride=s, padding=pad) if need_dx else None
    return dx, dw, db


def ip_fwd(x2d, weight, bias):
    """weight {vdim, hdim}: y = x W + b  (src/worker/layer.cc:186-196)."""
    with _mm_ctx(x2d):
        y = _f(x2d @ weight)
    return y + bias if bias is not None else y


def ip_bwd(x2d, weight, dy, need_dx=True):
    db = dy.sum(dim=0)
    with _mm_ctx(x2d):
        dw = _f(x2d.t() @ dy)
        dx = _f(dy @ weight.t()) if need_dx else None
    return dx, dw, db


# ------------------------------------------------------------------------------------------- softmax loss
def softmax_loss(logits, labels, topk=1, scale=1.0):
    """returns (metric[2], grad): metric = scale*[mean NLL, top-k accuracy], grad=(p-onehot)*scale/N."""
    n, dim = logits.shape
    p = torch.softmax(logits.float(), dim=1)
    lab = labels.long().clamp(0, dim - 1)
    pl = p.gather(1, lab[:, None]).squeeze(1)
    loss = -torch.log(torch.clamp_min(pl, FLT_MIN)).mean()
    xl = logits.float().gather(1, lab[:, None])
    rank = (logits.float() > xl).sum(dim=1)
    acc = (rank < topk).float().mean()
    grad = p.clone()
    grad[torch.arange(n, device=logits.device), lab] -= 1.0
    grad = grad * (scale / n)
    return torch.stack([loss * scale, acc * scale]), grad


# ------------------------------------------------------------------------------------------- LR schedules / updaters
def learning_rate(up, step: int) -> float:
    """UpdaterProto -> LR(step)  (src/utils/updater.cc:11-51)."""
    from ..proto import schema as S
    base = up.base_learning_rate
    m = up.learning_rate_change_method
    name = S.enum_name("UpdaterProto.ChangeProto", m)
    freq = up.learning_rate_change_frequency
    if name == "kFixed":
        return base
    if name == "kLinear":
        r = step * 1.0 / freq
        return (1.0 - r) * base + r * up.final_learning_rate
    if name == "kExponential":
        return base / math.pow(2, step * 1.0 / freq)
    if name == "kInverse_t":
        return base / (1.0 + step * 1.0 / up.final_learning_rate)
    if name == "kInverse":
        return base * math.pow(1.0 + up.gamma * step, -up.pow)
    if name == "kStep":
        return base * math.pow(up.gamma, step // freq)
    raise ValueError(f"unknown learning-rate change method {m}")


def update_ref(kind: str, w, g, h, u, *, lr, momentum=0.0, wd=0.0, delta=1e-7, rho=0.9, grad_scale=1.0):
    """In-place update of fp32 tensors; same math as csrc/updaters.cu::update_one."""
    sg = g * grad_scale
    if kind == "kSGD":
        gg = sg + wd * w if wd > 0 else sg
        if momentum > 0:
            h.mul_(momentum).add_(gg, alpha=lr)
            w.sub_(h)
        else:
            w.sub_(gg, alpha=lr)
    elif kind == "kNesterov":
        gg = sg + wd * w if wd > 0 else sg
        hold = h.clone()
        h.mul_(momentum).add_(gg, alpha=lr)
        w.sub_(h * (1 + momentum) - hold * momentum)
    elif kind == "kAdaGrad":
        h.add_(sg * sg)
        gg = sg + wd * w if wd > 0 else sg
        w.sub_(lr * gg / torch.sqrt(h + delta))
    elif kind == "kRMSProp":
        h.mul_(rho).add_(sg * sg, alpha=1 - rho)
        gg = sg + wd * w if wd > 0 else sg
        w.sub_(lr * gg / torch.sqrt(h + delta))
    elif kind == "kAdaDelta":
        gg = sg + wd * w if wd > 0 else sg
        h.mul_(rho).add_(gg * gg, alpha=1 - rho)
        t = gg * torch.sqrt(u + delta) / torch.sqrt(h + delta)
        u.mul_(rho).add_(t * t, alpha=1 - rho)
        w.sub_(t)
    else:
        raise ValueError(kind)


# ------------------------------------------------------------------------------------------- sync rules
def easgd_ref(w, s, alpha):
    """d = (w - s) alpha; s += d; w -= d  (src/utils/param.cc:244-284)."""
    d = (w - s) * alpha
    s.add_(d)
    w.sub_(d)


def randomsync_ref(w, snap, s, ratio, seed):
    """Shared-seed subset exchange (src/utils/param.cc:130-228) on {i: philox(seed,i) < ratio}."""
    n = w.numel()
    sel = torch.ones(n, dtype=torch.bool, device=w.device) if ratio >= 1 else \
        (philox_uniform(seed, n, w.device) < ratio)
    wf, sf, snf = w.view(-1), s.view(-1), snap.view(-1)
    delta = wf[sel] - snf[sel]
    old = sf[sel].clone()
    sf[sel] += delta
    wf[sel] += old - snf[sel]
    snf[sel] = wf[sel]
    return int(sel.sum())
