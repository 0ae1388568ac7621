"""Implicit-GEMM convolution kernels (TMA im2col tiles) against torch's fp32 convolution of the same inputs:
forward with bias + ReLU, data gradient (with a fused activation gradient), weight gradient; bf16 and tf32 operands;
grey / RGB inputs (padded pixels), 11x11 stride 4 (AlexNet conv1), 5x5 pad 2, channel counts that need zero-filled
chunks."""
import pytest
import torch
import torch.nn.functional as F_

from singa_b200.ops import gemm as G
from singa_b200.ops.conv import ConvPlan

pytestmark = pytest.mark.gpu
DEV = "cuda"

CASES = [
    # N, C, H, W, F, k, stride, pad
    (2, 3, 35, 35, 24, 11, 4, 0),      # AlexNet conv1 shape family: RGB, 11x11, stride 4
    (3, 1, 28, 28, 20, 5, 1, 0),       # LeNet conv1: grey input
    (2, 20, 12, 12, 50, 5, 1, 0),      # LeNet conv2: 20 channels (not a multiple of 8)
    (2, 3, 32, 32, 32, 5, 1, 2),       # CIFAR conv1: pad 2
    (2, 32, 16, 16, 32, 5, 1, 2),      # CIFAR conv2
    (2, 96, 13, 13, 64, 5, 1, 2),      # AlexNet conv2 family: 96 channels (1.5 chunks of 64 bf16)
    (1, 64, 9, 11, 136, 3, 1, 1),      # 3x3, ragged width, F > 128
    (2, 16, 15, 15, 16, 3, 2, 1),      # strided with a data gradient (explicit col2im path)
]


def _nhwc(t):
    return t.contiguous(memory_format=torch.channels_last)


def _tols(dtype, K):
    if dtype == torch.bfloat16:
        return dict(rtol=3e-2, atol=3e-2 * (K ** 0.5) * 0.1 + 2e-2)
    return dict(rtol=2e-3, atol=2.0 ** -10 * (K ** 0.5) * 6.0 * 0.25 + 1e-4)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", CASES)
def test_conv_forward_dgrad_wgrad(case, dtype):
    N, C, H, W, F, k, s, p = case
    torch.manual_seed(sum(case))
    x32 = torch.randn(N, C, H, W, device=DEV) * 0.5
    w32 = torch.randn(F, C * k * k, device=DEV) * 0.2
    b32 = torch.randn(F, device=DEV)
    x = _nhwc(x32.to(dtype))
    plan = ConvPlan(N, C, H, W, F, k, s, p, dtype, DEV, need_dgrad=True)
    plan.pack(w32)
    OH, OW = plan.OH, plan.OW
    xr, wr = x.float(), w32.to(dtype).float().view(F, C, k, k)
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        # ---- forward, bias + ReLU in the epilogue
        y = _nhwc(torch.full((N, F, OH, OW), float("nan"), device=DEV, dtype=dtype))
        y2d = y.permute(0, 2, 3, 1).reshape(plan.P, F)
        assert y2d.data_ptr() == y.data_ptr()
        plan.forward(x if C > 1 else x[:, 0], y2d, bias=b32, act=G.ACT_RELU)
        torch.cuda.synchronize()
        ref = torch.relu(F_.conv2d(xr, wr, b32, stride=s, padding=p))
        torch.testing.assert_close(y.float(), ref, **_tols(dtype, C * k * k))
        # ---- weight gradient
        dy32 = torch.randn(N, F, OH, OW, device=DEV) * 0.5
        dy = _nhwc(dy32.to(dtype))
        dw = torch.full((F, C * k * k), float("nan"), device=DEV)
        plan.wgrad(dy, dw)
        torch.cuda.synchronize()
        dw_ref = torch.nn.grad.conv2d_weight(xr, wr.shape, dy.float(), stride=s, padding=p).reshape(F, -1)
        tol = _tols(dtype, plan.P)
        torch.testing.assert_close(dw, dw_ref, **tol)
        dw2 = torch.ones_like(dw)
        plan.wgrad(dy, dw2, accumulate=True)
        torch.cuda.synchronize()
        torch.testing.assert_close(dw2, dw_ref + 1.0, **tol)
        # ---- data gradient with the activation gradient of the layer below fused in
        aux32 = torch.tanh(torch.randn(N, C, H, W, device=DEV)) * 1.7
        aux = _nhwc(aux32.to(dtype))
        dx = _nhwc(torch.full((N, C, H, W), float("nan"), device=DEV, dtype=dtype))
        dx2d = dx.permute(0, 2, 3, 1).reshape(N * H * W, C)
        a2d = aux.permute(0, 2, 3, 1).reshape(N * H * W, C)
        plan.dgrad(dy, dx2d, grad_act=G.ACT_STANH, aux2d=a2d)
        torch.cuda.synchronize()
        dx_ref = torch.nn.grad.conv2d_input(xr.shape, wr, dy.float(), stride=s, padding=p)
        dx_ref = dx_ref * G.act_grad_from_output_ref(aux.float(), G.ACT_STANH)
        torch.testing.assert_close(dx.float(), dx_ref, **_tols(dtype, F * k * k))
    finally:
        torch.backends.cudnn.allow_tf32 = prev


def test_conv_larger_batch_many_tiles():
    """More pixels than one wave of tiles, tile boundaries inside images, tail tile beyond the last image."""
    N, C, H, W, F, k, s, p = 37, 32, 17, 19, 96, 3, 1, 1
    dtype = torch.float32
    x = _nhwc(torch.randn(N, C, H, W, device=DEV))
    w32 = torch.randn(F, C * k * k, device=DEV) * 0.1
    plan = ConvPlan(N, C, H, W, F, k, s, p, dtype, DEV, need_dgrad=False)
    plan.pack(w32)
    y = _nhwc(torch.empty(N, F, plan.OH, plan.OW, device=DEV))
    plan.forward(x, y.permute(0, 2, 3, 1).reshape(plan.P, F))
    torch.cuda.synchronize()
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        ref = F_.conv2d(x, w32.view(F, C, k, k), None, stride=s, padding=p)
    finally:
        torch.backends.cudnn.allow_tf32 = prev
    torch.testing.assert_close(y, ref, **_tols(dtype, C * k * k))


S2D_CASES = [
    # N, C, H, W, F, k, stride, pad
    (2, 3, 35, 35, 24, 11, 4, 0),      # AlexNet conv1 family: 3x3 over 48 channels after space-to-depth
    (3, 3, 31, 29, 40, 7, 2, 3),       # padded, ragged, stride 2 (4x4 over 12 channels, padded to 16-byte pixels)
    (2, 1, 33, 33, 16, 5, 4, 1),       # grey input, k not a multiple of the stride
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", S2D_CASES)
def test_conv_space_to_depth_plan(case, dtype):
    """Strided convolution of a thin input layer as a stride-1 convolution over the space-to-depth image."""
    from singa_b200.ops.conv import S2dConvPlan, make_plan
    N, C, H, W, F, k, s, p = case
    torch.manual_seed(sum(case))
    x32 = torch.randn(N, C, H, W, device=DEV) * 0.5
    w32 = torch.randn(F, C * k * k, device=DEV) * 0.2
    b32 = torch.randn(F, device=DEV)
    x = _nhwc(x32.to(dtype))
    plan = S2dConvPlan(N, C, H, W, F, k, s, p, dtype, DEV)
    if C * (2 if dtype == torch.bfloat16 else 4) * s * s >= 64:
        assert isinstance(make_plan(N, C, H, W, F, k, s, p, dtype, DEV, need_dgrad=False), S2dConvPlan)
    plan.pack(w32)
    OH, OW = plan.OH, plan.OW
    xr, wr = x.float(), w32.to(dtype).float().view(F, C, k, k)
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        y = _nhwc(torch.full((N, F, OH, OW), float("nan"), device=DEV, dtype=dtype))
        y2d = y.permute(0, 2, 3, 1).reshape(plan.P, F)
        plan.forward(x if C > 1 else x[:, 0], y2d, bias=b32, act=G.ACT_RELU)
        torch.cuda.synchronize()
        ref = torch.relu(F_.conv2d(xr, wr, b32, stride=s, padding=p))
        assert ref.shape == y.shape
        torch.testing.assert_close(y.float(), ref, **_tols(dtype, C * k * k))
        dy32 = torch.randn(N, F, OH, OW, device=DEV) * 0.5
        dy = _nhwc(dy32.to(dtype))
        dw = torch.full((F, C * k * k), float("nan"), device=DEV)
        plan.wgrad(dy, dw)
        torch.cuda.synchronize()
        dw_ref = torch.nn.grad.conv2d_weight(xr, wr.shape, dy.float(), stride=s, padding=p).reshape(F, -1)
        tol = _tols(dtype, plan.P)
        torch.testing.assert_close(dw, dw_ref, **tol)
        dw2 = torch.ones_like(dw)
        plan.wgrad(dy, dw2, accumulate=True)
        torch.cuda.synchronize()
        torch.testing.assert_close(dw2, dw_ref + 1.0, **tol)
    finally:
        torch.backends.cudnn.allow_tf32 = prev


PAIR_CASES = [
    # N, C, H, W, F, k, stride, pad — 128-byte channel chunks in both precisions
    (5, 64, 13, 13, 256, 3, 1, 1),     # 845 pixels: 4 pair tiles, the last one with an all-padding second CTA
    (3, 128, 12, 10, 136, 3, 1, 1),    # F = 136: ragged filter tile, half of it beyond the matrix for the second CTA
    (2, 64, 17, 17, 128, 5, 1, 2),     # 5x5
    (3, 96, 13, 13, 160, 5, 1, 2),     # bf16: 64-byte chunks (two boxes per k-block)
    (4, 48, 14, 14, 384, 3, 1, 1),     # fp32: 64-byte chunks, bf16: 32-byte chunks; 384 filters = 1.5 pair tiles
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", PAIR_CASES)
def test_conv_cta_pair_tiles(case, dtype):
    """Forward and data gradient on CTA-pair (cta_group::2) tiles, forced, against the fp32 convolution."""
    N, C, H, W, F, k, s, p = case
    torch.manual_seed(sum(case))
    x = _nhwc((torch.randn(N, C, H, W, device=DEV) * 0.5).to(dtype))
    w32 = torch.randn(F, C * k * k, device=DEV) * 0.1
    b32 = torch.randn(F, device=DEV)
    plan = ConvPlan(N, C, H, W, F, k, s, p, dtype, DEV, need_dgrad=True)
    plan.pack(w32)
    OH, OW = plan.OH, plan.OW
    xr, wr = x.float(), w32.to(dtype).float().view(F, C, k, k)
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        ref = torch.relu(F_.conv2d(xr, wr, b32, stride=s, padding=p))
        dy = _nhwc((torch.randn(N, F, OH, OW, device=DEV) * 0.5).to(dtype))
        dx_ref = torch.nn.grad.conv2d_input(xr.shape, wr, dy.float(), stride=s, padding=p)
        for bn in (128, 256):
            y = _nhwc(torch.full((N, F, OH, OW), float("nan"), device=DEV, dtype=dtype))
            plan.forward(x, y.permute(0, 2, 3, 1).reshape(plan.P, F), bias=b32, act=G.ACT_RELU, tile=(bn, 2))
            torch.cuda.synchronize()
            torch.testing.assert_close(y.float(), ref, **_tols(dtype, C * k * k))
            dx = _nhwc(torch.full((N, C, H, W), float("nan"), device=DEV, dtype=dtype))
            plan.dgrad(dy, dx.permute(0, 2, 3, 1).reshape(N * H * W, C), tile=(bn, 2))
            torch.cuda.synchronize()
            torch.testing.assert_close(dx.float(), dx_ref, **_tols(dtype, F * k * k))
        # weight gradient: CTA pairs are the library's own choice for more than 128 filters (split-K over the pixels)
        dw = torch.full((F, C * k * k), float("nan"), device=DEV)
        plan.wgrad(dy, dw)
        torch.cuda.synchronize()
        dw_ref = torch.nn.grad.conv2d_weight(xr, wr.shape, dy.float(), stride=s, padding=p).reshape(F, -1)
        torch.testing.assert_close(dw, dw_ref, **_tols(dtype, plan.P))
        dw2 = torch.ones_like(dw)
        plan.wgrad(dy, dw2, accumulate=True)
        torch.cuda.synchronize()
        torch.testing.assert_close(dw2, dw_ref + 1.0, **_tols(dtype, plan.P))
    finally:
        torch.backends.cudnn.allow_tf32 = prev
