"""The convolutions of AlexNet (batch 256) and the CIFAR-10 net (batch 128): forward / data gradient / weight gradient
of this library's implicit-GEMM kernels against cuDNN through torch at equal precision (bf16, and fp32 with TF32
allowed), channels-last in both.  CUDA events around 10 launches after warm-up.
Usage: python tools/conv_bench.py   -> table (committed as profiles/conv_vs_cudnn.log)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn.functional as F_  # noqa: E402

from singa_b200.ops.conv import _mat, make_plan  # noqa: E402

dev = "cuda"
torch.backends.cudnn.allow_tf32 = True
torch.backends.cudnn.benchmark = True


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1000


LAYERS = [  # name, N, C, H, W, F, k, stride, pad, needs data gradient
    ("alexnet conv1", 256, 3, 227, 227, 96, 11, 4, 0, False),
    ("alexnet conv2", 256, 96, 27, 27, 256, 5, 1, 2, True),
    ("alexnet conv3", 256, 256, 13, 13, 384, 3, 1, 1, True),
    ("alexnet conv4", 256, 384, 13, 13, 384, 3, 1, 1, True),
    ("alexnet conv5", 256, 384, 13, 13, 256, 3, 1, 1, True),
    ("cifar conv1", 128, 3, 32, 32, 32, 5, 1, 2, False),
    ("cifar conv2", 128, 32, 16, 16, 32, 5, 1, 2, True),
    ("cifar conv3", 128, 32, 8, 8, 64, 5, 1, 2, True),
]
for dtype, name in ((torch.float32, "fp32 storage, tf32 products"), (torch.bfloat16, "bf16")):
    print(f"\n== {name}   us per launch (TFLOP/s):  ours | cuDNN")
    print(f"{'layer':16s} {'forward':>26s} {'data gradient':>26s} {'weight gradient':>26s}")
    for lname, N, C, H, W, F, k, s, p, need_dx in LAYERS:
        x = (torch.randn(N, C, H, W, device=dev) * 0.5).to(dtype).contiguous(memory_format=torch.channels_last)
        w32 = torch.randn(F, C * k * k, device=dev) * 0.05
        plan = make_plan(N, C, H, W, F, k, s, p, dtype, dev, need_dgrad=need_dx)   # what the layer uses
        plan.pack(w32)
        y = torch.empty(N, F, plan.OH, plan.OW, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last)
        dy = torch.randn_like(y)
        dw = torch.empty(F, C * k * k, device=dev)
        dx = torch.empty_like(x)
        y2d, dx2d = _mat(y, plan.P, F, F), _mat(dx, N * H * W, C, C)
        w4 = w32.view(F, C, k, k).to(dtype).contiguous(memory_format=torch.channels_last)
        flops = 2.0 * plan.P * F * C * k * k
        xin = x if C > 1 else x[:, 0]
        t_f = timeit(lambda: plan.forward(xin, y2d))
        t_fl = timeit(lambda: F_.conv2d(x, w4, None, stride=s, padding=p))
        t_w = timeit(lambda: plan.wgrad(dy, dw))
        t_wl = timeit(lambda: torch.nn.grad.conv2d_weight(x, w4.shape, dy, stride=s, padding=p))
        cell = lambda a, b: f"{a:7.0f} ({flops / a / 1e6:5.0f}) |{b:7.0f} ({flops / b / 1e6:5.0f})"   # noqa: E731
        if need_dx:
            t_d = timeit(lambda: plan.dgrad(dy, dx2d))
            t_dl = timeit(lambda: torch.nn.grad.conv2d_input(x.shape, w4, dy, stride=s, padding=p))
            dcell = cell(t_d, t_dl)
        else:
            dcell = f"{'(input layer)':>26s}"
        print(f"{lname:16s} {cell(t_f, t_fl)} {dcell} {cell(t_w, t_wl)}")
