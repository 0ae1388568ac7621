"""A few launches of ONE convolution (for ncu captures and A/B timing):
    python tools/conv_case.py <fp32|bf16> <fwd|dgrad|wgrad> <N> <C> <H> <W> <F> <k> <stride> <pad> [reps]
Prints the CUDA-event time per launch.  Example (AlexNet conv2): fp32 fwd 256 96 27 27 256 5 1 2"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from singa_b200.ops.conv import ConvPlan, _mat  # noqa: E402

dt = torch.float32 if sys.argv[1] == "fp32" else torch.bfloat16
what = sys.argv[2]
N, C, H, W, F, k, s, p = (int(v) for v in sys.argv[3:11])
reps = int(sys.argv[11]) if len(sys.argv) > 11 else 5
dev = "cuda"
torch.manual_seed(0)
x = (torch.randn(N, C, H, W, device=dev) * 0.5).to(dt).contiguous(memory_format=torch.channels_last)
w32 = torch.randn(F, C * k * k, device=dev) * 0.05
plan = ConvPlan(N, C, H, W, F, k, s, p, dt, dev, need_dgrad=what == "dgrad")
plan.pack(w32)
y = torch.empty(N, F, plan.OH, plan.OW, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
dy = torch.randn_like(y)
dw = torch.empty(F, C * k * k, device=dev)
dx = torch.empty_like(x)
y2d, dx2d = _mat(y, plan.P, F, F), _mat(dx, N * H * W, C, C)
xin = x if C > 1 else x[:, 0]
plan.forward(xin, y2d)
fn = {"fwd": lambda: plan.forward(xin, y2d), "dgrad": lambda: plan.dgrad(dy, dx2d),
      "wgrad": lambda: plan.wgrad(dy, dw)}[what]
for _ in range(3):
    fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    fn()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / reps * 1e3
flops = 2.0 * plan.P * F * C * k * k
print(f"{sys.argv[1]} {what} N={N} C={C} {H}x{W} F={F} k={k} s={s} p={p}: {us:.2f} us  {flops / us / 1e6:.1f} TFLOP/s")
