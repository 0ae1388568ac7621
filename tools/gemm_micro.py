"""Isolated timing of the MNIST-MLP GEMM shapes (CUDA events, back-to-back launches)."""
import sys
import torch
from singa_b200.ops import kernels as K

dev = torch.device("cuda:0")
torch.manual_seed(0)

def bench(name, fn, n=300):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:40s} {e0.elapsed_time(e1) / n * 1000:8.2f} us")

B = 1000
for (kin, nout) in [(784, 2500), (2500, 2000), (2000, 1500), (1500, 1000), (1000, 500), (500, 10)]:
    ldk = (kin + 7) // 8 * 8
    ldn = (nout + 7) // 8 * 8
    x = torch.randn(B, ldk, device=dev).bfloat16()
    w = torch.randn(kin, ldn, device=dev).bfloat16()
    b = torch.randn(nout, device=dev)
    y = torch.empty(B, ldn, device=dev, dtype=torch.bfloat16)
    dy = torch.randn(B, ldn, device=dev).bfloat16()
    dw = torch.empty(kin, ldn, device=dev)
    dx = torch.empty(B, ldk, device=dev, dtype=torch.bfloat16)
    bench(f"fwd  {B}x{nout}x{kin}", lambda: K.gemm(x[:, :kin], w[:, :nout], B, nout, kin, a_mn=False, b_mn=True, out16=y[:, :nout], bias=b, act=1))
    bench(f"wgrd {kin}x{nout}x{B}", lambda: K.gemm(x[:, :kin], dy[:, :nout], kin, nout, B, a_mn=True, b_mn=True, out32=dw[:, :nout]))
    bench(f"dgrd {B}x{kin}x{nout}", lambda: K.gemm(dy[:, :nout], w[:, :nout], B, kin, nout, a_mn=False, b_mn=False, out16=dx[:, :kin]))
