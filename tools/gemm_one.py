"""Run a few launches of chosen GEMM problems (for ncu captures): python tools/gemm_one.py"""
import torch
from singa_b200.ops import kernels as K

dev = torch.device("cuda:0")
torch.manual_seed(0)
B = 1000


def mk(kin, nout):
    ldk, ldn = (kin + 7) // 8 * 8, (nout + 7) // 8 * 8
    x = torch.randn(B, ldk, device=dev).bfloat16()
    w = torch.randn(kin, ldn, device=dev).bfloat16()
    dy = torch.randn(B, ldn, device=dev).bfloat16()
    dw = torch.empty(kin, ldn, device=dev)
    dx = torch.empty(B, ldk, device=dev, dtype=torch.bfloat16)
    y = torch.empty(B, ldn, device=dev, dtype=torch.bfloat16)
    b = torch.randn(nout, device=dev)
    return x[:, :kin], w[:, :nout], dy[:, :nout], dw[:, :nout], dx[:, :kin], y[:, :nout], b


for _ in range(3):
    x, w, dy, dw, dx, y, b = mk(500, 10)
    K.gemm(dy, w, B, 500, 10, a_mn=False, b_mn=False, out16=dx)            # tiny: fixed overhead
    x, w, dy, dw, dx, y, b = mk(2500, 2000)
    K.gemm(x, dy, 2500, 2000, B, a_mn=True, b_mn=True, out32=dw)           # wgrad, 160 tiles of 128x256
    K.gemm(x, w, B, 2000, 2500, a_mn=False, b_mn=True, out16=y, bias=b, act=2)   # forward
    torch.cuda.synchronize()
print("ok")
