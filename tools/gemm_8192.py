"""Three launches of the 8192^3 bf16 GEMM (CTA-pair kernel) for an ncu capture: python tools/gemm_8192.py"""
import torch
from singa_b200.ops import kernels as K

dev = torch.device("cuda:0")
M = N = Kd = 8192
a = torch.randn(M, Kd, device=dev).bfloat16()
b = torch.randn(N, Kd, device=dev).bfloat16()
o = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
bias = torch.randn(N, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(3):
    e0.record()
    K.gemm(a, b, M, N, Kd, out16=o, bias=bias, act=1)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"launch {i}: {ms:.3f} ms  {2 * M * N * Kd / ms / 1e9:.0f} TFLOP/s")
