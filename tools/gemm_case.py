"""A few launches of ONE GEMM problem (for ncu captures and A/B timing):
    python tools/gemm_case.py <fp32|bf16> <M> <N> <K> <a_mn 0/1> <b_mn 0/1> [bn] [pair 0/1/2] [reps]
Prints the CUDA-event time per launch."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from singa_b200.ops import kernels as K  # noqa: E402

dt = torch.float32 if sys.argv[1] == "fp32" else torch.bfloat16
M, N, Kd, a_mn, b_mn = (int(v) for v in sys.argv[2:7])
bn = int(sys.argv[7]) if len(sys.argv) > 7 else 0
pair = int(sys.argv[8]) if len(sys.argv) > 8 else 0
reps = int(sys.argv[9]) if len(sys.argv) > 9 else 5
q = 8 if dt == torch.bfloat16 else 4
pad = lambda n: (n + q - 1) // q * q   # noqa: E731
dev = "cuda"
torch.manual_seed(0)
a = (torch.randn(Kd if a_mn else M, pad(M if a_mn else Kd), device=dev) * 0.05).to(dt)[:, :(M if a_mn else Kd)]
b = (torch.randn(Kd if b_mn else N, pad(N if b_mn else Kd), device=dev) * 0.05).to(dt)[:, :(N if b_mn else Kd)]
out = torch.empty(M, pad(N), device=dev, dtype=dt)[:, :N]
kw = {"out32": out} if dt == torch.float32 else {"out16": out}
tile = (bn, pair) if (bn or pair) else None
for _ in range(3):
    K.gemm(a, b, M, N, Kd, a_mn=bool(a_mn), b_mn=bool(b_mn), tile=tile, **kw)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    K.gemm(a, b, M, N, Kd, a_mn=bool(a_mn), b_mn=bool(b_mn), tile=tile, **kw)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / reps * 1e3
print(f"{sys.argv[1]} {M}x{N}x{Kd} a_mn={a_mn} b_mn={b_mn} tile={tile}: {us:.2f} us  {2.0 * M * N * Kd / us / 1e6:.1f} TFLOP/s")
