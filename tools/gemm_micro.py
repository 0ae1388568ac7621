"""Isolated timing of the MNIST-MLP GEMM shapes for every tile configuration: 20 dependent launches captured in
one CUDA graph (so the host is out of the picture)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from singa_b200.ops import kernels as K  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
REP = 20
CFGS = [(0, 0), (64, 1), (128, 1), (256, 1), (128, 2), (256, 2)]


def bench(fn, n=20):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2):
            fn()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(REP):
                fn()
        for _ in range(3):
            g.replay()
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(n):
            g.replay()
        e1.record(s)
        s.synchronize()
    return e0.elapsed_time(e1) / (n * REP) * 1000


def sweep(label, call, N):
    out = []
    for cfg in CFGS:
        if cfg[0] > 64 and N <= cfg[0] // 2:
            out.append("     -")
            continue
        out.append(f"{bench(lambda: call(cfg)):6.2f}")
    print(f"{label:28s} " + " ".join(out))


B = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
DT = torch.float32 if (len(sys.argv) > 2 and sys.argv[2] == "fp32") else torch.bfloat16
Q = 8 if DT == torch.bfloat16 else 4
print(f"operands: {DT}")
print(f"{'case (us per launch)':28s} " + " ".join(f"{str(c[0]) + ('p' if c[1] == 2 else ''):>6s}" for c in CFGS))
for (kin, nout) in [(784, 2500), (2500, 2000), (2000, 1500), (1500, 1000), (1000, 500), (500, 10)]:
    ldk = (kin + Q - 1) // Q * Q
    ldn = (nout + Q - 1) // Q * Q
    x = torch.randn(B, ldk, device=dev).to(DT)
    w = torch.randn(kin, ldn, device=dev).to(DT)
    b = torch.randn(nout, device=dev)
    y = torch.empty(B, ldn, device=dev, dtype=DT)
    dy = torch.randn(B, ldn, device=dev).to(DT)
    dw = torch.empty(kin, (nout + 3) // 4 * 4, device=dev)
    dx = torch.empty(B, ldk, device=dev, dtype=DT)
    OUT = (lambda t: {"out32": t}) if DT == torch.float32 else (lambda t: {"out16": t})
    xs, ws, dys = x[:, :kin], w[:, :nout], dy[:, :nout]
    sweep(f"fwd  {B}x{nout}x{kin}", lambda c: K.gemm(xs, ws, B, nout, kin, a_mn=False, b_mn=True, bias=b, act=2, tile=c, **OUT(y[:, :nout])), nout)
    sweep(f"wgrd {kin}x{nout}x{B}", lambda c: K.gemm(xs, dys, kin, nout, B, a_mn=True, b_mn=True, out32=dw[:, :nout], tile=c), nout)
    sweep(f"dgrd {B}x{kin}x{nout}", lambda c: K.gemm(dys, ws, B, kin, nout, a_mn=False, b_mn=False, tile=c, **OUT(dx[:, :kin])), kin)
