"""Every GEMM of the MNIST MLP step (6 layers x forward / weight gradient / data gradient, batch 1000) and a large
square problem: this library's tcgen05 kernel (default tile choice, fused epilogue included) against cuBLAS through
torch.matmul on the same operands, at equal precision — bf16 operands, and fp32 operands multiplied as tf32.
20 dependent launches per CUDA graph, graph replayed 20 times, CUDA events (host out of the picture).
Usage: python tools/gemm_vs_cublas.py [batch]   -> table on stdout (committed as profiles/gemm_vs_cublas.log)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from singa_b200.ops import kernels as K  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
REP = 20


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


def pad(n, q):
    return (n + q - 1) // q * q


def mk(rows, cols, dtype):
    q = 8 if dtype == torch.bfloat16 else 4
    return (torch.randn(rows, pad(cols, q), device=dev) * 0.05).to(dtype)[:, :cols]


def row(label, flops, t_ours, t_lib):
    print(f"{label:30s} {t_ours:8.2f} {flops / t_ours / 1e6:8.1f}   {t_lib:8.2f} {flops / t_lib / 1e6:8.1f}   {t_lib / t_ours:6.2f}x")


B = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
for dtype, name in ((torch.float32, "fp32 operands, tf32 products"), (torch.bfloat16, "bf16 operands")):
    torch.backends.cuda.matmul.allow_tf32 = True
    print(f"\n== {name}   (us per launch, TFLOP/s; ours = tcgen05 kernel incl. fused epilogue, lib = cuBLAS via torch.matmul)")
    print(f"{'GEMM (M x N x K)':30s} {'ours us':>8s} {'TFLOP/s':>8s}   {'lib us':>8s} {'TFLOP/s':>8s}   {'lib/ours':>7s}")
    tot_o = tot_l = 0.0
    for (kin, nout) in [(784, 2500), (2500, 2000), (2000, 1500), (1500, 1000), (1000, 500), (500, 10)]:
        x, w, dy = mk(B, kin, dtype), mk(kin, nout, dtype), mk(B, nout, dtype)
        b = torch.randn(nout, device=dev)
        y, dx = mk(B, nout, dtype), mk(B, kin, dtype)
        dw = torch.empty(kin, pad(nout, 4), device=dev)[:, :nout]
        out = (lambda t: {"out32": t}) if dtype == torch.float32 else (lambda t: {"out16": t})
        # the library gets the same operands: row strides padded to 16 bytes (unpadded bf16 rows of 2500 elements are
        # misaligned for cuBLAS and run 2-3x slower), outputs contiguous
        yl, dxl = torch.empty(B, nout, device=dev, dtype=dtype), torch.empty(B, kin, device=dev, dtype=dtype)
        dwl = torch.empty(kin, nout, device=dev, dtype=dtype)
        xc, wc, dyc = x, w, dy
        cases = [
            (f"fwd  {B}x{nout}x{kin}", 2.0 * B * nout * kin,
             lambda: K.gemm(x, w, B, nout, kin, a_mn=False, b_mn=True, bias=b, act=2, **out(y)),
             lambda: torch.matmul(xc, wc, out=yl)),
            (f"wgrd {kin}x{nout}x{B}", 2.0 * B * nout * kin,
             lambda: K.gemm(x, dy, kin, nout, B, a_mn=True, b_mn=True, out32=dw),
             lambda: torch.matmul(xc.t(), dyc, out=dwl)),
            (f"dgrd {B}x{kin}x{nout}", 2.0 * B * nout * kin,
             lambda: K.gemm(dy, w, B, kin, nout, a_mn=False, b_mn=False, **out(dx)),
             lambda: torch.matmul(dyc, wc.t(), out=dxl)),
        ]
        for label, fl, ours, lib in cases:
            to, tl = bench(ours), bench(lib)
            tot_o += to
            tot_l += tl
            row(label, fl, to, tl)
    print(f"{'sum of the 18 GEMMs':30s} {tot_o:8.2f} {'':8s}   {tot_l:8.2f} {'':8s}   {tot_l / tot_o:6.2f}x")
    n = 8192
    a, bb = mk(n, n, dtype), mk(n, n, dtype)
    o = mk(n, n, dtype)
    ol = torch.empty(n, n, device=dev, dtype=dtype)
    out = {"out32": o} if dtype == torch.float32 else {"out16": o}
    to = bench(lambda: K.gemm(a, bb, n, n, n, **out), n=3)
    tl = bench(lambda: torch.matmul(a, bb.t(), out=ol), n=3)
    row(f"square {n}^3 (NT)", 2.0 * n ** 3, to, tl)
