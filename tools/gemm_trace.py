"""Phase timeline of CTA 0 of the GEMM kernel (needs a library built with -DSG_GEMM_TRACE)."""
import ctypes
import torch
from singa_b200.ops import kernels as K, native

lib = native.require()
lib.sg_gemm_trace_read.argtypes = [ctypes.c_void_p]
dev = torch.device("cuda:0")
B = 1000
NAMES = ["entry", "setup done", "1st TMA issued", "all TMA issued", "1st stage landed", "last MMA issued",
         "accumulator ready", "epilogue done", "final sync", "tmem freed",
         "chunk0 tmem loaded", "chunk0 done", "chunk1 tmem loaded", "chunk1 done",
         "h0 start", "h0 staged", "h0 stored", "h1 stored"]


def trace(label, fn):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(10):
                fn()
        for _ in range(3):
            g.replay()
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(20):
            g.replay()
        e1.record(s)
        s.synchronize()
    per = e0.elapsed_time(e1) / 200 * 1000
    buf = (ctypes.c_ulonglong * 32)()
    assert lib.sg_gemm_trace_read(ctypes.addressof(buf)) == 0
    t = list(buf)[:18]
    print(f"{label}: {per:.2f} us per launch; CTA 0 lifetime {t[9] - t[0]} cycles")
    for i in range(1, 18):
        print(f"    {NAMES[i]:20s} +{t[i] - t[0]:7d}")


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


x, w, dy, dw, dx, y, b = mk(500, 10)
trace("dgrad 1000x500x10 (1 k-block, BN=64)", lambda: K.gemm(dy, w, B, 500, 10, a_mn=False, b_mn=False, out16=dx))
x, w, dy, dw, dx, y, b = mk(1000, 500)
trace("fwd 1000x500x1000", lambda: K.gemm(x, w, B, 500, 1000, a_mn=False, b_mn=True, out16=y, bias=b, act=2))
trace("wgrad 1000x500x1000 fp32", lambda: K.gemm(x, dy, 1000, 500, B, a_mn=True, b_mn=True, out32=dw))
x, w, dy, dw, dx, y, b = mk(2500, 2000)
trace("fwd 1000x2000x2500", lambda: K.gemm(x, w, B, 2000, 2500, a_mn=False, b_mn=True, out16=y, bias=b, act=2))
trace("wgrad 2500x2000x1000 fp32", lambda: K.gemm(x, dy, 2500, 2000, B, a_mn=True, b_mn=True, out32=dw))
trace("dgrad 1000x2500x2000", lambda: K.gemm(dy, w, B, 2500, 2000, a_mn=False, b_mn=False, out16=dx))
