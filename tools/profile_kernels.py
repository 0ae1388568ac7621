#!/usr/bin/env python
"""Stand-alone launches of the hot kernels for CUDA-event timing and for ncu captures (profiles/README.md).

  python tools/profile_kernels.py gemm        8192^3 + the MNIST-MLP layer shapes, CUDA-event timed, L2 flushed
  python tools/profile_kernels.py push        the fused parameter-server push kernel (ranks emulated on one GPU)
  python tools/profile_kernels.py membound    element-wise / reduction family achieved bandwidth
"""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from singa_b200.ops import gemm as G  # noqa: E402
from singa_b200.ops import kernels as K  # noqa: E402
from singa_b200.ops import native  # noqa: E402

DEV = "cuda"
PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
except Exception:  # noqa: BLE001
    pass
FLUSH = None


def timed(fn, iters=20, warm=3, flush=True):
    global FLUSH
    if FLUSH is None:
        FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    total = 0.0
    for _ in range(iters):
        if flush:
            FLUSH.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        total += a.elapsed_time(b)
    return total / iters


def gemm_cases():
    peak = PEAKS.get("bf16_tflops", 1590.0)
    which = "measured" if PEAKS else "fallback"
    rows = []
    shapes = [("8192^3 NT", 8192, 8192, 8192, False, False)]
    dims = [784, 2500, 2000, 1500, 1000, 500]
    for i in range(5):
        k, n = dims[i], dims[i + 1]
        shapes.append((f"mlp fc{i+1} fwd", 1000, n, k, False, True))
        shapes.append((f"mlp fc{i+1} dgrad", 1000, k, n, False, False))
        shapes.append((f"mlp fc{i+1} wgrad", k, n, 1000, True, True))
    shapes.append(("wide fwd 4096x16384x16384", 4096, 16384, 16384, False, True))
    for name, M, N, Kd, a_mn, b_mn in shapes:
        pad = lambda x: (x + 7) // 8 * 8
        a = (torch.randn(Kd if a_mn else M, pad(M if a_mn else Kd), device=DEV) * 0.05).to(torch.bfloat16)[:, :(M if a_mn else Kd)]
        b = (torch.randn(Kd if b_mn else N, pad(N if b_mn else Kd), device=DEV) * 0.05).to(torch.bfloat16)[:, :(N if b_mn else Kd)]
        out = torch.empty(M, pad(N), device=DEV, dtype=torch.bfloat16)[:, :N]
        ms = timed(lambda: G.gemm(a, b, M, N, Kd, a_mn=a_mn, b_mn=b_mn, out16=out))
        cfg = G.tuned_configs().get((M, N, Kd, a_mn, b_mn))
        tf = 2.0 * M * N * Kd / ms / 1e9
        rows.append((name, M, N, Kd, ms * 1e3, tf, tf / peak))
        print(f"{name:28s} M={M:5d} N={N:5d} K={Kd:5d}  {ms*1e3:8.1f} us  {tf:7.1f} TFLOP/s  {tf/peak*100:5.1f}% of {which} cuBLAS bf16 peak  tile={cfg}")
    return rows


def push_case():
    from singa_b200.ops._decl import Updater as CUpdater
    from singa_b200.parallel.fused_ps import PsPush
    lib = native.require()
    world, nchunks = 2, 11700            # the MNIST-MLP arena: 11.98 M elements
    n = nchunks * 1024
    grads = [torch.randn(n, device=DEV) for _ in range(world)]
    mirrors = [torch.zeros(n, device=DEV, dtype=torch.bfloat16) for _ in range(world)]
    ready = torch.full((world,), 1 << 40, dtype=torch.int64, device=DEV)
    version = torch.zeros(world, dtype=torch.int64, device=DEV)
    w, h = torch.randn(n, device=DEV), torch.zeros(n, device=DEV)
    cp = torch.zeros(nchunks, dtype=torch.int32, device=DEV)
    one = torch.ones(1, device=DEV)
    scratch = torch.zeros(2, dtype=torch.int64, device=DEV)
    a = PsPush()
    a.world = world
    for r in range(world):
        a.grad[r], a.w16[r] = grads[r].data_ptr(), mirrors[r].data_ptr()
        a.ready[r], a.version[r] = ready[r:r + 1].data_ptr(), version[r:r + 1].data_ptr()
    a.w, a.h, a.u = w.data_ptr(), h.data_ptr(), None
    a.chunk_param, a.lr_mult, a.wd_mult = cp.data_ptr(), one.data_ptr(), one.data_ptr()
    a.first_chunk, a.nchunks = 0, nchunks
    a.up = CUpdater(0, 0.01, 0.9, 0.0, 1e-7, 0.9, 0.5, None)
    a.scratch, a.self_count, a.max_ctas = scratch.data_ptr(), 1, 592
    s = torch.cuda.current_stream().cuda_stream
    ms = timed(lambda: native.check(lib.sg_ps_push_update(ctypes.byref(a), s)))
    bytes_moved = n * (4 * world + 4 * 2 + 4 * 2 + 2 * world)   # grads + w rw + h rw + mirrors
    bw = bytes_moved / ms / 1e6
    peak = PEAKS.get("hbm_gbs", 6650.0)
    print(f"ps_push_update world={world} (emulated on one GPU) n={n}: {ms*1e3:.1f} us, {bw:.0f} GB/s "
          f"= {bw/peak*100:.1f}% of {'measured' if PEAKS else 'fallback'} copy bandwidth")


def membound_cases():
    """Achieved bandwidth of the memory-bound kernels, counted on the COMPULSORY traffic (every input read once, every
    output written once), against the measured copy bandwidth."""
    peak = PEAKS.get("hbm_gbs", 6650.0)

    def line(name, ms, nbytes):
        print(f"{name:34s} {ms*1e3:8.1f} us  {nbytes/ms/1e6:7.0f} GB/s  {nbytes/ms/1e6/peak*100:5.1f}% of copy bandwidth")

    cl = lambda t: t.contiguous(memory_format=torch.channels_last)   # noqa: E731
    for dt, dn in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
        eb = 4 if dt == torch.float32 else 2
        n = 128 << 20
        x = torch.randn(n, device=DEV).to(dt)
        y = torch.empty_like(x)
        line(f"act_fwd stanh {dn}", timed(lambda: K.act_fwd(x, y, 2), iters=5, flush=False), 2 * eb * n)
        line(f"act_bwd stanh {dn}", timed(lambda: K.act_bwd(x, y, y, 2), iters=5, flush=False), 3 * eb * n)
        line(f"dropout {dn}", timed(lambda: K.dropout(x, y, 0.5, 1), iters=5, flush=False), 2 * eb * n)
        del x, y
        # AlexNet pool1 (after conv1): 256 x 96 x 55 x 55 -> 27 x 27, max 3/2
        N, C, H, W, k, s = 256, 96, 55, 55, 3, 2
        OH = OW = 27
        xi = cl(torch.randn(N, C, H, W, device=DEV).to(dt))
        yo = cl(torch.empty(N, C, OH, OW, device=DEV, dtype=dt))
        dy = cl(torch.randn(N, C, OH, OW, device=DEV).to(dt))
        dx = cl(torch.empty_like(xi))
        nin, nout = xi.numel() * eb, yo.numel() * eb
        line(f"pool_fwd max 3/2 96x55x55 {dn}",
             timed(lambda: K.pool_fwd(xi, yo, N, H, W, C, OH, OW, k, s, 0, True), iters=5), nin + nout)
        line(f"pool_bwd max 3/2 96x55x55 {dn}",
             timed(lambda: K.pool_bwd(xi, yo, dy, dx, N, H, W, C, OH, OW, k, s, 0, True), iters=5), 2 * nin + 2 * nout)
        del xi, dx
        # AlexNet norm1: 256 x 96 x 27 x 27, local size 5
        norm = torch.empty(N, OH, OW, C, device=DEV)
        y2 = cl(torch.empty_like(yo))
        px = N * OH * OW
        yo.normal_()
        line(f"lrn_fwd 5 96x27x27 {dn}", timed(lambda: K.lrn_fwd(yo, y2, norm, px, C, 5, 1e-4, 0.75, 1.0), iters=5),
             2 * nout + 4 * yo.numel())
        line(f"lrn_bwd 5 96x27x27 {dn}", timed(lambda: K.lrn_bwd(yo, norm, dy, y2, px, C, 5, 1e-4, 0.75), iters=5),
             3 * nout + 4 * yo.numel())
        # bias gradient of AlexNet conv1 / conv2: column sums of [pixels][F]
        for F, pix in ((96, 256 * 55 * 55), (256, 256 * 27 * 27)):
            m = torch.randn(pix, F, device=DEV).to(dt)
            out = torch.zeros(F, device=DEV)
            line(f"colsum [{pix}][{F}] {dn}", timed(lambda: K.colsum(m, pix, F, F, out), iters=5), pix * F * eb)
            del m
        # AlexNet input: space-to-depth of 256 x 227 x 227 x 3 (stride 4, 3x3 taps -> 57 x 57 x 48)
        xin = torch.randn(256, 227, 227, 3, device=DEV).to(dt)
        xs = torch.empty(256, 57, 57, 48, device=DEV, dtype=dt)
        line(f"space_to_depth 227x227x3 /4 {dn}",
             timed(lambda: K.space_to_depth(xin, xs, 256, 227, 227, 3, 57, 57, 48, 4, 0), iters=5),
             (xin.numel() + xs.numel()) * eb)
        del xin, xs
    rows, cols = 65536, 4096
    m = torch.randn(rows, cols, device=DEV).to(torch.bfloat16)
    out = torch.zeros(cols, device=DEV)
    line("colsum [65536][4096] bf16", timed(lambda: K.colsum(m, rows, cols, cols, out), iters=5, flush=False), 2 * rows * cols)
    del m
    nch = 200000
    w = torch.randn(nch * 1024, device=DEV); g = torch.randn_like(w); h = torch.zeros_like(w)
    w16 = torch.zeros(nch * 1024, device=DEV, dtype=torch.bfloat16)
    cp = torch.zeros(nch, dtype=torch.int32, device=DEV); one = torch.ones(1, device=DEV)
    ms = timed(lambda: K.update_arena(w, g, h, None, w16, cp, one, one, 0, nch, kind=0, lr=0.01, momentum=0.9),
               iters=5, flush=False)
    line("update_arena sgd+mom (+bf16 mirror)", ms, nch * 1024 * (4 * 2 + 4 + 4 * 2 + 2))


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "gemm"
    if len(sys.argv) > 2:      # second argument: tile policy (0 single CTA, 1 CTA pairs, 2 auto)
        native.require().sg_gemm_set_pair(int(sys.argv[2]))
        print(f"# sg_gemm_set_pair({sys.argv[2]})")
    {"gemm": gemm_cases, "push": push_case, "membound": membound_cases}[what]()
