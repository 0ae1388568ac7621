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
    peak = PEAKS.get("hbm_gbs", 6650.0)
    n = 256 << 20
    x = torch.randn(n, device=DEV).to(torch.bfloat16)
    y = torch.empty_like(x)
    for name, fn, nbytes in [
        ("act_fwd stanh bf16", lambda: K.act_fwd(x, y, 2), 4 * n),
        ("act_bwd stanh bf16", lambda: K.act_bwd(x, y, y, 2), 6 * n),
        ("dropout bf16", lambda: K.dropout(x, y, 0.5, 1), 4 * n),
    ]:
        ms = timed(fn, iters=5, flush=False)
        print(f"{name:22s} {ms*1e3:8.1f} us  {nbytes/ms/1e6:7.0f} GB/s  {nbytes/ms/1e6/peak*100:5.1f}% of copy bandwidth")
    rows, cols = 65536, 4096
    m = torch.randn(rows, cols, device=DEV).to(torch.bfloat16)
    out = torch.zeros(cols, device=DEV)
    ms = timed(lambda: K.colsum(m, rows, cols, cols, out), iters=5, flush=False)
    print(f"{'colsum bf16':22s} {ms*1e3:8.1f} us  {2*rows*cols/ms/1e6:7.0f} GB/s")
    nch = 200000
    w = torch.randn(nch * 1024, device=DEV); g = torch.randn_like(w); h = torch.zeros_like(w)
    w16 = torch.zeros(nch * 1024, device=DEV, dtype=torch.bfloat16)
    cp = torch.zeros(nch, dtype=torch.int32, device=DEV); one = torch.ones(1, device=DEV)
    ms = timed(lambda: K.update_arena(w, g, h, None, w16, cp, one, one, 0, nch, kind=0, lr=0.01, momentum=0.9),
               iters=5, flush=False)
    nb = nch * 1024 * (4 * 2 + 4 + 4 * 2 + 2)
    print(f"{'update_arena sgd+mom':22s} {ms*1e3:8.1f} us  {nb/ms/1e6:7.0f} GB/s  {nb/ms/1e6/peak*100:5.1f}% of copy bandwidth")


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "gemm"
    if len(sys.argv) > 2:      # second argument: tile policy (0 single CTA, 1 CTA pairs, 2 auto)
        native.require().sg_gemm_set_pair(int(sys.argv[2]))
        print(f"# sg_gemm_set_pair({sys.argv[2]})")
    {"gemm": gemm_cases, "push": push_case, "membound": membound_cases}[what]()
