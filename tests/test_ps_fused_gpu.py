"""Fused push kernel on ONE GPU: the ranks are emulated as separate local buffers (every 'peer pointer' is a
local pointer), flags pre-set so nothing waits — numerics of reduce + update + broadcast + version publish."""
import ctypes

import pytest
import torch

from singa_b200.ops import native
from singa_b200.ops import reference as R

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mirror", [torch.bfloat16, torch.float32, None])
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("kind,kid", [("kSGD", 0), ("kAdaGrad", 2)])
def test_push_update_emulated(world, kind, kid, mirror):
    from singa_b200.ops._decl import Updater as CUpdater
    from singa_b200.parallel.fused_ps import PsPush
    lib = native.require()
    dev = "cuda"
    nchunks = 40
    n = nchunks * 1024
    grads = [torch.randn(n, device=dev) for _ in range(world)]
    mirrors = [torch.zeros(n, device=dev, dtype=mirror) for _ in range(world)] if mirror is not None else []
    ready = torch.full((world,), 2, dtype=torch.int64, device=dev)
    version = torch.zeros(world, dtype=torch.int64, device=dev)
    w = torch.randn(n, device=dev)
    h = torch.rand(n, device=dev)
    chunk_param = torch.zeros(nchunks, dtype=torch.int32, device=dev)
    chunk_param[20:] = 1
    lr_mult = torch.tensor([1.0, 0.5], device=dev)
    wd_mult = torch.tensor([1.0, 1.0], device=dev)
    scratch = torch.zeros(2, dtype=torch.int64, device=dev)
    wr, hr = w.clone(), h.clone()
    first, cnt = 8, 25       # only a sub-range is owned by "this server"
    a = PsPush()
    a.world = world
    for r in range(world):
        a.grad[r] = grads[r].data_ptr()
        a.w16[r] = mirrors[r].data_ptr() if mirrors else None
        a.ready[r] = ready[r:r + 1].data_ptr()
        a.version[r] = version[r:r + 1].data_ptr()
    a.w, a.h, a.u = w.data_ptr(), h.data_ptr(), None
    a.chunk_param, a.lr_mult, a.wd_mult = chunk_param.data_ptr(), lr_mult.data_ptr(), wd_mult.data_ptr()
    a.first_chunk, a.nchunks = first, cnt
    a.up = CUpdater(kid, 0.01, 0.9, 0.001, 1e-7, 0.9, 1.0 / world, None)
    a.step_dev, a.value_add = None, 0
    a.scratch = scratch.data_ptr()
    a.self_count = 1
    a.max_ctas = 7
    a.mirror_fp32 = 1 if mirror == torch.float32 else 0
    for launch in range(2):   # two launches: the arrival counter must be back to zero after the first
        native.check(lib.sg_ps_push_update(ctypes.byref(a), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        g = sum(grads) / world
        for pid, (lo, hi) in enumerate(((0, 20 * 1024), (20 * 1024, n))):
            lo2, hi2 = max(lo, first * 1024), min(hi, (first + cnt) * 1024)
            if lo2 < hi2:
                R.update_ref(kind, wr[lo2:hi2], g[lo2:hi2], hr[lo2:hi2], None, lr=0.01 * float(lr_mult[pid]),
                             momentum=0.9, wd=0.001, delta=1e-7, rho=0.9)
        torch.testing.assert_close(w, wr, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(h, hr, rtol=1e-5, atol=1e-6)
        lo, hi = first * 1024, (first + cnt) * 1024
        for m in mirrors:
            assert torch.equal(m[lo:hi], w[lo:hi].to(mirror))
            assert float(m[:lo].abs().sum()) == 0 and float(m[hi:].abs().sum()) == 0
        assert scratch.tolist() == [0, launch + 1]
        assert version.tolist() == [launch + 1] * world


def test_flag_set_wait():
    lib = native.require()
    flags = torch.zeros(3, dtype=torch.int64, device="cuda")
    step = torch.tensor([9], dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for i in range(3):
        native.check(lib.sg_ps_flag_set(ctypes.c_void_p(flags[i:i + 1].data_ptr()),
                                        ctypes.c_void_p(step.data_ptr()), 1, s))
    arr = (ctypes.c_void_p * 32)()
    for i in range(3):
        arr[i] = flags[i:i + 1].data_ptr()
    native.check(lib.sg_ps_flag_wait(arr, 3, ctypes.c_void_p(step.data_ptr()), 1, None, s))   # already satisfied
    seq = torch.tensor([10], dtype=torch.int64, device="cuda")
    native.check(lib.sg_ps_flag_wait(arr, 3, None, 0, ctypes.c_void_p(seq.data_ptr()), s))   # self-counting
    for i in range(3):
        native.check(lib.sg_ps_flag_bump(ctypes.c_void_p(flags[i:i + 1].data_ptr()),
                                         ctypes.c_void_p(seq.data_ptr()) if i == 0 else None, s))
    native.check(lib.sg_ps_flag_wait(arr, 3, None, 0, ctypes.c_void_p(seq.data_ptr()), s))
    torch.cuda.synchronize()
    assert flags.tolist() == [11, 11, 11] and int(seq.item()) == 11


def test_gemm_fused_wait_passes_when_version_reached():
    from singa_b200.ops import gemm as G
    dev = "cuda"
    a = torch.randn(256, 128, device=dev).to(torch.bfloat16)
    b = torch.randn(128, 256, device=dev).to(torch.bfloat16)
    out = torch.empty(256, 256, device=dev)
    flags = torch.tensor([3, 4], dtype=torch.int64, device=dev)
    seq = torch.tensor([3], dtype=torch.int64, device=dev)
    G.gemm(a, b, 256, 256, 128, b_mn=True, out32=out,
           wait=(torch.tensor([flags[0:1].data_ptr(), flags[1:2].data_ptr()], dtype=torch.int64, device=dev),
                 seq.data_ptr()))
    torch.cuda.synchronize()
    torch.testing.assert_close(out, a.float() @ b.float(), rtol=2e-3, atol=2e-2)


@pytest.mark.parametrize("config", ["lenet", "mnist_mlp"])
def test_fused_exchange_graph_capture_world1(config):
    """The whole fused step (forward with fused pull waits, backward, bucketed push) must capture into a CUDA graph —
    a one-bucket net (lenet) once failed because an idle comm stream was joined during capture."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, PYTHONPATH=root)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_PORT"):
        env.pop(k, None)
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "fused_world1_check.py"), config], env=env,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "PASS fused world=1" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
