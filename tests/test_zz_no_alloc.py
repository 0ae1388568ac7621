"""The allocation-free push entry point against the ordinary one.

``sg_ps_push_update_noalloc`` exists for callers that are not allowed to create device memory (an embedding
application whose memory manager owns all of HBM, a callback in which cudaMalloc is forbidden, a control loop
that must never wait on the allocator): its scratch words are a device array compiled into the library instead
of a caller-provided buffer.  Here it is called through the C ABI (ctypes on the shared object) next to
``sg_ps_push_update`` on identical inputs, twice in a row on the same slot — the second call sees whatever the
first one left in the shared scratch — and every output (fp32 master weights, history, every rank's bf16
mirror, the version flags) must agree within the tolerance the other push tests use.  The file sorts last so a
run that stops at the first failure has already executed every other test.
"""
import ctypes

import pytest
import torch

from singa_b200 import build as B

pytestmark = pytest.mark.gpu


def _lib():
    lib = ctypes.CDLL(str(B.LIB_PATH))
    lib.sg_last_error.restype = ctypes.c_char_p
    for name in ("sg_ps_push_update", "sg_ps_push_update_noalloc"):
        getattr(lib, name).restype = ctypes.c_int
    lib.sg_ps_push_update.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.sg_ps_push_update_noalloc.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    return lib


def _args(world, grads, mirrors, ready, version, w, h, chunk_param, lr_mult, wd_mult, first, cnt, step_dev,
          scratch):
    from singa_b200.ops._decl import Updater as CUpdater
    from singa_b200.parallel.fused_ps import PsPush
    a = PsPush()
    a.world = world
    for r in range(world):
        a.grad[r] = grads[r].data_ptr()
        a.w16[r] = mirrors[r].data_ptr()
        a.ready[r] = ready[r:r + 1].data_ptr()
        a.version[r] = version[r:r + 1].data_ptr()
    a.w, a.h, a.u = w.data_ptr(), h.data_ptr(), None
    a.chunk_param, a.lr_mult, a.wd_mult = chunk_param.data_ptr(), lr_mult.data_ptr(), wd_mult.data_ptr()
    a.first_chunk, a.nchunks = first, cnt
    a.up = CUpdater(0, 0.01, 0.9, 0.001, 1e-7, 0.9, 1.0 / world, None)      # SGD with momentum
    a.step_dev, a.value_add = step_dev.data_ptr(), 1                         # explicit flag value
    a.scratch = scratch.data_ptr() if scratch is not None else None
    a.self_count = 0
    a.max_ctas = 9
    return a


def test_noalloc_push_matches_regular_push_twice():
    lib = _lib()
    dev = "cuda"
    world, nchunks, first, cnt, slot = 2, 48, 5, 37, 3
    n = nchunks * 1024
    torch.manual_seed(1234)
    chunk_param = torch.zeros(nchunks, dtype=torch.int32, device=dev)
    chunk_param[24:] = 1
    lr_mult = torch.tensor([1.0, 2.0], device=dev)
    wd_mult = torch.tensor([1.0, 0.0], device=dev)
    w0, h0 = torch.randn(n, device=dev), torch.zeros(n, device=dev)
    state = {}
    for tag in ("regular", "noalloc"):
        state[tag] = dict(w=w0.clone(), h=h0.clone(),
                          mirrors=[torch.zeros(n, device=dev, dtype=torch.bfloat16) for _ in range(world)],
                          version=torch.zeros(world, dtype=torch.int64, device=dev))
    ready = torch.full((world,), 100, dtype=torch.int64, device=dev)     # gradients are "ready": nothing waits
    scratch = torch.zeros(2, dtype=torch.int64, device=dev)
    step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):
        grads = [torch.randn(n, device=dev) for _ in range(world)]
        step_dev.fill_(call)
        st = state["regular"]
        a = _args(world, grads, st["mirrors"], ready, st["version"], st["w"], st["h"], chunk_param, lr_mult,
                  wd_mult, first, cnt, step_dev, scratch)
        rc = lib.sg_ps_push_update(ctypes.byref(a), stream)
        assert rc == 0, lib.sg_last_error()
        st = state["noalloc"]
        b = _args(world, grads, st["mirrors"], ready, st["version"], st["w"], st["h"], chunk_param, lr_mult,
                  wd_mult, first, cnt, step_dev, None)
        rc = lib.sg_ps_push_update_noalloc(ctypes.byref(b), slot, stream)
        assert rc == 0, lib.sg_last_error()
        torch.cuda.synchronize()
        ra, rb = state["regular"], state["noalloc"]
        torch.testing.assert_close(rb["w"], ra["w"], rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(rb["h"], ra["h"], rtol=1e-5, atol=1e-6)
        for ma, mb in zip(ra["mirrors"], rb["mirrors"]):
            assert torch.equal(ma, mb)
        assert rb["version"].tolist() == ra["version"].tolist() == [call + 1] * world
        assert not torch.equal(ra["w"], w0)                                # the update really happened


def test_noalloc_push_rejects_out_of_range_slot_and_self_count():
    lib = _lib()
    dev = "cuda"
    n = 4 * 1024
    t = torch.zeros(n, device=dev)
    m = torch.zeros(n, device=dev, dtype=torch.bfloat16)
    flags = torch.zeros(1, dtype=torch.int64, device=dev)
    cp = torch.zeros(4, dtype=torch.int32, device=dev)
    one = torch.ones(1, device=dev)
    a = _args(1, [t], [m], flags, flags.clone(), t.clone(), t.clone(), cp, one, one, 0, 4, flags.clone(), None)
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.sg_ps_push_update_noalloc(ctypes.byref(a), 64, stream) != 0
    assert b"outside the built-in scratch" in lib.sg_last_error()
    assert lib.sg_ps_push_update_noalloc(ctypes.byref(a), -1, stream) != 0
    a.self_count = 1
    assert lib.sg_ps_push_update_noalloc(ctypes.byref(a), 0, stream) != 0
    assert b"self-counted" in lib.sg_last_error()
    torch.cuda.synchronize()
