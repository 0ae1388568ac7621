"""tcgen05 GEMM vs a plain PyTorch fp32 reference of the same op (inputs rounded to bf16 in both)."""
import pytest
import torch

from singa_b200.ops import gemm as G
from singa_b200.ops import native

pytestmark = pytest.mark.gpu


def _pad8(n):
    return (n + 7) // 8 * 8


def _mk(rows, cols, dev, scale=1.0):
    """bf16 [rows][cols] view of a padded buffer (ld multiple of 8, padding zero)."""
    buf = torch.zeros(rows, _pad8(cols), device=dev, dtype=torch.bfloat16)
    buf[:, :cols] = (torch.randn(rows, cols, device=dev) * scale).to(torch.bfloat16)
    return buf[:, :cols]


def _run(M, N, K, a_mn, b_mn, bn=0, **kw):
    dev = "cuda"
    torch.manual_seed(M * 7 + N * 3 + K)
    a = _mk(K, M, dev) if a_mn else _mk(M, K, dev)
    b = _mk(K, N, dev) if b_mn else _mk(N, K, dev)
    out32 = torch.full((M, _pad8(N) if N % 4 else N), float("nan"), device=dev)[:, :N]
    native.require().sg_gemm_set_tile(bn)
    try:
        G.gemm(a, b, M, N, K, a_mn=a_mn, b_mn=b_mn, out32=out32, **kw)
    finally:
        native.require().sg_gemm_set_tile(0)
    torch.cuda.synchronize()
    kwr = {k: v for k, v in kw.items() if k in ("bias", "act", "grad_act", "aux")}
    ref = G.gemm_ref(a, b, a_mn, b_mn, **kwr)
    return out32, ref


@pytest.mark.parametrize("bn", [64, 128, 256])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_majors_tiles(a_mn, b_mn, bn):
    out, ref = _run(256, 512, 384, a_mn, b_mn, bn)
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-2)


@pytest.mark.parametrize("M,N,K", [(1000, 2500, 784), (1000, 10, 500), (64, 500, 800), (1000, 500, 1000),
                                   (130, 70, 72), (1, 8, 8), (777, 1500, 2000)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (False, False), (True, True)])
def test_gemm_ragged(M, N, K, a_mn, b_mn):
    out, ref = _run(M, N, K, a_mn, b_mn)
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=3e-2)


def test_gemm_epilogue_bias_act():
    M, N, K = 300, 200, 136
    bias = torch.randn(N, device="cuda")
    for act in (G.ACT_RELU, G.ACT_STANH):
        out, ref = _run(M, N, K, False, True, bias=bias, act=act)
        torch.testing.assert_close(out, ref, rtol=5e-3, atol=3e-2)


def test_gemm_epilogue_gradact_bf16_out_accumulate():
    M, N, K = 257, 192, 320
    dev = "cuda"
    a, b = _mk(M, K, dev, 0.5), _mk(N, K, dev, 0.5)
    aux = torch.tanh(torch.randn(M, N, device=dev)).to(torch.bfloat16)
    out16 = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
    out32 = torch.ones(M, N, device=dev)
    G.gemm(a, b, M, N, K, out32=out32, out16=out16, grad_act=G.ACT_STANH, aux=aux, accumulate=True)
    torch.cuda.synchronize()
    ref = G.gemm_ref(a, b, grad_act=G.ACT_STANH, aux=aux) + 1.0
    torch.testing.assert_close(out32, ref, rtol=2e-3, atol=2e-2)
    torch.testing.assert_close(out16.float(), ref, rtol=2e-2, atol=5e-2)
    # fp32 aux, relu gradient
    aux32 = torch.randn(M, N, device=dev)
    out32 = torch.empty(M, N, device=dev)
    G.gemm(a, b, M, N, K, out32=out32, grad_act=G.ACT_RELU, aux=aux32)
    torch.cuda.synchronize()
    torch.testing.assert_close(out32, G.gemm_ref(a, b, grad_act=G.ACT_RELU, aux=aux32), rtol=2e-3, atol=2e-2)


def test_gemm_perm_columns():
    # conv wgrad: GEMM columns ordered (ky,kx,c) are stored in the reference's (c,ky,kx) order
    F, C, kk, P = 48, 6, 9, 640
    dev = "cuda"
    a, b = _mk(P, F, dev), _mk(P, C * kk, dev)      # both stored [K=P][mn]
    out = torch.zeros(F, C * kk, device=dev)
    G.gemm(a, b, F, C * kk, P, a_mn=True, b_mn=True, out32=out, perm=(C, kk))
    torch.cuda.synchronize()
    ref = G.gemm_ref(a, b, True, True, perm=(C, kk))
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=3e-2)


def test_gemm_large_timing():
    """8192^3: sanity on numerics of a big persistent run plus a throughput print (not asserted)."""
    M = N = K = 8192
    dev = "cuda"
    a = (torch.randn(M, K, device=dev) * 0.05).to(torch.bfloat16)
    b = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
    out16 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        G.gemm(a, b, M, N, K, out16=out16)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        G.gemm(a, b, M, N, K, out16=out16)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"\n[gemm 8192^3 bf16] {ms:.3f} ms  {2 * M * N * K / ms / 1e9:.1f} TFLOP/s")
    ref = (a[:256].float() @ b[:512].float().t())
    torch.testing.assert_close(out16[:256, :512].float(), ref, rtol=2e-2, atol=5e-2)


# ---------------------------------------------------------------------------------------------- CTA pairs
def _run_pair(M, N, K, a_mn, b_mn, bn, **kw):
    lib = native.require()
    lib.sg_gemm_set_pair(1)
    try:
        return _run(M, N, K, a_mn, b_mn, bn, **kw)
    finally:
        lib.sg_gemm_set_pair(0)


@pytest.mark.parametrize("bn", [128, 256])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_pair_majors_tiles(a_mn, b_mn, bn):
    out, ref = _run_pair(512, 512, 384, a_mn, b_mn, bn)
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-2)


@pytest.mark.parametrize("M,N,K", [(1000, 2500, 784), (1000, 10, 500), (130, 70, 72), (777, 1500, 2000),
                                   (4096, 1024, 520)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (False, False), (True, True)])
def test_gemm_pair_ragged(M, N, K, a_mn, b_mn):
    out, ref = _run_pair(M, N, K, a_mn, b_mn, 0)
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=3e-2)


def test_gemm_pair_epilogue_and_timing():
    bias = torch.randn(300, device="cuda")
    out, ref = _run_pair(700, 300, 264, False, True, 0, bias=bias, act=G.ACT_STANH)
    torch.testing.assert_close(out, ref, rtol=5e-3, atol=3e-2)
    M = N = K = 8192
    a = (torch.randn(M, K, device="cuda") * 0.05).to(torch.bfloat16)
    b = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    o = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    lib = native.require()
    res = {}
    for mode in (0, 1):
        lib.sg_gemm_set_pair(mode)
        for _ in range(3):
            G.gemm(a, b, M, N, K, out16=o)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            G.gemm(a, b, M, N, K, out16=o)
        e1.record()
        torch.cuda.synchronize()
        res[mode] = e0.elapsed_time(e1) / 10
        torch.testing.assert_close(o[:256, :512].float(), a[:256].float() @ b[:512].float().t(), rtol=2e-2, atol=5e-2)
    lib.sg_gemm_set_pair(0)
    print(f"\n[gemm 8192^3] single-CTA tiles {res[0]:.3f} ms ({2*M*N*K/res[0]/1e9:.0f} TFLOP/s), "
          f"CTA pairs {res[1]:.3f} ms ({2*M*N*K/res[1]/1e9:.0f} TFLOP/s)")


# ---------------------------------------------------------------------------------------------- split-K
@pytest.mark.parametrize("split", [0, 1, 7])
@pytest.mark.parametrize("accumulate", [False, True])
def test_gemm_split_k_wgrad_shapes(split, accumulate):
    # conv-wgrad-like: tiny output, huge K (= batch x pixels), both operands MN-major, permuted columns
    F, C, kk, P = 96, 3, 121, 20000
    dev = "cuda"
    a, b = _mk(P, F, dev, 0.1), _mk(P, C * kk, dev, 0.1)
    out = torch.full((F, C * kk), 0.5, device=dev)
    G.gemm(a, b, F, C * kk, P, a_mn=True, b_mn=True, out32=out, perm=(C, kk), split_k=split, accumulate=accumulate)
    torch.cuda.synchronize()
    ref = G.gemm_ref(a, b, True, True, perm=(C, kk)) + (0.5 if accumulate else 0.0)
    torch.testing.assert_close(out, ref, rtol=3e-3, atol=5e-2)


def test_gemm_split_k_plain_and_timing():
    M, N, K = 256, 2400, 46656            # AlexNet conv2 weight gradient at batch 64
    dev = "cuda"
    a, b = _mk(K, M, dev, 0.05), _mk(K, N, dev, 0.05)
    outs, times = {}, {}
    for split in (1, 0):
        out = torch.empty(M, N, device=dev)
        for _ in range(2):
            G.gemm(a, b, M, N, K, a_mn=True, b_mn=True, out32=out, split_k=split)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            G.gemm(a, b, M, N, K, a_mn=True, b_mn=True, out32=out, split_k=split)
        e1.record()
        torch.cuda.synchronize()
        outs[split], times[split] = out, e0.elapsed_time(e1) / 5
    torch.testing.assert_close(outs[0], outs[1], rtol=2e-3, atol=5e-2)
    print(f"\n[wgrad 256x2400x46656] no split {times[1]*1e3:.0f} us, split-K {times[0]*1e3:.0f} us")


# ---------------------------------------------------------------------------------------------------------
# tf32 operands (fp32 in memory): the fp32-parity path.  The oracle is the fp32 product of the same fp32 inputs;
# the tensor core keeps 10 mantissa bits of each operand, so the tolerance is ~2^-10 relative per product term.
# ---------------------------------------------------------------------------------------------------------
def _mk32(rows, cols, dev, scale=1.0):
    buf = torch.zeros(rows, (cols + 3) // 4 * 4, device=dev, dtype=torch.float32)
    buf[:, :cols] = torch.randn(rows, cols, device=dev) * scale
    return buf[:, :cols]


def _run32(M, N, K, a_mn, b_mn, bn=0, pair=0, **kw):
    dev = "cuda"
    torch.manual_seed(M * 5 + N * 11 + K)
    a = _mk32(K, M, dev) if a_mn else _mk32(M, K, dev)
    b = _mk32(K, N, dev) if b_mn else _mk32(N, K, dev)
    out32 = torch.full((M, (N + 3) // 4 * 4), float("nan"), device=dev)[:, :N]
    G.gemm(a, b, M, N, K, a_mn=a_mn, b_mn=b_mn, out32=out32, tile=(bn, pair) if (bn or pair) else None, **kw)
    torch.cuda.synchronize()
    kwr = {k: v for k, v in kw.items() if k in ("bias", "act", "grad_act", "aux")}
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        ref = G.gemm_ref(a, b, a_mn, b_mn, **kwr)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    return out32, ref


def _tf32_close(out, ref, K):
    # the tensor core truncates each operand to 10 mantissa bits (relative error < 2^-10 each, < 2^-9 per product);
    # products have random signs, so the sum's error grows like sqrt(K); the bound is ~8 sigma of that
    tol = 2.0 ** -10 * (K ** 0.5) * 6.0 + 1e-5
    err = (out - ref).abs().max().item()
    assert err <= tol, f"max abs error {err} > {tol}"
    assert torch.isfinite(out).all()


@pytest.mark.parametrize("bn", [64, 128, 256])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_tf32_majors_tiles(a_mn, b_mn, bn):
    out, ref = _run32(256, 512, 384, a_mn, b_mn, bn, 1)
    _tf32_close(out, ref, 384)


@pytest.mark.parametrize("bn", [128, 256])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_tf32_pair(a_mn, b_mn, bn):
    out, ref = _run32(512, 768, 416, a_mn, b_mn, bn, 2)
    _tf32_close(out, ref, 416)


@pytest.mark.parametrize("M,N,K", [(1000, 2500, 784), (1000, 10, 500), (64, 500, 800), (130, 70, 72), (1, 8, 8),
                                   (777, 1500, 2000), (33, 36, 20)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (False, False), (True, True)])
def test_gemm_tf32_ragged(M, N, K, a_mn, b_mn):
    out, ref = _run32(M, N, K, a_mn, b_mn)
    _tf32_close(out, ref, K)


def test_gemm_tf32_epilogues():
    dev = "cuda"
    M, N, K = 300, 200, 136
    bias = torch.randn(N, device=dev)
    for act in (G.ACT_RELU, G.ACT_STANH):
        out, ref = _run32(M, N, K, False, True, bias=bias, act=act)
        _tf32_close(out, ref, K)
    # data-gradient epilogue: fp32 output times act'(aux), aux fp32
    aux = torch.tanh(torch.randn(M, N, device=dev)) * 1.7
    for ga in (G.ACT_RELU, G.ACT_STANH):
        out, ref = _run32(M, N, K, False, False, grad_act=ga, aux=aux)
        _tf32_close(out, ref, K)
    # accumulate and split-K
    a, b = _mk32(4096, 96, dev), _mk32(4096, 200, dev)
    out32 = torch.ones(96, 200, device=dev)
    G.gemm(a, b, 96, 200, 4096, a_mn=True, b_mn=True, out32=out32, accumulate=True)
    torch.cuda.synchronize()
    _tf32_close(out32, a.t() @ b + 1.0, 4096)
