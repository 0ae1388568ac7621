"""tcgen05 GEMM wrapper (``sg_gemm``: bf16 or tf32 operands) and its plain-PyTorch restatement used on CPU and as
the oracle.

``gemm(a, b, ...)`` computes ``acc[m][n] = sum_k A(m,k) B(n,k)`` where ``a`` is stored ``[M][K]`` (or
``[K][M]`` when ``a_mn``) and ``b`` is stored ``[N][K]`` (or ``[K][N]`` when ``b_mn``), followed by the fused
epilogue (bias, activation, activation-gradient, accumulate, optional conv-wgrad column permutation).
Reference behaviour: ``mshadow::dot`` as used by InnerProduct (src/worker/layer.cc:186-213).
"""
from __future__ import annotations

import ctypes

import torch

from . import native

ACT_NONE, ACT_RELU, ACT_STANH = 0, 1, 2
DT_BF16, DT_TF32 = 0, 1
STANH_A, STANH_B = 1.7159047, 0.66666667


def act_fwd_ref(x: torch.Tensor, act: int) -> torch.Tensor:
    if act == ACT_RELU:
        return torch.clamp_min(x, 0)
    if act == ACT_STANH:
        return STANH_A * torch.tanh(STANH_B * x)
    return x


def act_grad_from_output_ref(y: torch.Tensor, act: int) -> torch.Tensor:
    if act == ACT_RELU:
        return (y > 0).to(y.dtype)
    if act == ACT_STANH:
        return STANH_A * STANH_B - (STANH_B / STANH_A) * y * y
    return torch.ones_like(y)


def gemm_ref(a, b, a_mn=False, b_mn=False, bias=None, act=ACT_NONE, grad_act=ACT_NONE, aux=None,
             accumulate_into=None, perm=None):
    """fp32 PyTorch restatement of the kernel (inputs are upcast; result is fp32)."""
    A = a.float().t() if a_mn else a.float()
    B = b.float() if b_mn else b.float().t()
    v = A @ B
    if bias is not None:
        v = v + bias.float()
    v = act_fwd_ref(v, act)
    if grad_act != ACT_NONE:
        v = v * act_grad_from_output_ref(aux.float(), grad_act)
    if perm is not None:
        c, kk = perm
        n = v.shape[1]
        idx = torch.arange(n, device=v.device)
        out = torch.empty_like(v)
        out[:, (idx % c) * kk + idx // c] = v
        v = out
    if accumulate_into is not None:
        v = v + accumulate_into.float()
    return v


# ---------------------------------------------------------------------------------------------------------
# Shape autotuner ("measure, don't guess"): the first time a (M, N, K, majors) problem is seen outside of graph
# capture, every tile configuration is timed with CUDA events on the caller's own operands and the fastest is
# remembered for the rest of the process.  The layer shapes of a net are static, so this costs a few
# milliseconds during the eager warm-up steps and nothing afterwards.
# ---------------------------------------------------------------------------------------------------------
import os as _os

# Opt-in: tuned in isolation the small MNIST-MLP GEMMs pick narrow tiles that fill every SM, which then starve the
# weight-gradient GEMMs running concurrently on the side streams (measured 0.366 ms/step tuned vs 0.307 ms with the
# built-in wide-tile model), so the model is the default.
_TUNE = _os.environ.get("SINGA_B200_GEMM_AUTOTUNE", "0") == "1"
_tuned: dict = {}
_CANDIDATES = [(64, 1), (128, 1), (256, 1), (128, 2), (256, 2)]


def tuned_configs() -> dict:
    return dict(_tuned)


def _autotune(key, launch) -> tuple:
    best, best_t = (0, 0), float("inf")
    for bn, pair in _CANDIDATES:
        if bn > 64 and key[1] <= bn // 2:
            continue
        try:
            launch(bn, pair)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(4):
                launch(bn, pair)
            e1.record()
            torch.cuda.synchronize()
            t = e0.elapsed_time(e1)
        except RuntimeError:
            continue
        if t < best_t:
            best, best_t = (bn, pair), t
    _tuned[key] = best
    return best


def gemm(a: torch.Tensor, b: torch.Tensor, M: int, N: int, K: int, *, a_mn=False, b_mn=False,
         out32: torch.Tensor | None = None, out16: torch.Tensor | None = None,
         bias: torch.Tensor | None = None, act=ACT_NONE, grad_act=ACT_NONE,
         aux: torch.Tensor | None = None, accumulate=False, perm: tuple[int, int] | None = None,
         wait: tuple | None = None, split_k: int = 0, tile: tuple | None = None):
    """Launch the sm_100a kernel.  ``a``/``b`` are 2-D CUDA tensors of one dtype whose last dim is contiguous:
    bf16 (row stride a multiple of 8) or fp32 (multiple of 4; multiplied as tf32 — the fp32-parity path).
    Outputs are written in place."""
    lib = native.require()
    assert a.dtype == b.dtype and a.dtype in (torch.bfloat16, torch.float32), (a.dtype, b.dtype)
    dtype = DT_BF16 if a.dtype == torch.bfloat16 else DT_TF32
    assert a.stride(-1) == 1 and b.stride(-1) == 1
    ep = native.GemmEpilogue()
    if out32 is not None:
        assert out32.dtype == torch.float32 and out32.stride(-1) == 1
        ep.d32, ep.ldd32 = out32.data_ptr(), out32.stride(0)
    if out16 is not None:
        assert out16.dtype == torch.bfloat16 and out16.stride(-1) == 1
        ep.d16, ep.ldd16 = out16.data_ptr(), out16.stride(0)
    if bias is not None:
        assert bias.dtype in (torch.float32, torch.bfloat16) and bias.is_contiguous()
        ep.bias = bias.data_ptr()
        ep.bias_bf16 = 1 if bias.dtype == torch.bfloat16 else 0
    ep.act, ep.grad_act = act, grad_act
    if grad_act != ACT_NONE:
        assert aux is not None and aux.stride(-1) == 1
        ep.aux, ep.ldaux = aux.data_ptr(), aux.stride(0)
        ep.aux_bf16 = 1 if aux.dtype == torch.bfloat16 else 0
    ep.accumulate = 1 if accumulate else 0
    if perm is not None:
        ep.perm_c, ep.perm_kk = perm
    ep.split_k = split_k
    if wait is not None:   # (device int64 tensor of flag addresses, address of the sequence number)
        flags, seq = wait
        assert flags.dtype == torch.int64 and flags.is_cuda
        ep.wait_flags, ep.nwait, ep.wait_seq = flags.data_ptr(), flags.numel(), seq
    def launch(bn=0, pair=0):
        ep.tile_bn, ep.tile_pair = bn, pair
        rc = lib.sg_gemm(dtype, a.data_ptr(), 1 if a_mn else 0, a.stride(0), b.data_ptr(), 1 if b_mn else 0,
                         b.stride(0), M, N, K, ctypes.byref(ep), native.stream_ptr())
        native.check(rc, "sg_gemm")

    cfg = tile or (0, 0)    # explicit (tile_bn, tile_pair) override, e.g. from a benchmark sweep
    if _TUNE and tile is None:
        key = (M, N, K, bool(a_mn), bool(b_mn), dtype)
        cfg = _tuned.get(key)
        if cfg is None:
            # tuning re-runs the kernel: only when that is harmless (no accumulation, no flag wait, not capturing)
            if accumulate or wait is not None or torch.cuda.is_current_stream_capturing():
                cfg = (0, 0)
            else:
                cfg = _autotune(key, launch)
    launch(*cfg)
    native.count_launch()
