"""Execution context: which device the worker runs on and which backend executes the layer math.

* ``native`` — hand-written sm_100a kernels from ``libsinga_b200.so``.  The product path; selected automatically
  when a GPU and the library exist.  Two compute precisions (``Context.dtype``):
    - ``fp32`` (default; the reference computes in fp32, src/worker/layer.cc:186-213): fp32 activations, gradients
      and weights; the tcgen05 GEMMs multiply them as tf32 with fp32 accumulation; no low-precision weight copy.
    - ``bf16``: bf16 activations and a bf16 weight mirror, fp32 accumulation and fp32 master parameters.
* ``torch``  — plain PyTorch ops in fp32.  On CPU this is the oracle / plumbing path (BASELINE config 1);
  on a GPU it is the cuBLAS/cuDNN *library baseline* the product is compared against.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch

from .ops import native


DTYPES = ("fp32", "bf16")


def _default_dtype() -> str:
    d = os.environ.get("SINGA_B200_DTYPE", "fp32")
    if d not in DTYPES:
        raise ValueError(f"SINGA_B200_DTYPE must be one of {DTYPES}, got {d!r}")
    return d


@dataclass
class Context:
    device: torch.device
    backend: str  # "native" | "torch"
    dtype: str = "fp32"  # compute precision of the native backend: "fp32" (tf32 tensor cores) | "bf16"

    @property
    def is_native(self) -> bool:
        return self.backend == "native"

    @property
    def is_bf16(self) -> bool:
        return self.is_native and self.dtype == "bf16"

    @property
    def act_dtype(self) -> torch.dtype:
        return torch.bfloat16 if self.is_bf16 else torch.float32

    def pad_ld(self, n: int) -> int:
        """Row stride (elements) that is TMA-legal for an operand of the compute dtype: 16-byte multiple."""
        return (n + 7) // 8 * 8 if self.is_bf16 else (n + 3) // 4 * 4


def _default() -> Context:
    want = os.environ.get("SINGA_B200_BACKEND", "")
    if torch.cuda.is_available():
        dev = torch.device("cuda", torch.cuda.current_device())
        if want == "torch":
            return Context(dev, "torch", _default_dtype())
        native.require()  # on a GPU box a missing library is an error, never a silent fallback
        return Context(dev, "native", _default_dtype())
    return Context(torch.device("cpu"), "torch", _default_dtype())


_ctx: Context | None = None


def ctx() -> Context:
    global _ctx
    if _ctx is None:
        _ctx = _default()
    return _ctx


def set_context(device: str | torch.device | None = None, backend: str | None = None,
                dtype: str | None = None) -> Context:
    """Override the context (tests, ``--impl`` / ``--dtype`` switches).  ``backend='native'`` requires CUDA + the
    library."""
    global _ctx
    cur = _default() if _ctx is None else _ctx
    dev = torch.device(device) if device is not None else cur.device
    be = backend or cur.backend
    if dev.type == "cuda" and dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    if be == "native":
        if dev.type != "cuda":
            raise RuntimeError("the native backend needs a CUDA device")
        native.require()
    dt = dtype or cur.dtype
    if dt not in DTYPES:
        raise ValueError(f"dtype must be one of {DTYPES}, got {dt!r}")
    _ctx = Context(dev, be, dt)
    return _ctx


def reset_context() -> None:
    global _ctx
    _ctx = None


def pad8(n: int) -> int:
    return (n + 7) // 8 * 8


_step_state = None
_RING = 512      # pinned staging slots: the host may run hundreds of (graph-replayed) steps ahead of the GPU


def step_state() -> dict:
    """Per-step scalars kept in DEVICE memory so that a captured CUDA graph follows them without
    re-capture: ``lr`` (fp32[1], read by the fused updater), ``step`` (int64[1], mixed into the dropout seeds of the
    forward pass) and ``step_bwd`` (int64[1], the same for the backward pass — it differs from ``step`` only when the
    executor runs the rotated schedule, where one graph holds backward(t) and forward(t+1)).  All live in one 32-byte
    device buffer that the executor refreshes with ONE tiny async copy per step from a ring of pinned host slots (a
    slot is not rewritten until ``_RING`` steps later, so a pending copy never reads a newer step's values)."""
    global _step_state
    c = ctx()
    if _step_state is None or _step_state["lr"].device != c.device:
        pin = c.device.type == "cuda"
        dev = torch.zeros(32, dtype=torch.uint8, device=c.device)
        host = torch.zeros(_RING, 32, dtype=torch.uint8, pin_memory=pin)
        _step_state = {
            "buf": dev,
            "step": dev.view(torch.int64)[0:1],
            "lr": dev.view(torch.float32)[2:3],
            "step_bwd": dev.view(torch.int64)[2:3],
            "host": host,
            "host_step_view": host.view(torch.int64).view(_RING, 4),
            "host_lr_view": host.view(torch.float32).view(_RING, 8),
            "host_step": torch.zeros(1, dtype=torch.int64),      # last values published (host-side readers)
            "host_step_bwd": torch.zeros(1, dtype=torch.int64),
            "host_lr": torch.zeros(1, dtype=torch.float32),
            "slot": 0,
        }
    return _step_state


def set_step_state(step: int, lr: float, step_bwd: int | None = None) -> None:
    st = step_state()
    k = st["slot"]
    st["slot"] = (k + 1) % _RING
    sb = step if step_bwd is None else step_bwd
    st["host_step_view"][k, 0] = step
    st["host_step_view"][k, 2] = sb
    st["host_lr_view"][k, 2] = lr
    st["host_step"][0] = step
    st["host_step_bwd"][0] = sb
    st["host_lr"][0] = lr
    st["buf"].copy_(st["host"][k], non_blocking=True)
