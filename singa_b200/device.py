"""Execution context: which device the worker runs on and which backend executes the layer math.

* ``native`` — hand-written sm_100a kernels from ``libsinga_b200.so`` (bf16 activations, fp32 accumulation,
  fp32 master parameters).  The product path; selected automatically when a GPU and the library exist.
* ``torch``  — plain PyTorch ops in fp32.  On CPU this is the oracle / plumbing path (BASELINE config 1);
  on a GPU it is the cuBLAS/cuDNN *library baseline* the product is compared against.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch

from .ops import native


@dataclass
class Context:
    device: torch.device
    backend: str  # "native" | "torch"

    @property
    def is_native(self) -> bool:
        return self.backend == "native"

    @property
    def act_dtype(self) -> torch.dtype:
        return torch.bfloat16 if self.is_native else torch.float32


def _default() -> Context:
    want = os.environ.get("SINGA_B200_BACKEND", "")
    if torch.cuda.is_available():
        dev = torch.device("cuda", torch.cuda.current_device())
        if want == "torch":
            return Context(dev, "torch")
        native.require()  # on a GPU box a missing library is an error, never a silent fallback
        return Context(dev, "native")
    return Context(torch.device("cpu"), "torch")


_ctx: Context | None = None


def ctx() -> Context:
    global _ctx
    if _ctx is None:
        _ctx = _default()
    return _ctx


def set_context(device: str | torch.device | None = None, backend: str | None = None) -> Context:
    """Override the context (tests, ``--impl`` switches).  ``backend='native'`` requires CUDA + the library."""
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
    _ctx = Context(dev, be)
    return _ctx


def reset_context() -> None:
    global _ctx
    _ctx = None


def pad8(n: int) -> int:
    return (n + 7) // 8 * 8


_step_state = None


def step_state() -> dict:
    """Per-step scalars kept in DEVICE memory so that a captured CUDA graph follows them without
    re-capture: ``lr`` (fp32[1], read by the fused updater) and ``step`` (int64[1], mixed into dropout seeds).
    The executor refreshes them with one tiny async copy before each step."""
    global _step_state
    c = ctx()
    if _step_state is None or _step_state["lr"].device != c.device:
        pin = c.device.type == "cuda"
        _step_state = {
            "lr": torch.zeros(1, dtype=torch.float32, device=c.device),
            "step": torch.zeros(1, dtype=torch.int64, device=c.device),
            "host_lr": torch.zeros(1, dtype=torch.float32, pin_memory=pin),
            "host_step": torch.zeros(1, dtype=torch.int64, pin_memory=pin),
        }
    return _step_state


def set_step_state(step: int, lr: float) -> None:
    st = step_state()
    st["host_lr"][0] = lr
    st["host_step"][0] = step
    st["lr"].copy_(st["host_lr"], non_blocking=True)
    st["step"].copy_(st["host_step"], non_blocking=True)
