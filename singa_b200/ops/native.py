"""ctypes binding of ``libsinga_b200.so`` (C ABI declared in ``include/singa_b200.h``).

The library is built in-tree by :mod:`singa_b200.build`; if the shared object is missing or stale and
``nvcc`` is present it is (re)built on first use.  On a machine with a GPU a missing library is an error
(``require()``), never a silent fallback: the CUDA path must be the one that runs there.
"""
from __future__ import annotations

import ctypes
import os
import threading

import torch

from .. import build as _build

_lock = threading.Lock()
_lib = None
_load_error: str | None = None


class GemmEpilogue(ctypes.Structure):
    _fields_ = [
        ("d32", ctypes.c_void_p), ("ldd32", ctypes.c_longlong),
        ("d16", ctypes.c_void_p), ("ldd16", ctypes.c_longlong),
        ("bias", ctypes.c_void_p),
        ("act", ctypes.c_int), ("grad_act", ctypes.c_int),
        ("aux", ctypes.c_void_p), ("ldaux", ctypes.c_longlong), ("aux_bf16", ctypes.c_int),
        ("accumulate", ctypes.c_int),
        ("perm_c", ctypes.c_int), ("perm_kk", ctypes.c_int),
        ("wait_flags", ctypes.c_void_p), ("nwait", ctypes.c_int), ("wait_seq", ctypes.c_void_p),
        ("bias_bf16", ctypes.c_int),
        ("tile_bn", ctypes.c_int), ("tile_pair", ctypes.c_int), ("split_k", ctypes.c_int),
    ]


def _declare(lib):
    lib.sg_last_error.restype = ctypes.c_char_p
    lib.sg_version.restype = ctypes.c_int
    lib.sg_device_cc.argtypes = [ctypes.c_int]
    lib.sg_gemm_bf16.argtypes = [
        ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong,
        ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong,
        ctypes.c_int, ctypes.c_int, ctypes.c_int,
        ctypes.POINTER(GemmEpilogue), ctypes.c_void_p,
    ]
    lib.sg_gemm.argtypes = [ctypes.c_int] + list(lib.sg_gemm_bf16.argtypes)
    from ._decl import ConvGeom
    lib.sg_conv_chunk_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.sg_conv_rows.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ConvGeom), ctypes.c_void_p,
                                 ctypes.c_longlong, ctypes.c_int, ctypes.POINTER(GemmEpilogue), ctypes.c_void_p]
    lib.sg_conv_wgrad.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ConvGeom), ctypes.c_void_p,
                                  ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int,
                                  ctypes.c_void_p]
    lib.sg_gemm_set_tile.argtypes = [ctypes.c_int]
    lib.sg_gemm_set_pair.argtypes = [ctypes.c_int]


def load():
    """Return the loaded library or ``None`` (reason in :func:`load_error`)."""
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return _lib
    with _lock:
        if _lib is not None or _load_error is not None:
            return _lib
        try:
            if _build.needs_build() and _build.nvcc_path() and os.environ.get("SINGA_B200_NO_AUTOBUILD") != "1":
                _build.build()
            lib = ctypes.CDLL(str(_build.LIB_PATH))
            _declare(lib)
            from . import _decl
            _decl.declare_all(lib)
            _lib = lib
        except Exception as e:  # noqa: BLE001 - recorded and surfaced by require()
            _load_error = f"{type(e).__name__}: {e}"
    return _lib


def load_error() -> str | None:
    return _load_error


def require():
    lib = load()
    if lib is None:
        raise RuntimeError(f"libsinga_b200.so is not available ({_load_error}); run `python -m singa_b200.build`")
    return lib


def available() -> bool:
    """True when the native kernels can run: library loaded and a CUDA device present."""
    return torch.cuda.is_available() and load() is not None


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = require().sg_last_error().decode(errors="replace")
        raise RuntimeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


# count of native kernel launches issued through this module (bench.py reports it as gpu_launches)
launch_count = 0


def count_launch(n: int = 1) -> None:
    global launch_count
    launch_count += n
