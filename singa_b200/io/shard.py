"""Python face of the native Shard (``csrc/shard.cc``): same operations as the reference class
(include/utils/shard.h:33-142) — ``insert`` (duplicate keys ignored), ``next``, ``flush``, ``seek_to_first``,
``count`` — plus ``read_image_batch`` which decodes image records directly into pinned batch arrays."""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import numpy as np

from ..ops import native

_declared = False


def _lib():
    global _declared
    lib = native.require()
    if not _declared:
        C = ctypes
        lib.sg_shard_open.restype = C.c_void_p
        lib.sg_shard_open.argtypes = [C.c_char_p, C.c_char, C.c_longlong]
        lib.sg_shard_close.argtypes = [C.c_void_p]
        lib.sg_shard_close.restype = None
        lib.sg_shard_insert.argtypes = [C.c_void_p, C.c_char_p, C.c_longlong, C.c_char_p, C.c_longlong]
        lib.sg_shard_flush.argtypes = [C.c_void_p]
        lib.sg_shard_next.argtypes = [C.c_void_p]
        for f in (lib.sg_shard_key, lib.sg_shard_val):
            f.restype = C.c_void_p
            f.argtypes = [C.c_void_p, C.POINTER(C.c_longlong)]
        lib.sg_shard_seek_to_first.argtypes = [C.c_void_p]
        lib.sg_shard_seek_to_first.restype = None
        lib.sg_shard_count.argtypes = [C.c_void_p]
        lib.sg_shard_count.restype = C.c_longlong
        lib.sg_shard_read_image_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_longlong]
        lib.sg_shard_read_image_batch_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_longlong]
        _declared = True
    return lib


class Shard:
    kRead, kCreate, kAppend = "r", "c", "a"

    def __init__(self, folder: str, mode: str = "r", capacity: int = 104857600):
        self.lib = _lib()
        self.folder, self.mode = folder, mode
        self.h = self.lib.sg_shard_open(folder.encode(), mode.encode(), capacity)
        if not self.h:
            raise IOError(self.lib.sg_last_error().decode())

    def insert(self, key, value) -> bool:
        k = key.encode() if isinstance(key, str) else bytes(key)
        v = value if isinstance(value, (bytes, bytearray)) else value.SerializeToString()
        rc = self.lib.sg_shard_insert(self.h, k, len(k), bytes(v), len(v))
        if rc < 0:
            raise IOError(self.lib.sg_last_error().decode())
        return rc == 1

    def next(self) -> Optional[Tuple[bytes, bytes]]:
        rc = self.lib.sg_shard_next(self.h)
        if rc < 0:
            raise IOError(self.lib.sg_last_error().decode())
        if rc == 0:
            return None
        n = ctypes.c_longlong()
        kp = self.lib.sg_shard_key(self.h, ctypes.byref(n))
        key = ctypes.string_at(kp, n.value)
        vp = self.lib.sg_shard_val(self.h, ctypes.byref(n))
        return key, ctypes.string_at(vp, n.value)

    def flush(self) -> None:
        self.lib.sg_shard_flush(self.h)

    def seek_to_first(self) -> None:
        self.lib.sg_shard_seek_to_first(self.h)

    def count(self) -> int:
        return int(self.lib.sg_shard_count(self.h))

    def close(self) -> None:
        if self.h:
            self.lib.sg_shard_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def read_image_batch(shard: Shard, pixels: np.ndarray, labels: np.ndarray) -> None:
    """Decode the next ``batch`` image records into ``pixels`` — uint8 for byte-valued records, float32 for
    float-valued ones (mean-subtracted images; a uint8 batch refuses them rather than clamping)."""
    assert pixels.dtype in (np.uint8, np.float32) and pixels.flags.c_contiguous and labels.dtype == np.int32
    batch, dim = pixels.shape
    fn = shard.lib.sg_shard_read_image_batch if pixels.dtype == np.uint8 else shard.lib.sg_shard_read_image_batch_f32
    rc = fn(shard.h, pixels.ctypes.data, labels.ctypes.data, batch, dim)
    if rc < 0:
        raise IOError(shard.lib.sg_last_error().decode())
