"""LMDB input (``kLMDBData``, src/worker/layer.cc:237-328) without liblmdb.

* ``LMDBEnv`` — Python face of the native read-only reader (``csrc/lmdb_reader.cc``): B+tree walk over an mmap of
  ``<dir>/data.mdb`` in key order, the operations the reference uses (``mdb_env_open``, cursor FIRST / NEXT, stat
  entries).
* ``LMDBSource`` — batch source of the data layer: Caffe ``Datum`` values are decoded natively straight into the
  pinned batch buffers, wrapping around at the end of the database.
* ``write_lmdb`` — minimal writer of a valid single-transaction LMDB file (sorted keys, leaf / branch / overflow
  pages, two meta pages).  Used by the tests and by ``python -m singa_b200.tools.loader --format lmdb``; the
  reference relies on Caffe's ``convert_imageset`` for this.
"""
from __future__ import annotations

import ctypes
import os
import struct
from typing import Iterable, List, Optional, Tuple

import numpy as np

from ..layers.data import BatchSource
from ..ops import native
from ..proto import schema as S

_declared = False


def _lib():
    global _declared
    lib = native.require()
    if not _declared:
        C = ctypes
        lib.sg_lmdb_open.restype = C.c_void_p
        lib.sg_lmdb_open.argtypes = [C.c_char_p]
        lib.sg_lmdb_close.argtypes = [C.c_void_p]
        lib.sg_lmdb_close.restype = None
        lib.sg_lmdb_count.argtypes = [C.c_void_p]
        lib.sg_lmdb_count.restype = C.c_longlong
        lib.sg_lmdb_seek_to_first.argtypes = [C.c_void_p]
        lib.sg_lmdb_seek_to_first.restype = None
        lib.sg_lmdb_next.argtypes = [C.c_void_p]
        for f in (lib.sg_lmdb_key, lib.sg_lmdb_val):
            f.restype = C.c_void_p
            f.argtypes = [C.c_void_p, C.POINTER(C.c_longlong)]
        lib.sg_lmdb_first_shape.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        lib.sg_lmdb_read_datum_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_longlong]
        lib.sg_lmdb_read_datum_batch_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_longlong]
        lib.sg_lmdb_first_is_float.argtypes = [C.c_void_p]
        _declared = True
    return lib


class LMDBEnv:
    """Read-only environment + cursor."""

    def __init__(self, path: str):
        self.lib = _lib()
        self.h = self.lib.sg_lmdb_open(path.encode())
        if not self.h:
            raise RuntimeError(self.lib.sg_last_error().decode(errors="replace") or f"cannot open LMDB {path}")

    def close(self) -> None:
        if self.h:
            self.lib.sg_lmdb_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def count(self) -> int:
        return int(self.lib.sg_lmdb_count(self.h))

    def seek_to_first(self) -> None:
        self.lib.sg_lmdb_seek_to_first(self.h)

    def next(self) -> Optional[Tuple[bytes, bytes]]:
        """Next (key, value) in key order; None at the end (the following call starts over)."""
        rc = self.lib.sg_lmdb_next(self.h)
        if rc < 0:
            raise RuntimeError(self.lib.sg_last_error().decode(errors="replace"))
        if rc == 0:
            return None
        n = ctypes.c_longlong()
        kp = self.lib.sg_lmdb_key(self.h, ctypes.byref(n))
        key = ctypes.string_at(kp, n.value)
        vp = self.lib.sg_lmdb_val(self.h, ctypes.byref(n))
        return key, ctypes.string_at(vp, n.value)

    def first_shape(self) -> List[int]:
        """{channels, height, width} declared by the first Datum; ``self.encoded`` tells whether its data field holds
        a compressed image (``Datum.encoded``) instead of raw pixels."""
        chw = (ctypes.c_int * 3)()
        rc = self.lib.sg_lmdb_first_shape(self.h, chw)
        self.encoded = rc == 2
        if rc not in (0, 2):
            native.check(rc, "sg_lmdb_first_shape")
        return [int(chw[0]), int(chw[1]), int(chw[2])]

    def read_datum_batch(self, pixels: np.ndarray, labels: np.ndarray) -> None:
        assert pixels.dtype in (np.uint8, np.float32) and pixels.flags.c_contiguous and labels.dtype == np.int32
        fn = self.lib.sg_lmdb_read_datum_batch if pixels.dtype == np.uint8 else self.lib.sg_lmdb_read_datum_batch_f32
        native.check(fn(self.h, pixels.ctypes.data, labels.ctypes.data, pixels.shape[0], pixels.shape[1]),
                     "sg_lmdb_read_datum_batch")

    def first_is_float(self) -> bool:
        rc = self.lib.sg_lmdb_first_is_float(self.h)
        if rc < 0:
            raise IOError(self.lib.sg_last_error().decode())
        return rc == 1


def decode_image(buf: bytes) -> np.ndarray:
    """Compressed image bytes -> {3, H, W} uint8 in Caffe's channel order (BGR), via OpenCV or Pillow."""
    try:
        import cv2
        img = cv2.imdecode(np.frombuffer(buf, dtype=np.uint8), cv2.IMREAD_COLOR)
        if img is None:
            raise ValueError("cv2.imdecode failed")
    except ImportError:
        import io
        from PIL import Image
        img = np.asarray(Image.open(io.BytesIO(buf)).convert("RGB"))[:, :, ::-1]
    return np.ascontiguousarray(img.transpose(2, 0, 1))


class LMDBSource(BatchSource):
    def __init__(self, path: str):
        if not os.path.exists(path):
            raise RuntimeError(
                f"kLMDBData layer: {path!r} does not exist.  Set `data_param {{ source: \"synthetic\" "
                "synthetic_shape: ... }` or export SINGA_B200_SYNTHETIC=1 to substitute synthetic records.")
        self.env = LMDBEnv(path)
        c, h, w = self.env.first_shape()
        self.encoded = self.env.encoded
        if self.encoded:                       # compressed images: the shape is that of the first decoded image
            self.env.seek_to_first()
            c, h, w = self._decode_record(self.env.next()[1])[0].shape
        self.sample_shape = [c, h, w] if c > 1 else [h, w]
        # Datum.float_data records (mean-subtracted images) are staged as float32, unchanged
        self.is_float = (not self.encoded) and self.env.first_is_float()
        self.env.seek_to_first()

    @staticmethod
    def _decode_record(value: bytes):
        d = S.Datum()
        d.ParseFromString(value)
        return decode_image(d.data), int(d.label)

    def next_batch(self, pixels, labels):
        if not self.encoded:
            self.env.read_datum_batch(pixels, labels)     # native: raw pixels memcpy'd into the batch buffers
            return
        shape = tuple(self.sample_shape) if len(self.sample_shape) == 3 else (1, *self.sample_shape)
        for i in range(pixels.shape[0]):                  # encoded Datums: per-record image decode on the host
            kv = self.env.next() or self.env.next()
            img, labels[i] = self._decode_record(kv[1])
            if img.shape != shape:
                import cv2
                img = cv2.resize(img.transpose(1, 2, 0), (shape[2], shape[1])).transpose(2, 0, 1)
            pixels[i, :] = img.reshape(-1)

    def skip(self, n):
        for _ in range(n):
            if self.env.next() is None:
                self.env.next()

    def count(self):
        return self.env.count()


# ----------------------------------------------------------------------------------------------------------
# writer
# ----------------------------------------------------------------------------------------------------------
_PSIZE = 4096
_PAGEHDR = 16
_P_BRANCH, _P_LEAF, _P_OVERFLOW, _P_META = 0x01, 0x02, 0x04, 0x08
_F_BIGDATA = 0x01
_NODEMAX = ((((_PSIZE - _PAGEHDR) // 2) & ~1) - 2)      # larger leaf nodes move their data to overflow pages
_P_INVALID = (1 << 64) - 1


def _even(n: int) -> int:
    return (n + 1) & ~1


def _page(pgno: int, flags: int, nodes: List[bytes]) -> bytes:
    """A branch / leaf page: node offsets grow up from the header, node bodies grow down from the page end."""
    buf = bytearray(_PSIZE)
    upper = _PSIZE
    ptrs = []
    for nd in nodes:
        upper -= _even(len(nd))
        buf[upper:upper + len(nd)] = nd
        ptrs.append(upper)
    lower = _PAGEHDR + 2 * len(nodes)
    assert lower <= upper, "page overflow"
    struct.pack_into("<QHHHH", buf, 0, pgno, 0, flags, lower, upper)
    struct.pack_into(f"<{len(ptrs)}H", buf, _PAGEHDR, *ptrs)
    return bytes(buf)


def _pack_level(entries: List[Tuple[bytes, bytes]], flags: int, pages: List[bytes], first_pgno: int):
    """Greedy packing of pre-built nodes (first key, node bytes) into pages; returns [(first key, pgno)]."""
    out = []
    cur: List[bytes] = []
    cur_key = None
    used = _PAGEHDR
    for key, nd in entries:
        need = _even(len(nd)) + 2
        if cur and used + need > _PSIZE:
            pgno = first_pgno + len(pages)
            pages.append(_page(pgno, flags, cur))
            out.append((cur_key, pgno))
            cur, used, cur_key = [], _PAGEHDR, None
        if flags == _P_BRANCH and not cur:
            nd = nd[:6] + struct.pack("<H", 0)        # the first key of a branch page is implicit (empty)
            need = _even(len(nd)) + 2
        if cur_key is None:
            cur_key = key
        cur.append(nd)
        used += need
    if cur:
        pgno = first_pgno + len(pages)
        pages.append(_page(pgno, flags, cur))
        out.append((cur_key, pgno))
    return out


def write_lmdb(path: str, items: Iterable[Tuple[bytes, bytes]]) -> int:
    """Write ``items`` ((key, value) pairs, any order, unique keys) as ``<path>/data.mdb`` (+ an empty ``lock.mdb``).
    Returns the number of records."""
    items = sorted(items, key=lambda kv: kv[0])
    os.makedirs(path, exist_ok=True)
    pages: List[bytes] = []                      # pages 2.. in file order
    first = 2
    leaf_nodes = []
    n_overflow = 0
    for key, val in items:
        assert 0 < len(key) <= 511, "LMDB keys are 1..511 bytes"
        if 8 + len(key) + len(val) > _NODEMAX:
            npg = (_PAGEHDR - 1 + len(val)) // _PSIZE + 1
            pgno = first + len(pages)
            body = bytearray(npg * _PSIZE)
            struct.pack_into("<QHHI", body, 0, pgno, 0, _P_OVERFLOW, npg)
            body[_PAGEHDR:_PAGEHDR + len(val)] = val
            for i in range(npg):
                pages.append(bytes(body[i * _PSIZE:(i + 1) * _PSIZE]))
            n_overflow += npg
            nd = struct.pack("<HHHH", len(val) & 0xFFFF, len(val) >> 16, _F_BIGDATA, len(key)) + key + struct.pack("<Q", pgno)
        else:
            nd = struct.pack("<HHHH", len(val) & 0xFFFF, len(val) >> 16, 0, len(key)) + key + val
        leaf_nodes.append((key, nd))
    depth, root, n_leaf, n_branch = 0, _P_INVALID, 0, 0
    if leaf_nodes:
        level = _pack_level(leaf_nodes, _P_LEAF, pages, first)
        n_leaf = len(level)
        depth = 1
        while len(level) > 1:
            nodes = [(k, struct.pack("<HHHH", pg & 0xFFFF, (pg >> 16) & 0xFFFF, (pg >> 32) & 0xFFFF, len(k)) + k)
                     for k, pg in level]
            level = _pack_level(nodes, _P_BRANCH, pages, first)
            n_branch += len(level)
            depth += 1
        root = level[0][1]
    last_pg = first + len(pages) - 1 if pages else 1
    mapsize = max(1 << 20, (last_pg + 1) * _PSIZE)

    def meta(pgno: int, txnid: int) -> bytes:
        buf = bytearray(_PSIZE)
        struct.pack_into("<QHHI", buf, 0, pgno, 0, _P_META, 0)
        off = _PAGEHDR
        struct.pack_into("<IIQQ", buf, off, 0xBEEFC0DE, 1, 0, mapsize)
        off += 24
        struct.pack_into("<IHHQQQQQ", buf, off, _PSIZE, 0x08, 0, 0, 0, 0, 0, _P_INVALID)            # FREE db
        off += 48
        struct.pack_into("<IHHQQQQQ", buf, off, 0, 0, depth, n_branch, n_leaf, n_overflow, len(items), root)  # MAIN db
        off += 48
        struct.pack_into("<QQ", buf, off, last_pg, txnid)
        return bytes(buf)

    tmp = os.path.join(path, "data.mdb.tmp")
    with open(tmp, "wb") as f:
        f.write(meta(0, 0 if not items else 2))
        f.write(meta(1, 1))
        for p in pages:
            f.write(p)
    os.replace(tmp, os.path.join(path, "data.mdb"))
    open(os.path.join(path, "lock.mdb"), "ab").close()
    return len(items)


def datum_bytes(pixels: np.ndarray, label: int) -> bytes:
    """Serialised Caffe Datum for a {C,H,W} (or {H,W}) image: uint8 pixels go to ``data``, floating-point pixels
    (e.g. a mean-subtracted image) to ``float_data``."""
    d = S.Datum()
    shp = list(pixels.shape)
    if len(shp) == 2:
        shp = [1] + shp
    d.channels, d.height, d.width = shp
    if np.issubdtype(np.asarray(pixels).dtype, np.floating):
        d.float_data.extend(np.asarray(pixels, dtype=np.float32).reshape(-1).tolist())
    else:
        d.data = np.ascontiguousarray(pixels, dtype=np.uint8).tobytes()
    d.label = int(label)
    return d.SerializeToString()
