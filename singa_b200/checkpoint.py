"""Checkpoint / resume in the Param-proto layout, one file per parameter-server shard.

The reference declares checkpointing but never implements it (``Worker::Resume`` is an empty TODO,
src/worker/worker.cc:65-67; ``ParamProto`` has its value field commented out, src/proto/model.proto:69-70, with a
warning about protobuf's message-size limit; ``kPretrained`` has no case in ``Param::Init``).  This module supplies the
intended design (the table-server predecessor dumped per shard too, src/test/test_table_server.cc:148-181):

  file   = binary ``ParamCheckpoint { step, model_name, shard_id, nshards, repeated ParamProto param }``
  param  = name / id / shape + ``blob[]`` where each ``BlobProto`` carries one segment of at most
           ``split_threshold`` floats starting at element ``offset`` of the flattened parameter: ``data`` = value,
           ``diff`` = updater history; ``aux_blob[]`` likewise holds ``data`` = AdaDelta update, ``diff`` = RandomSync
           snapshot.

Who writes what: every rank writes the element ranges whose fp32 master copy (and optimizer state) it OWNS —
``<folder>/step<N>.shard<r>of<W>.bin`` — straight from its own memory; nothing is gathered.  A job without sharded
servers writes one file, ``step<N>.bin``.  Resuming reads every shard file of the step and scatters the segments by
``offset``, so the number of ranks may change between save and resume.

The files are standard protobuf wire format, but the float payloads are neither built nor parsed through Python
protobuf objects: segments are emitted as packed fields directly from NumPy buffers and read back with
``numpy.frombuffer`` (a 1.09 G-parameter model would otherwise go through per-element Python lists and hit the 2 GB
message limit).  Device->host copies run on a side stream into pinned memory.  Files appear atomically (written to
``.tmp``, then renamed).
"""
from __future__ import annotations

import glob
import os
import re
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np
import torch

from .proto import schema as S

# ---------------------------------------------------------------------------------------------------------
# protobuf wire helpers (varint, length-delimited fields)
# ---------------------------------------------------------------------------------------------------------
_BLOB_NUM, _BLOB_DATA, _BLOB_DIFF, _BLOB_OFFSET = 1, 5, 6, 7
_PARAM_BLOB, _PARAM_AUX = 6, 15
_CK_STEP, _CK_PARAM, _CK_SHARD, _CK_NSHARDS, _CK_NAME = 1, 2, 3, 4, 5


def _varint(v: int) -> bytes:
    out = bytearray()
    v &= (1 << 64) - 1
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _tag(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def _read_varint(buf, pos: int) -> Tuple[int, int]:
    v, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, pos
        shift += 7


def _fields(buf, lo: int, hi: int) -> Iterator[Tuple[int, int, int, int]]:
    """Yield (field, wire, start, end) of every field in buf[lo:hi]; for varints ``start`` is the VALUE."""
    pos = lo
    while pos < hi:
        key, pos = _read_varint(buf, pos)
        field, wire = key >> 3, key & 7
        if wire == 0:
            v, pos = _read_varint(buf, pos)
            yield field, wire, v, pos
        elif wire == 2:
            n, pos = _read_varint(buf, pos)
            yield field, wire, pos, pos + n
            pos += n
        elif wire == 5:
            yield field, wire, pos, pos + 4
            pos += 4
        elif wire == 1:
            yield field, wire, pos, pos + 8
            pos += 8
        else:
            raise ValueError(f"unsupported wire type {wire} in checkpoint")


def _blob_bytes(offset: int, a: Optional[np.ndarray], b: Optional[np.ndarray]) -> List[bytes]:
    """Serialised BlobProto { num, offset, data (packed), diff (packed) } as a list of buffers (no float copies)."""
    n = a.size if a is not None else b.size
    parts = [_tag(_BLOB_NUM, 0), _varint(n), _tag(_BLOB_OFFSET, 0), _varint(offset)]
    for field, arr in ((_BLOB_DATA, a), (_BLOB_DIFF, b)):
        if arr is not None:
            raw = memoryview(np.ascontiguousarray(arr, dtype="<f4")).cast("B")
            parts += [_tag(field, 2), _varint(len(raw)), raw]
    return parts


def _len(parts: List) -> int:
    return sum(len(p) for p in parts)


# ---------------------------------------------------------------------------------------------------------
# what this rank owns
# ---------------------------------------------------------------------------------------------------------
def _ownership(pm) -> Tuple[int, int, Optional[List[Tuple[int, int]]]]:
    """(shard id, number of shards, owned element ranges of the arena or None = everything)."""
    be = getattr(pm, "backend", None) if pm is not None else None
    world = getattr(be, "world", 1) if be is not None else 1
    if be is not None and world > 1 and hasattr(be, "owned_ranges"):
        return be.rank, world, sorted(be.owned_ranges())
    return 0, 1, None


def _layer_of(net, param) -> str:
    for l in net.layers:
        if any(q is param for q in l.get_params()):
            return l.name
    return ""


class _HostArena:
    """Pinned host copies of the OWNED ranges of the arena buffers, filled by async copies on a side stream."""

    def __init__(self, arena, ranges: List[Tuple[int, int]]):
        self.ranges = ranges
        self.bufs: Dict[str, List[np.ndarray]] = {}
        cuda = arena.data.is_cuda
        stream = torch.cuda.Stream() if cuda else None
        if cuda:
            stream.wait_stream(torch.cuda.current_stream())
        keep = []
        for name in ("data", "history", "update", "snapshot"):
            t = getattr(arena, name, None)
            if t is None:
                continue
            parts = []
            for lo, hi in ranges:
                h = torch.empty(hi - lo, dtype=torch.float32, pin_memory=cuda)
                if cuda:
                    with torch.cuda.stream(stream):
                        h.copy_(t[lo:hi], non_blocking=True)
                else:
                    h.copy_(t[lo:hi])
                keep.append(h)
                parts.append(h.numpy())
            self.bufs[name] = parts
        if cuda:
            stream.synchronize()
        self._keep = keep

    def slice(self, name: str, lo: int, hi: int) -> Optional[np.ndarray]:
        """Arena elements [lo, hi) of buffer ``name`` (must lie inside one owned range), or None if absent."""
        parts = self.bufs.get(name)
        if parts is None:
            return None
        for (a, b), arr in zip(self.ranges, parts):
            if a <= lo and hi <= b:
                return arr[lo - a:hi - a]
        raise KeyError("range not owned")


def _param_segments(p, arena, ranges) -> List[Tuple[int, int, int]]:
    """[(logical offset, arena lo, arena hi)] of the pieces of parameter ``p`` inside the owned ``ranges``."""
    o, n = arena.offsets[id(p)], arena.slot[id(p)]
    out = []
    for a, b in ranges:
        lo, hi = max(a, o), min(b, o + n)
        if lo < hi:
            out.append((lo - o, lo, hi))       # offset in SLOT coordinates; padded rows are resolved by the caller
    return out


def _logical(p, arr: Optional[np.ndarray], slot_off: int) -> Tuple[int, Optional[np.ndarray]]:
    """Map a piece of a parameter's arena slot (starting at slot element ``slot_off``) to (logical offset, values):
    identity for dense parameters, padding columns dropped for weights stored with a padded row stride."""
    cols = p.shape[-1] if p.shape else 1
    ld = getattr(p, "ld", cols)
    if ld == cols or len(p.shape) != 2:
        if arr is None:
            return slot_off, None
        keep = max(0, min(arr.size, p.size - slot_off))
        return slot_off, arr[:keep]
    idx = np.arange(slot_off, slot_off + (arr.size if arr is not None else 0))
    valid = (idx % ld) < cols
    first = idx[valid][0] if valid.any() else slot_off
    off = int(first // ld) * cols + int(first % ld)
    return off, (arr[valid] if arr is not None else None)


# ---------------------------------------------------------------------------------------------------------
# save
# ---------------------------------------------------------------------------------------------------------
def shard_path(folder: str, step: int, shard: int, nshards: int) -> str:
    if nshards == 1:
        return os.path.join(folder, f"step{int(step)}.bin")
    return os.path.join(folder, f"step{int(step)}.shard{shard}of{nshards}.bin")


def save(folder: str, net, pm, step: int, model_name: str = "") -> Optional[str]:
    """Write this rank's shard of the checkpoint of ``step``; returns its path.  Collective only in the sense that
    every owner must call it for the checkpoint to be complete — no communication takes place."""
    shard, nshards, ranges = _ownership(pm)
    arena = pm.arena if pm is not None else None
    os.makedirs(folder, exist_ok=True)
    path = shard_path(folder, step, shard, nshards)
    owners = [p for p in net.params if p.is_owner]
    host = None
    if arena is not None:
        host = _HostArena(arena, ranges if ranges is not None else [(0, arena.total)])
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(_tag(_CK_STEP, 0) + _varint(int(step)))
        f.write(_tag(_CK_SHARD, 0) + _varint(shard) + _tag(_CK_NSHARDS, 0) + _varint(nshards))
        name_b = model_name.encode()
        f.write(_tag(_CK_NAME, 2) + _varint(len(name_b)) + name_b)
        for p in owners:
            hdr = S.ParamProto()
            hdr.CopyFrom(p.proto)
            del hdr.blob[:]
            del hdr.aux_blob[:]
            lname = _layer_of(net, p)
            hdr.name = f"{lname}/{p.name}" if lname else p.name
            hdr.id = p.id
            del hdr.shape[:]
            hdr.shape.extend(p.shape)
            hdr.version = int(p.version)
            thr = max(1, hdr.split_threshold)
            body: List = [hdr.SerializeToString()]
            if host is not None and id(p) in arena.offsets:
                segs = _param_segments(p, arena, host.ranges)
                pieces = []
                for slot_off, lo, hi in segs:
                    main = [_logical(p, host.slice(k, lo, hi), slot_off) for k in ("data", "history")]
                    aux = [_logical(p, host.slice(k, lo, hi), slot_off) for k in ("update", "snapshot")]
                    pieces.append((main[0][0], main[0][1], main[1][1], aux[0][1], aux[1][1]))
            else:   # no arena (plain tensors): one piece with everything
                flat = lambda t: None if t is None else t.detach().to("cpu", torch.float32).contiguous().view(-1).numpy()  # noqa: E731
                pieces = [(0, flat(p.data), flat(p.history), flat(p.update), flat(p.snapshot))]
            for off, val, hist, upd, snap in pieces:
                n = val.size if val is not None else 0
                for lo in range(0, n, thr):
                    hi = min(n, lo + thr)
                    cut = lambda a: None if a is None else a[lo:hi]   # noqa: E731
                    blob = _blob_bytes(off + lo, cut(val), cut(hist))
                    body += [_tag(_PARAM_BLOB, 2), _varint(_len(blob))] + blob
                    if upd is not None or snap is not None:
                        ab = _blob_bytes(off + lo, cut(upd), cut(snap))
                        body += [_tag(_PARAM_AUX, 2), _varint(_len(ab))] + ab
            f.write(_tag(_CK_PARAM, 2) + _varint(_len(body)))
            for part in body:
                f.write(part)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)
    return path


# ---------------------------------------------------------------------------------------------------------
# load
# ---------------------------------------------------------------------------------------------------------
def _parse_blob(buf, lo: int, hi: int, next_off: int) -> Tuple[int, Optional[np.ndarray], Optional[np.ndarray]]:
    off, data, diff = None, None, None
    for field, wire, a, b in _fields(buf, lo, hi):
        if field == _BLOB_OFFSET and wire == 0:
            off = a
        elif field in (_BLOB_DATA, _BLOB_DIFF):
            if wire == 2:       # packed
                arr = np.frombuffer(buf, dtype="<f4", count=(b - a) // 4, offset=a)
            else:               # one unpacked element (files written by other protobuf encoders)
                arr = np.frombuffer(buf, dtype="<f4", count=1, offset=a)
            if field == _BLOB_DATA:
                data = arr if data is None else np.concatenate([data, arr])
            else:
                diff = arr if diff is None else np.concatenate([diff, arr])
    return (next_off if off is None else off), data, diff     # files without offsets: segments follow one another


def _read_shard(path: str):
    """-> (step, shard, nshards, [(ParamProto header, [(offset, value, history)], [(offset, update, snapshot)])])."""
    buf = np.fromfile(path, dtype=np.uint8)
    mv = memoryview(buf)
    step, shard, nshards, params = 0, 0, 1, []
    for field, wire, a, b in _fields(mv, 0, len(mv)):
        if wire == 0:
            if field == _CK_STEP:
                step = a
            elif field == _CK_SHARD:
                shard = a
            elif field == _CK_NSHARDS:
                nshards = a
        elif field == _CK_PARAM and wire == 2:
            blobs, aux = [], []
            nb = na = 0
            spans = []                       # byte spans of the blob fields: everything else is the small header
            for f2, w2, a2, b2 in _fields(mv, a, b):
                if w2 == 2 and f2 in (_PARAM_BLOB, _PARAM_AUX):
                    off, x, y = _parse_blob(mv, a2, b2, nb if f2 == _PARAM_BLOB else na)
                    (blobs if f2 == _PARAM_BLOB else aux).append((off, x, y))
                    n = x.size if x is not None else (y.size if y is not None else 0)
                    if f2 == _PARAM_BLOB:
                        nb = off + n
                    else:
                        na = off + n
                    spans.append((a2, b2))
            # header = the param message with the blob payloads cut out; their (tag, length) prefixes are cut as well
            header = bytearray()
            pos = a
            for a2, b2 in spans:
                ln = _varint(b2 - a2)
                header += bytes(mv[pos:a2 - len(ln) - 1])
                pos = b2
            header += bytes(mv[pos:b])
            pp = S.ParamProto()
            pp.ParseFromString(bytes(header))
            params.append((pp, blobs, aux))
    return step, shard, nshards, params


def _scatter(dst: Optional[torch.Tensor], p, off: int, arr: Optional[np.ndarray], what: str, name: str) -> None:
    if dst is None or arr is None or arr.size == 0:
        return
    assert off + arr.size <= p.size, f"{name}: {what} segment [{off}, {off + arr.size}) exceeds {p.size} values"
    src = torch.from_numpy(np.ascontiguousarray(arr))
    if dst.is_contiguous():
        dst.view(-1)[off:off + arr.size].copy_(src)
    else:       # padded row stride in the arena (small weights): go through a dense copy
        dense = dst.detach().to("cpu").contiguous()
        dense.view(-1)[off:off + arr.size].copy_(src)
        dst.copy_(dense.to(dst.device))


def step_files(path: str) -> List[str]:
    """All shard files of the checkpoint that ``path`` (any one of its files) belongs to."""
    folder, base = os.path.split(path)
    m = re.match(r"step(\d+)(?:\.shard\d+of(\d+))?\.bin$", base)
    if not m:
        raise ValueError(f"not a checkpoint file name: {path}")
    step = int(m.group(1))
    if m.group(2) is None:
        return [path]
    n = int(m.group(2))
    files = [shard_path(folder, step, r, n) for r in range(n)]
    missing = [f for f in files if not os.path.exists(f)]
    if missing:
        raise FileNotFoundError(f"checkpoint of step {step} is incomplete: missing {missing}")
    return files


def load(path: str, net, pm=None) -> int:
    """Restore parameters (+ optimizer state) from the checkpoint ``path`` belongs to (every shard file is read, so the
    number of ranks may differ from the run that saved); returns the step to continue from."""
    owners = [p for p in net.params if p.is_owner]
    by_name = {f"{_layer_of(net, p)}/{p.name}": p for p in owners}
    by_id = {p.id: p for p in owners}
    seen: Dict[int, int] = {id(p): 0 for p in owners}
    step = 0
    for f in step_files(path):
        step, _shard, _n, params = _read_shard(f)
        for pp, blobs, aux in params:
            p = by_name.get(pp.name) or by_id.get(pp.id)
            if p is None:
                continue
            assert list(pp.shape) == list(p.shape), \
                f"checkpoint shape {list(pp.shape)} != param shape {p.shape} ({pp.name})"
            for off, val, hist in blobs:
                _scatter(p.data, p, off, val, "value", pp.name)
                _scatter(p.history, p, off, hist, "history", pp.name)
                seen[id(p)] += val.size if val is not None else 0
            for off, upd, snap in aux:
                _scatter(p.update, p, off, upd, "update", pp.name)
                _scatter(p.snapshot, p, off, snap, "snapshot", pp.name)
            p.version = pp.version
    for p in owners:
        if seen[id(p)] != p.size:
            raise KeyError(f"checkpoint {path}: parameter {p.id} ({p.name}) has {seen[id(p)]} of {p.size} values")
    if pm is not None:
        pm.arena.sync_mirror()
    return int(step)


def latest(folder: str) -> Optional[str]:
    """Newest COMPLETE checkpoint in ``folder`` (a path usable with ``load``), or None."""
    steps: Dict[int, str] = {}
    for f in glob.glob(os.path.join(folder, "step*.bin")):
        m = re.search(r"step(\d+)(?:\.shard\d+of\d+)?\.bin$", f)
        if m:
            steps.setdefault(int(m.group(1)), f)
    for s in sorted(steps, reverse=True):
        try:
            return sorted(step_files(steps[s]))[0]
        except FileNotFoundError:
            continue
    return None
