"""Checkpoint / resume in the Param-proto layout.

The reference declares checkpointing but never implements it (``Worker::Resume`` is an empty TODO,
src/worker/worker.cc:65-67; ``ParamProto`` has its value field commented out, src/proto/model.proto:69-70;
``kPretrained`` has no case in ``Param::Init``).  This module supplies the intended design:

  file   = binary ``ParamCheckpoint { step, model_name, repeated ParamProto param }``
  param  = name / id / shape + ``blob[]`` where each ``BlobProto`` carries one segment of at most
           ``split_threshold`` floats: ``data`` = value, ``diff`` = updater history; ``aux_blob[]`` likewise
           holds ``data`` = AdaDelta update, ``diff`` = RandomSync snapshot.

Optimizer state is included so a resumed run continues exactly.  Files are written atomically
(``write_proto_to_binary_file``) as ``<folder>/step<N>.bin``; ``latest`` finds the newest one.  With a sharded
parameter server the fp32 master copy is first gathered from the shard owners.
"""
from __future__ import annotations

import glob
import os
import re
from typing import Optional

import numpy as np
import torch

from .proto import schema as S
from .utils.common import read_proto_from_binary_file, write_proto_to_binary_file


def _flat_cpu(t: Optional[torch.Tensor]) -> Optional[np.ndarray]:
    if t is None:
        return None
    return t.detach().to("cpu", torch.float32).contiguous().view(-1).numpy()


def _fill_blobs(blobs, a: Optional[np.ndarray], b: Optional[np.ndarray], threshold: int) -> None:
    n = a.size if a is not None else (b.size if b is not None else 0)
    for lo in range(0, n, threshold):
        hi = min(n, lo + threshold)
        bp = blobs.add()
        bp.num = hi - lo
        if a is not None:
            bp.data.extend(a[lo:hi].tolist())
        if b is not None:
            bp.diff.extend(b[lo:hi].tolist())


def param_to_proto(p, layer_name: str = "") -> "S.ParamProto":
    pp = S.ParamProto()
    pp.CopyFrom(p.proto)
    pp.name = f"{layer_name}/{p.name}" if layer_name else p.name
    pp.id = p.id
    del pp.shape[:]
    pp.shape.extend(p.shape)
    pp.version = int(p.version)
    thr = max(1, pp.split_threshold)
    _fill_blobs(pp.blob, _flat_cpu(p.data), _flat_cpu(p.history), thr)
    if p.update is not None or p.snapshot is not None:
        _fill_blobs(pp.aux_blob, _flat_cpu(p.update), _flat_cpu(p.snapshot), thr)
    return pp


def _gather(blobs, field: str) -> Optional[np.ndarray]:
    parts = [np.asarray(getattr(b, field), dtype=np.float32) for b in blobs]
    parts = [x for x in parts if x.size]
    return np.concatenate(parts) if parts else None


def proto_to_param(pp, p) -> None:
    assert list(pp.shape) == list(p.shape), f"checkpoint shape {list(pp.shape)} != param shape {p.shape} ({pp.name})"

    def put(dst: Optional[torch.Tensor], arr: Optional[np.ndarray], what: str):
        if arr is None or dst is None:
            return
        assert arr.size == p.size, f"{pp.name}: {what} has {arr.size} values, expected {p.size}"
        dst.copy_(torch.from_numpy(arr).view(p.shape).to(dst.device))

    put(p.data, _gather(pp.blob, "data"), "value")
    put(p.history, _gather(pp.blob, "diff"), "history")
    put(p.update, _gather(pp.aux_blob, "data"), "update")
    put(p.snapshot, _gather(pp.aux_blob, "diff"), "snapshot")
    p.version = pp.version


def _layer_of(net, param) -> str:
    for l in net.layers:
        if any(q is param for q in l.get_params()):
            return l.name
    return ""


def save(folder: str, net, pm, step: int, model_name: str = "") -> Optional[str]:
    """Write ``<folder>/step<step>.bin``; returns the path (None on ranks that do not write)."""
    if pm is not None and pm.backend is not None and hasattr(pm.backend, "gather_master"):
        pm.backend.gather_master()          # collective: every rank takes part
        import torch.distributed as dist
        if dist.is_initialized() and dist.get_rank() != 0:
            return None
    os.makedirs(folder, exist_ok=True)
    ck = S.ParamCheckpoint()
    ck.step = int(step)
    ck.model_name = model_name
    for p in net.params:
        if not p.is_owner:
            continue
        ck.param.add().CopyFrom(param_to_proto(p, _layer_of(net, p)))
    path = os.path.join(folder, f"step{int(step)}.bin")
    write_proto_to_binary_file(ck, path)
    return path


def load(path: str, net, pm=None) -> int:
    """Restore parameters (+ optimizer state) from ``path``; returns the step to continue from."""
    ck = read_proto_from_binary_file(path, S.ParamCheckpoint())
    by_id = {pp.id: pp for pp in ck.param}
    by_name = {pp.name: pp for pp in ck.param}
    for p in net.params:
        if not p.is_owner:
            continue
        pp = by_name.get(f"{_layer_of(net, p)}/{p.name}") or by_id.get(p.id)
        if pp is None:
            raise KeyError(f"checkpoint {path} has no entry for parameter {p.id} ({p.name})")
        proto_to_param(pp, p)
    if pm is not None:
        pm.arena.sync_mirror()
    return int(ck.step)


def latest(folder: str) -> Optional[str]:
    best, best_step = None, -1
    for f in glob.glob(os.path.join(folder, "step*.bin")):
        m = re.search(r"step(\d+)\.bin$", f)
        if m and int(m.group(1)) > best_step:
            best, best_step = f, int(m.group(1))
    return best
