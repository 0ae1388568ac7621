"""Updaters and learning-rate schedules (reference: include/utils/updater.h:15-76, src/utils/updater.cc).

``Updater.update(step, param)`` is the reference's per-parameter call (CPU path, parameter-server side);
``Updater.update_arena(step, arena, first_chunk, nchunks)`` is the B200 form: one fused multi-tensor kernel
over a chunk range of the flat arena that also refreshes the bf16 mirror.
"""
from __future__ import annotations

from typing import Sequence

import torch

from .ops import reference as R
from .proto import schema as S

_KIND_ID = {"kSGD": 0, "kNesterov": 1, "kAdaGrad": 2, "kRMSProp": 3, "kAdaDelta": 4}


class Updater:
    kind = "kSGD"

    def __init__(self, proto):
        self.proto = S.UpdaterProto()
        self.proto.CopyFrom(proto)
        self.momentum = proto.momentum
        self.weight_decay = proto.weight_decay
        self.delta = proto.delta
        self.rho = proto.rho
        if self.needs_lr and not proto.base_learning_rate > 0:
            raise ValueError(f"{self.kind}: base_learning_rate must be > 0")

    needs_lr = True
    needs_history = True
    needs_update = False

    def learning_rate(self, step: int) -> float:
        return R.learning_rate(self.proto, step) if self.needs_lr else 0.0

    # reference API -----------------------------------------------------------------------------
    def update(self, step: int, param, grad_scale: float = 1.0) -> None:
        if self.needs_history and param.history is None:
            param.history = torch.zeros_like(param.data)
        if self.needs_update and param.update is None:
            param.update = torch.zeros_like(param.data)
        if step == 0:
            if param.history is not None:
                param.history.zero_()
            if param.update is not None:
                param.update.zero_()
        lr = self.learning_rate(step) * param.learning_rate_multiplier
        wd = self.weight_decay * param.weight_decay_multiplier
        R.update_ref(self.kind, param.data, param.grad, param.history, param.update, lr=lr,
                     momentum=self.momentum, wd=wd, delta=self.delta, rho=self.rho, grad_scale=grad_scale)
        param.version = step + 1

    # arena API ---------------------------------------------------------------------------------
    def update_arena(self, step: int, arena, first_chunk: int = 0, nchunks: int | None = None,
                     grad_scale: float = 1.0, params: Sequence | None = None) -> None:
        if nchunks is None:
            nchunks = arena.nchunks - first_chunk
        if nchunks <= 0:
            return
        if arena.use16:
            from .ops import kernels as K
            K.update_arena(arena.data, arena.grad, arena.history, arena.update, arena.data16,
                           arena.chunk_param, arena.lr_mult, arena.wd_mult, first_chunk, nchunks,
                           kind=_KIND_ID[self.kind], lr=self.learning_rate(step), momentum=self.momentum,
                           weight_decay=self.weight_decay, delta=self.delta, rho=self.rho,
                           grad_scale=grad_scale)
            lo, hi = first_chunk * 1024, (first_chunk + nchunks) * 1024
            for p in (params if params is not None else arena.params):
                if p.pack16 is not None and lo <= arena.offsets[id(p if p.is_owner else p.owner)] < hi:
                    p.pack16()
        else:
            lo, hi = first_chunk * 1024, (first_chunk + nchunks) * 1024
            for p in arena.params:
                if lo <= arena.offsets[id(p)] < hi:
                    self.update(step, p, grad_scale)


class SGDUpdater(Updater):
    kind = "kSGD"


class NesterovUpdater(Updater):
    kind = "kNesterov"


class AdaGradUpdater(Updater):
    kind = "kAdaGrad"


class RMSPropUpdater(Updater):
    kind = "kRMSProp"


class AdaDeltaUpdater(Updater):
    kind = "kAdaDelta"
    needs_lr = False
    needs_update = True


_CLASSES = {"kSGD": SGDUpdater, "kNesterov": NesterovUpdater, "kAdaGrad": AdaGradUpdater,
            "kRMSProp": RMSPropUpdater, "kAdaDelta": AdaDeltaUpdater}


def create_updater(proto) -> Updater:
    """Pick the updater class from ``UpdaterProto.type`` (src/utils/param_manager.cc:19-38)."""
    name = S.enum_name("UpdaterProto.Type", proto.type)
    return _CLASSES[name](proto)
