"""Bridge transport: moves activations (forward) and gradients (backward) across a location boundary of a
partitioned net.

Reference: the PUSH/PULL sockets of src/worker/worker.cc:139-155,216-261,293-299 (messages M7/M8 of the
survey).  On one box the two ends are either in the same process (several partitions executed by one
rank: a device copy) or on two GPU ranks of the group (``torch.distributed`` point-to-point over
NCCL/NVLink, or gloo on CPU).  A location id maps to rank ``group_base + location // nthreads_per_procs``.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist


class BridgeTransport:
    def __init__(self, cluster, local_threadid: int = 0):
        self.cluster = cluster
        self.local_threadid = local_threadid
        self.distributed = dist.is_available() and dist.is_initialized() and cluster is not None \
            and cluster.nprocs_per_group > 1
        self.rank = dist.get_rank() if self.distributed else 0

    def _rank_of(self, layer) -> int:
        c = self.cluster
        base = c.global_procsid - c.group_procsid
        return base + c.group_procsid_of(layer.locationid)

    def is_local(self, layer) -> bool:
        if not self.distributed:
            return True
        return self._rank_of(layer) == self.rank

    # forward: BridgeSrc (on the producer's rank) -> BridgeDst (on the consumer's rank)
    def send_data(self, src_layer) -> None:
        dst_layer = src_layer.dstlayers[0]
        if not self.distributed or self.is_local(dst_layer):
            dst_layer.data.copy_(src_layer.data)
            dst_layer.set_ready(True)
        else:
            dist.send(src_layer.data.contiguous(), self._rank_of(dst_layer))

    def recv_data(self, dst_layer) -> None:
        src_layer = dst_layer.srclayers[0]
        if self.distributed and not self.is_local(src_layer):
            buf = dst_layer.data if dst_layer.data.is_contiguous() else torch.empty_like(dst_layer.data).contiguous()
            dist.recv(buf, self._rank_of(src_layer))
            if buf is not dst_layer.data:
                dst_layer.data.copy_(buf)
            dst_layer.set_ready(True)
        # same process: the producer already copied into our buffer

    # backward: BridgeDst -> BridgeSrc
    def send_grad(self, dst_layer) -> None:
        src_layer = dst_layer.srclayers[0]
        if not self.distributed or self.is_local(src_layer):
            src_layer.grad.copy_(dst_layer.grad)
            src_layer.set_ready(True)
        else:
            dist.send(dst_layer.grad.contiguous(), self._rank_of(src_layer))

    def recv_grad(self, src_layer) -> None:
        dst_layer = src_layer.dstlayers[0]
        if self.distributed and not self.is_local(dst_layer):
            buf = src_layer.grad if src_layer.grad.is_contiguous() else torch.empty_like(src_layer.grad).contiguous()
            dist.recv(buf, self._rank_of(dst_layer))
            if buf is not src_layer.grad:
                src_layer.grad.copy_(buf)
            src_layer.set_ready(True)
