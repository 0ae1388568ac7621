"""Asynchronous worker groups with bounded staleness over peer memory (BASELINE config 3).

Each rank is a worker group running LOCAL updates every step (the reference's main mode: "local SGD on the worker
plus periodic Elastic-averaging or RandomSync synchronisation with the servers", src/utils/param_manager.cc:158-234,
src/utils/param.cc:130-284).  The servers are shards of a peer-mapped fp32 "center" arena: shard s lives on GPU s.
Every ``sync_frequency`` steps a group exchanges with every shard by launching the EASGD (or RandomSync) kernel
directly on the owner's memory — ``atomicAdd`` over NVLink, so concurrent groups are safe and no host message or
server thread exists.

Bounded staleness without host messages: every group release-stores its step counter into its own ``progress``
flag after each step; before starting step t a group acquire-waits (on its compute stream) until every group's
progress is at least ``t - staleness``.  ``staleness = 0`` degenerates to lock-step; the reference's implicit bound
is ``sync_frequency`` local steps.
"""
from __future__ import annotations

import ctypes
from typing import List, Tuple

import torch
import torch.distributed as dist

from ..device import ctx, step_state
from ..ops import kernels as K
from ..ops import native
from .fused_ps import shard_bounds


def async_regions(total_elems: int, world: int):
    return [("center", total_elems * 4), ("progress", 64)]


class AsyncExchange:
    handles_update = False

    def __init__(self, pm, heap):
        self.pm, self.heap = pm, heap
        self.lib = native.require()
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.arena = pm.arena
        self.staleness = max(0, int(pm.cluster.proto.staleness)) if pm.cluster is not None else 0
        self.shards: List[Tuple[int, int]] = [(lo * 1024, hi * 1024) for lo, hi in
                                             shard_bounds(self.arena.nchunks, self.world)]
        self.center_local = heap.tensor("center", torch.float32)
        self.center_peer = [heap.tensor("center", torch.float32, r) for r in range(self.world)]
        self.step_dev = step_state()["step"]
        self._wait = (ctypes.c_void_p * 32)()
        for r in range(self.world):
            self._wait[r] = heap.ptr("progress", r)
        self.nsync = 0
        self.seed_base = 0x5EED

    def seed_center(self) -> None:
        """Every owner copies its shard of the (identically initialised) weights into the center."""
        lo, hi = self.shards[self.rank]
        self.center_local[lo:hi].copy_(self.arena.data[lo:hi])
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()

    # -- reference API ---------------------------------------------------------------------------------
    def put_all(self) -> None:
        self.seed_center()

    def get_all(self) -> None:
        """A joining group pulls every shard of the center (kGet)."""
        for o, (lo, hi) in enumerate(self.shards):
            if hi > lo:
                self.arena.data[lo:hi].copy_(self.center_peer[o][lo:hi])
        self.arena.sync_mirror()

    def begin_step(self, step: int) -> None:
        # bounded staleness: wait until every group has finished step (step - 1 - staleness)
        need = -self.staleness
        native.check(self.lib.sg_ps_flag_wait(self._wait, self.world, ctypes.c_void_p(self.step_dev.data_ptr()),
                                              need, None, torch.cuda.current_stream().cuda_stream),
                     "sg_ps_flag_wait(progress)")
        native.count_launch()

    def end_step(self, step: int) -> None:
        native.check(self.lib.sg_ps_flag_set(ctypes.c_void_p(self.heap.ptr("progress", self.rank)),
                                             ctypes.c_void_p(self.step_dev.data_ptr()), 1,
                                             torch.cuda.current_stream().cuda_stream), "sg_ps_flag_set(progress)")
        native.count_launch()

    def sync_all(self, step: int) -> None:
        ar = self.arena
        for o, (lo, hi) in enumerate(self.shards):
            if hi <= lo:
                continue
            w = ar.data[lo:hi]
            w16 = ar.data16[lo:hi] if ar.data16 is not None else None
            server = self.center_peer[o][lo:hi]
            if self.pm.param_type == "RandomSync":
                K.randomsync(w, ar.snapshot[lo:hi], server, self.pm.sample_ratio,
                             self.seed_base + step * 131 + self.rank, w16)
            else:
                K.easgd_sync(w, server, self.pm.moving_rate, w16)
        for p in ar.params:
            if p.pack16 is not None:
                p.pack16()
        self.nsync += 1

    def wait_layer(self, layer, step: int) -> None:
        pass

    def progress(self) -> List[int]:
        torch.cuda.synchronize()
        return [int(self.heap.tensor("progress", torch.int64, r)[0].item()) for r in range(self.world)]

    def close(self) -> None:
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        self.heap.close()
