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
        cproto = pm.cluster.proto if pm.cluster is not None else None
        self.staleness = max(0, int(cproto.staleness)) if cproto is not None else 0
        # Server groups (src/server/pm_server.cc:105-127, Topology.server_group): the ranks are cut into `ngroups`
        # groups; each group keeps its own center copy, sharded over its `gsize` ranks, and the workers of a group
        # exchange with their own group's shards only.  Shard l of group g lives on rank g * gsize + l.
        self.ngroups = max(1, int(cproto.nserver_groups)) if cproto is not None else 1
        assert self.world % self.ngroups == 0, "nserver_groups must divide the number of ranks"
        self.gsize = self.world // self.ngroups
        self.group, self.local = self.rank // self.gsize, self.rank % self.gsize
        self.sync_interval = max(1, int(cproto.server_sync_interval)) if cproto is not None else 1
        assert self.arena.total % self.gsize == 0
        self.per = self.arena.total // self.gsize
        self.shards: List[Tuple[int, int]] = [(k * self.per, (k + 1) * self.per) for k in range(self.gsize)]
        self.center_local = heap.tensor("center", torch.float32)
        self.center_peer = [heap.tensor("center", torch.float32, self.group * self.gsize + k)
                            for k in range(self.gsize)]
        # ONE launch exchanges the whole arena with every shard of the group (pointer table in the kernel parameters)
        self.table = K.shard_table([heap.ptr("center", self.group * self.gsize + k) for k in range(self.gsize)],
                                   self.per)
        # the same shard in the other server groups: what the neighbour synchronisation averages with
        self.neighbours = K.shard_table([heap.ptr("center", g * self.gsize + self.local)
                                         for g in range(self.ngroups) if g != self.group], self.per) \
            if self.ngroups > 1 else None
        self.step_dev = step_state()["step"]
        self._wait = (ctypes.c_void_p * 32)()
        for r in range(self.world):
            self._wait[r] = heap.ptr("progress", r)
        self.nsync = 0
        self.ngroup_sync = 0
        self.seed_base = 0x5EED

    def seed_center(self) -> None:
        """Every owner copies its shard of the (identically initialised) weights into its group's center."""
        lo, hi = self.shards[self.local]
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
        w16 = ar.data16 if (ar.data16 is not None and ar.data16.dtype == torch.bfloat16) else None
        if self.pm.param_type == "RandomSync":
            # the subset changes per step through the DEVICE step scalar, so the launch can live in a captured graph
            K.randomsync_sharded(ar.data, ar.snapshot, self.table, self.pm.sample_ratio,
                                 self.seed_base + self.rank, self.step_dev, w16)
        else:
            K.easgd_sync_sharded(ar.data, self.table, self.pm.moving_rate, w16)
        for p in ar.params:
            if p.pack16 is not None:
                p.pack16()
        # server-group neighbour synchronisation (ParamShard::sync_now: every sync_interval updates of the shard);
        # decided from the step number alone so that a captured graph of "a step that synchronises" stays valid
        if self.group_sync_due(step):
            lo, hi = self.shards[self.local]
            K.group_average(self.center_local, self.neighbours, lo, hi - lo)
        self.account(step)

    def group_sync_due(self, step: int) -> bool:
        if self.neighbours is None:
            return False
        k = (step + 1) // self.pm.sync_frequency          # ordinal of this exchange
        return k % self.sync_interval == 0

    def account(self, step: int) -> None:
        """Host-side counters of one synchronising step (also called for steps replayed from a CUDA graph)."""
        self.nsync += 1
        if self.group_sync_due(step):
            self.ngroup_sync += 1

    def graph_variant(self, step: int):
        """Steps with the same variant launch the same kernels: plain / worker exchange / + neighbour averaging."""
        if not self.pm.sync_now(step):
            return 0
        return 2 if self.group_sync_due(step) else 1

    def after_replay(self, step: int) -> None:
        if self.pm.sync_now(step):
            self.account(step)

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
