"""Library baseline for the parameter server: NCCL reduce-scatter -> Updater on the owned shard -> NCCL
all-gather.  Same sharding and same updater kernel as the fused path, but the collectives
are plain ``torch.distributed`` calls — "a path that only calls NCCL is the baseline, not the product"."""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..device import ctx


class NcclExchange:
    handles_update = True

    def __init__(self, pm):
        self.pm = pm
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.arena = pm.arena
        assert self.arena.nchunks % self.world == 0, "arena must be padded to a multiple of the world size"
        self.per = self.arena.nchunks // self.world
        self.lo = self.rank * self.per * 1024
        self.hi = self.lo + self.per * 1024
        self.last_layer = pm._first_param_layer   # its gradient is the last one produced by backward
        if self.world > 1 and self.arena.data16 is not None:
            for p in self.arena.params:
                p.use_mirror = True
                if p.pack16 is not None:
                    p.pack_from_mirror = True
        self.gshard = torch.empty(self.per * 1024, dtype=torch.float32, device=self.arena.grad.device)

    def update_layer(self, layer, step: int, side_stream=None) -> None:
        if layer is not self.last_layer:
            return
        ar = self.arena
        if self.world > 1:
            dist.reduce_scatter_tensor(self.gshard, ar.grad, op=dist.ReduceOp.AVG)
            ar.grad[self.lo:self.hi].copy_(self.gshard)
        self.pm._update_range(self.rank * self.per, self.per, step)
        if self.world > 1:
            if ar.data16 is not None:
                dist.all_gather_into_tensor(ar.data16, ar.data16[self.lo:self.hi])
                for p in ar.params:
                    if p.pack16 is not None:
                        p.pack16()
            else:
                dist.all_gather_into_tensor(ar.data, ar.data[self.lo:self.hi])

    def wait_layer(self, layer, step: int) -> None:
        pass

    def gather_master(self) -> None:
        from .fused_ps import gather_owned
        gather_owned(self.arena, [(self.lo, self.hi)], self.world)

    def put_all(self):
        pass

    def get_all(self):
        if self.world > 1:
            dist.broadcast(self.arena.data, 0)
            self.arena.sync_mirror()

    def sync_all(self, step):
        pass

    def close(self):
        if ctx().device.type == "cuda":
            torch.cuda.synchronize()
