"""Library baseline for the parameter server: per push bucket, NCCL reduce-scatter -> Updater on the owned piece ->
NCCL all-gather, issued on a communication stream while backward continues on the compute stream (SURVEY §7.2
step 6).  Same buckets, same striped ownership and same updater kernel as the fused path
(``parallel/fused_ps.py``); only the collectives are plain ``torch.distributed`` calls — "a path that only calls
NCCL is the baseline, not the product".

Per bucket (a run of adjacent layers whose chunk count is a multiple of the world size, see ``ParamArena``):
  comm stream waits for the bucket's gradients (compute stream + weight-gradient side stream)
  reduce_scatter_tensor, in place:   grad[piece r]  <- mean over ranks of grad[bucket]
  update_arena over piece r          (fp32 master + optimizer state live on the owner only)
  all_gather_into_tensor, in place:  mirror[bucket] <- every owner's fresh piece  (fp32 compute: the values themselves)
The compute stream joins the comm stream at the end of the step (``end_step``), as the fused path does.
"""
from __future__ import annotations

import os
from typing import List, Tuple

import torch
import torch.distributed as dist

from ..device import ctx
from .fused_ps import make_buckets


class NcclExchange:
    handles_update = True

    def __init__(self, pm):
        self.pm = pm
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.arena = pm.arena
        assert self.arena.nchunks % self.world == 0, "arena must be padded to a multiple of the world size"
        self.layers = list(pm._param_layers)
        bucket_elems = int(os.environ.get("SINGA_B200_PUSH_BUCKET", str(6_000_000)))
        names = make_buckets([(l.name, *pm.layer_ranges[l.name]) for l in self.layers], bucket_elems,
                             first_alone=os.environ.get("SINGA_B200_PUSH_FIRST_ALONE", "1") != "0")
        by_name = {l.name: l for l in self.layers}
        self.buckets: List[List] = [[by_name[n] for n in b] for b in names]
        self.bucket_of = {l.name: b for b, ls in enumerate(self.buckets) for l in ls}
        self.trigger = {ls[-1].name for ls in self.buckets}
        self.ranges: List[Tuple[int, int, int]] = []     # per bucket: (first chunk, chunks, chunks per rank)
        for ls in self.buckets:
            lo = min(pm.layer_ranges[l.name][0] for l in ls)
            n = sum(pm.layer_ranges[l.name][1] for l in ls)
            assert n % self.world == 0, "bucket is not a multiple of the world size (arena padding rule broken)"
            self.ranges.append((lo, n, n // self.world))
        on_gpu = ctx().device.type == "cuda"
        self.comm = torch.cuda.Stream() if on_gpu else None
        self._used = False
        # per bucket: recorded on the comm stream behind the all-gather; a layer's forward waits for its bucket's event
        # (needed by the executor's rotated schedule, where the next forward pass starts before the comm stream is joined)
        self._gathered = [torch.cuda.Event() if on_gpu else None for _ in self.buckets]
        self._gathered_valid = [False] * len(self.buckets)
        if self.world > 1 and self.arena.data16 is not None:
            for p in self.arena.params:
                p.use_mirror = True
                if p.pack16 is not None:
                    p.pack_from_mirror = True

    # the buffer every rank reads its operands from: the mirror if there is one, else the fp32 values themselves
    def _gather_buf(self) -> torch.Tensor:
        return self.arena.data16 if self.arena.data16 is not None else self.arena.data

    def update_layer(self, layer, step: int, side_stream=None) -> None:
        b = self.bucket_of[layer.name]
        if self.comm is not None:
            self.comm.wait_stream(torch.cuda.current_stream())
            if side_stream is not None:
                self.comm.wait_stream(side_stream)
            self._used = True
        if layer.name not in self.trigger:
            return
        lo, n, per = self.ranges[b]
        ar = self.arena
        mine_lo = (lo + self.rank * per) * 1024
        mine_hi = mine_lo + per * 1024
        stream_ctx = torch.cuda.stream(self.comm) if self.comm is not None else _Null()
        with stream_ctx:
            if self.world > 1:
                dist.reduce_scatter_tensor(ar.grad[mine_lo:mine_hi], ar.grad[lo * 1024:(lo + n) * 1024],
                                           op=dist.ReduceOp.AVG)
            self.pm._update_range(lo + self.rank * per, per, step)
            if self.world > 1:
                buf = self._gather_buf()
                dist.all_gather_into_tensor(buf[lo * 1024:(lo + n) * 1024], buf[mine_lo:mine_hi])
                for l in self.buckets[b]:
                    for p in l.get_params():
                        if p.pack16 is not None:
                            p.pack16()
            if self._gathered[b] is not None:
                self._gathered[b].record(self.comm)
                self._gathered_valid[b] = True

    def end_step(self, step: int) -> None:
        if self.comm is not None and self._used:
            torch.cuda.current_stream().wait_stream(self.comm)
        self._used = False

    def wait_layer(self, layer, step: int) -> None:
        b = self.bucket_of.get(layer.name)
        if b is not None and self._gathered[b] is not None and self._gathered_valid[b]:
            torch.cuda.current_stream().wait_event(self._gathered[b])

    def pre_forward(self, step: int) -> None:
        pass

    def owned_ranges(self) -> List[Tuple[int, int]]:
        return [((lo + self.rank * per) * 1024, (lo + self.rank * per + per) * 1024) for lo, n, per in self.ranges]

    def gather_master(self) -> None:
        from .fused_ps import gather_owned
        gather_owned(self.arena, self.owned_ranges(), self.world)

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
            dist.barrier()


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
