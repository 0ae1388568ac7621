"""The B200 parameter server: gradient push and parameter pull as kernels over NVSwitch peer memory.

Every rank is both a worker group and the SERVER for one contiguous chunk range ("shard") of the flat
parameter arena.  Per layer, right after its backward (the reference's ``UpdateParam`` call site,
src/worker/worker.cc:290-292):

  1. the worker release-stores its ``ready[layer]`` flag (gradients complete, bf16 weights no longer read);
  2. the server side — on a separate comm stream, so it overlaps the backward of the layers below — launches
     ``sg_ps_push_update`` over (layer range ∩ own shard): reduce the peers' gradients through peer loads,
     apply the Updater to the fp32 master copy, broadcast the new bf16 weights into every rank's mirror,
     bump ``version[layer][server]`` on every rank.

Right before the layer's next forward (the ``WaitUpdate`` call site, worker.cc:249-253) the worker's compute
stream acquires the version flags of the shards that intersect the layer — a device-side wait, no host
message, no host synchronisation.  This is messages M2-M5/M9 of the survey without ZeroMQ.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Tuple

import torch
import torch.distributed as dist

from ..device import ctx, step_state
from ..ops import native
from ..ops._decl import Updater as CUpdater
from ..updater import _KIND_ID
from .symm import SymmetricHeap

MAX_WORLD = 8
MAX_WAIT = 32


class PsPush(ctypes.Structure):
    _fields_ = [
        ("world", ctypes.c_int),
        ("grad", ctypes.c_void_p * MAX_WORLD),
        ("w16", ctypes.c_void_p * MAX_WORLD),
        ("ready", ctypes.c_void_p * MAX_WORLD),
        ("version", ctypes.c_void_p * MAX_WORLD),
        ("w", ctypes.c_void_p), ("h", ctypes.c_void_p), ("u", ctypes.c_void_p),
        ("chunk_param", ctypes.c_void_p), ("lr_mult", ctypes.c_void_p), ("wd_mult", ctypes.c_void_p),
        ("first_chunk", ctypes.c_longlong), ("nchunks", ctypes.c_longlong),
        ("up", CUpdater),
        ("step_dev", ctypes.c_void_p), ("value_add", ctypes.c_longlong),
        ("scratch", ctypes.c_void_p),
        ("self_count", ctypes.c_int),
        ("max_ctas", ctypes.c_int),
        ("grad_mc", ctypes.c_void_p),
        ("w16_mc", ctypes.c_void_p),
        ("bump_ready", ctypes.c_void_p),
        ("bump_seq", ctypes.c_void_p),
        ("mirror_fp32", ctypes.c_int),
        ("trace", ctypes.c_void_p),
    ]


def shard_bounds(nchunks: int, world: int) -> List[Tuple[int, int]]:
    per = (nchunks + world - 1) // world
    return [(min(r * per, nchunks), min((r + 1) * per, nchunks)) for r in range(world)]


def make_buckets(layers, bucket_elems: int, chunk: int = 1024, first_alone: bool = False) -> List[List[str]]:
    """Group parameter layers into push buckets.  ``layers``: [(name, first_chunk, nchunks)] in FORWARD order.  Walks
    in backward order (last layer first) merging layers that are adjacent in the arena until a bucket holds at least
    ``bucket_elems`` parameters; returns the buckets as lists of names in backward order, so ``bucket[-1]`` is the
    layer whose gradient arrives last — the one that triggers the push.  A bucket always covers one contiguous chunk
    range (a gap in the arena closes the current bucket).  ``first_alone``: the first layer of the net — whose gradient
    is the last one backward produces, so its push is the one nothing can hide — always forms its own bucket, keeping
    the exposed push as small as possible."""
    buckets: List[List[str]] = []
    cur: List[str] = []
    cur_n, cur_lo = 0, None
    layers = list(layers)
    for idx, (name, lo, n) in enumerate(reversed(layers)):
        if cur and (cur_lo != lo + n or (first_alone and idx == len(layers) - 1)):
            buckets.append(cur)
            cur, cur_n = [], 0
        cur.append(name)
        cur_n += n * chunk
        cur_lo = lo
        if cur_n >= bucket_elems:
            buckets.append(cur)
            cur, cur_n = [], 0
    if cur:
        buckets.append(cur)
    return buckets


def symmetric_regions(total_elems: int, nlayers: int, world: int, mirror_bytes: int = 2):
    """Regions of the symmetric heap for an arena of ``total_elems`` elements.  ``mirror_bytes``: element size of the
    operand mirror the servers broadcast into (2 = bf16, 4 = fp32, 0 = none: one fp32 rank reads its master copy)."""
    regions = [("grad", total_elems * 4)]
    if mirror_bytes:
        regions.append(("data16", total_elems * mirror_bytes))
    return regions + [("ready", max(nlayers, 1) * 8), ("version", max(nlayers, 1) * world * 8)]


def gather_owned(arena, ranges, world: int) -> None:
    """Every rank contributes the ranges it owns (zeros elsewhere); a sum all-reduce rebuilds the full buffers."""
    torch.cuda.synchronize()
    if world <= 1:
        return
    for buf in (arena.data, arena.history, arena.update):
        if buf is None:
            continue
        tmp = torch.zeros_like(buf)
        for lo, hi in ranges:
            tmp[lo:hi].copy_(buf[lo:hi])
        dist.all_reduce(tmp)
        buf.copy_(tmp)
    torch.cuda.synchronize()


class FusedExchange:
    handles_update = True

    def __init__(self, pm, heap: SymmetricHeap):
        self.pm = pm
        self.heap = heap
        self.lib = native.require()
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        assert self.world <= MAX_WORLD, f"fused push/pull supports up to {MAX_WORLD} ranks per node"
        self.arena = pm.arena
        self.has_mirror = self.arena.data16 is not None
        self.mirror_fp32 = self.has_mirror and self.arena.data16.dtype == torch.float32
        self.layers = list(pm._param_layers)
        self.shards = shard_bounds(self.arena.nchunks, self.world)
        import os
        # two comm streams, alternating buckets: a push starts when ITS gradients are ready instead of queueing behind
        # the previous bucket's rendezvous
        self.comms = [torch.cuda.Stream() for _ in range(int(os.environ.get("SINGA_B200_COMM_STREAMS", "2")))]
        self.max_ctas = int(os.environ.get("SINGA_B200_PUSH_CTAS", "296"))
        # NVLS: in-switch reduction / broadcast when the heap has a multicast mapping (torch symmetric memory)
        # (measured on 8xB200, MNIST MLP: NVLS 0.520 ms/step vs peer loads 0.625 at 8 ranks; 0.469 vs 0.444 at 2)
        nvls_env = os.environ.get("SINGA_B200_NVLS", "auto")
        want = self.world >= 4 if nvls_env == "auto" else nvls_env != "0"
        self.use_nvls = want and bool(getattr(heap, "mc_ptr", None)) and heap.mc_ptr("grad") != 0
        # Push BUCKETS: every push costs ~15-20 us of pure rendezvous latency (peer flag poll, system-scope fences,
        # remote version stores) however small the layer is, and the pushes of a step serialise on the comm stream
        # (measured at 2 GPUs: 6 pushes = 210 us against a 150 us backward pass).  Adjacent layers are therefore
        # merged, walking in backward order, until a bucket holds SINGA_B200_PUSH_BUCKET parameters (default 6 M: measured
        # 0.4425 ms at 8 GPUs / 0.374 at 2, against 0.454 / 0.380 with 2 M and 0.520 / 0.414 unbucketed); the bucket is
        # pushed when its LAST-visited (lowest) layer has its gradient.
        bucket_elems = int(os.environ.get("SINGA_B200_PUSH_BUCKET", str(6_000_000)))
        first_alone = os.environ.get("SINGA_B200_PUSH_FIRST_ALONE", "1") != "0"
        names = make_buckets([(l.name, *pm.layer_ranges[l.name]) for l in self.layers], bucket_elems,
                             first_alone=first_alone)
        by_name = {l.name: l for l in self.layers}
        self.buckets: List[List] = [[by_name[n] for n in b] for b in names]
        self.bucket_of = {l.name: b for b, ls in enumerate(self.buckets) for l in ls}
        self.trigger = {ls[-1].name for ls in self.buckets}          # lowest layer: visited last in backward
        self.lidx = dict(self.bucket_of)                             # flag / scratch index of a layer = its bucket
        nb = max(len(self.buckets), 1)
        # per bucket: push scratch (arrival counter + sequence number) and the consumer's wait sequence number;
        # the flag values are self-counted on the device, so nothing depends on a host-updated step scalar
        self.scratch = torch.zeros(nb, 2, dtype=torch.int64, device=ctx().device)
        self.wait_seq = torch.zeros(nb, dtype=torch.int64, device=ctx().device)
        self.lr_dev = step_state()["lr"]
        # Ownership is striped PER BUCKET: its chunk range is cut into `world` contiguous pieces and rank r serves
        # piece r, so every push is spread over all GPUs (the last layers of backward would otherwise all land on
        # the rank owning the head of the arena).
        self.bucket_mine: List[Tuple[int, int]] = []
        self.bucket_owners: List[List[int]] = []
        for ls in self.buckets:
            lo = min(pm.layer_ranges[l.name][0] for l in ls)
            n = sum(pm.layer_ranges[l.name][1] for l in ls)
            pieces = [(lo + a, b - a) for a, b in shard_bounds(n, self.world)]
            self.bucket_mine.append(pieces[self.rank])
            self.bucket_owners.append([r for r, (_, cnt) in enumerate(pieces) if cnt > 0])
        self.mine: Dict[str, Tuple[int, int]] = {ls[-1].name: self.bucket_mine[b] for b, ls in enumerate(self.buckets)}
        self._wait_arrays = {}
        for l in self.layers:
            i = self.lidx[l.name]
            arr = (ctypes.c_void_p * MAX_WAIT)()
            owners = self.bucket_owners[i]
            assert len(owners) <= MAX_WAIT
            addrs = []
            for k, o in enumerate(owners):
                arr[k] = heap.ptr("version", self.rank, (i * self.world + o) * 8)
                addrs.append(arr[k])
            self._wait_arrays[l.name] = (arr, len(owners))
            # parameter pull fused into the consumer GEMM: layers that support it get the flag addresses and
            # the sequence number and acquire them inside their forward GEMM's TMA producer warp
            if getattr(l, "supports_fused_pull", False):
                flag_list = torch.tensor([int(a) for a in addrs], dtype=torch.int64, device=ctx().device)
                l.pull_wait = (flag_list, self.wait_seq[i:i + 1].data_ptr())
        self.launches = 0
        self._used_comms = set()
        # Rotated schedule (Executor.rotate): the next forward pass is enqueued while this step's pushes still run, so
        # the consumer-side wait sequence must be advanced on the COMPUTE stream (pre_forward) — a forward GEMM that
        # read it before the push kernel had started would accept the previous version of its weights.
        self.main_bumps_seq = False
        # the fp32 master copy of a parameter is current only on its server; every other consumer (bias in the GEMM
        # epilogue, repacked conv weights) must read this rank's bf16 mirror
        if self.world > 1:
            for p in self.arena.params:
                p.use_mirror = True
                if p.pack16 is not None:
                    p.pack_from_mirror = True
            dist.barrier()

    # -- worker side ---------------------------------------------------------------------------------
    def update_layer(self, layer, step: int, side_stream=None) -> None:
        i = self.lidx[layer.name]
        main = torch.cuda.current_stream()
        # everything below runs on the comm stream, which waits for BOTH gradients of the layer (data
        # gradient on the compute stream, parameter gradient on its side stream): the compute stream goes
        # straight on to the layer below
        comm = self.comms[i % len(self.comms)]
        self._used_comms.add(i % len(self.comms))
        comm.wait_stream(main)
        if side_stream is not None:
            comm.wait_stream(side_stream)
        if layer.name not in self.trigger:
            return                 # the bucket is pushed by its lowest layer; the comm stream has this gradient's events
        first, n = self.bucket_mine[i]
        ready_own = self.heap.ptr("ready", self.rank, i * 8)
        seq_own = None if self.main_bumps_seq else self.wait_seq[i:i + 1].data_ptr()
        if n <= 0:                 # no piece of this bucket to serve: only announce the local gradients
            native.check(self.lib.sg_ps_flag_bump(ctypes.c_void_p(ready_own), ctypes.c_void_p(seq_own),
                                                  comm.cuda_stream), "sg_ps_flag_bump")
            native.count_launch()
            return
        # -- server side: my piece of this layer
        a = PsPush()
        a.world = self.world
        for r in range(self.world):
            a.grad[r] = self.heap.ptr("grad", r)
            a.w16[r] = self.heap.ptr("data16", r) if self.has_mirror else None
            a.ready[r] = self.heap.ptr("ready", r, i * 8)
            a.version[r] = self.heap.ptr("version", r, (i * self.world + self.rank) * 8)
        ar, u = self.arena, self.pm.updater
        a.w = ar.data.data_ptr()
        a.h = ar.history.data_ptr() if ar.history is not None else None
        a.u = ar.update.data_ptr() if ar.update is not None else None
        a.chunk_param, a.lr_mult, a.wd_mult = ar.chunk_param.data_ptr(), ar.lr_mult.data_ptr(), ar.wd_mult.data_ptr()
        a.first_chunk, a.nchunks = first, n
        a.up = CUpdater(_KIND_ID[u.kind], getattr(self.pm, "lr", 0.0), u.momentum, u.weight_decay, u.delta,
                        u.rho, 1.0 / self.world, self.lr_dev.data_ptr())
        # LR is read from device memory so a captured graph follows the schedule; safe on the comm stream because
        # end_step joins it into the compute stream before the next step's LR copy is enqueued
        a.step_dev, a.value_add = None, 0
        a.scratch = self.scratch[i].data_ptr()
        a.self_count = 1
        a.max_ctas = self.max_ctas
        a.mirror_fp32 = 1 if self.mirror_fp32 else 0
        a.trace = self.trace_buf.data_ptr() if getattr(self, "trace_buf", None) is not None else None
        if self.use_nvls:
            a.grad_mc = self.heap.mc_ptr("grad")
            a.w16_mc = self.heap.mc_ptr("data16") if self.has_mirror else None
        a.bump_ready, a.bump_seq = ready_own, seq_own     # the kernel announces this rank's gradients itself
        native.check(self.lib.sg_ps_push_update(ctypes.byref(a), comm.cuda_stream), "sg_ps_push_update")
        native.count_launch()
        self.launches += 1

    def wait_layer(self, layer, step: int) -> None:
        if layer.name not in self._wait_arrays or getattr(layer, "pull_wait", None) is not None:
            return   # no parameters, or the wait happens inside the layer's forward GEMM
        arr, n = self._wait_arrays[layer.name]
        i = self.lidx[layer.name]
        native.check(self.lib.sg_ps_flag_wait(arr, n, None, 0, ctypes.c_void_p(self.wait_seq[i:i + 1].data_ptr()),
                                              torch.cuda.current_stream().cuda_stream), "sg_ps_flag_wait")
        native.count_launch()
        for p in layer.get_params():
            if p.pack16 is not None:     # repacked operands (convolution weights) are re-derived from the mirror
                p.pack_from_mirror = True
                p.pack16()

    def use_rotated_schedule(self) -> None:
        """Called by the executor when it runs backward(t) + push(t) || forward(t+1) in one graph.  The consumer-side
        sequence numbers are then advanced on the compute stream, and the weights of a layer are awaited by a
        one-warp kernel in front of its forward GEMM instead of inside it: a persistent GEMM that spins for its weights
        holds a whole SM (200 KB of shared memory, 54 K registers in the tf32 build), and when its grid covers every
        SM the push kernel it waits for may find no SM to run on — measured as a deadlock (trap after the bounded
        spin) on the 16k-wide MLP.  A one-warp waiter leaves the SMs to the push."""
        self.main_bumps_seq = True
        for l in self.layers:
            if getattr(l, "pull_wait", None) is not None:
                l.pull_wait = None

    def pre_forward(self, step: int) -> None:
        assert self.main_bumps_seq, "pre_forward is part of the rotated schedule (the executor sets main_bumps_seq)"
        self.wait_seq.add_(1)

    def end_step(self, step: int) -> None:
        # only streams that took work this step: joining an idle stream during graph capture is an error
        # ("dependency created on uncaptured work in another stream")
        for k in sorted(self._used_comms):
            torch.cuda.current_stream().wait_stream(self.comms[k])
        self._used_comms.clear()

    def owned_ranges(self) -> List[Tuple[int, int]]:
        """Element ranges of the arena whose fp32 master copy (and optimizer state) lives on this rank."""
        return [(a * 1024, (a + n) * 1024) for a, n in self.bucket_mine if n > 0]

    def gather_master(self) -> None:
        """Collective: assemble the full fp32 master weights / history on every rank (checkpointing)."""
        gather_owned(self.arena, self.owned_ranges(), self.world)

    # -- reference API no-ops in synchronous mode -----------------------------------------------------------
    def put_all(self):
        pass

    def get_all(self):
        pass

    def sync_all(self, step):
        pass

    def close(self):
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        self.heap.close()
