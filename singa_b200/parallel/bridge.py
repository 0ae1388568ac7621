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


class PeerBridgeTransport(BridgeTransport):
    """Bridges over peer memory (messages M7/M8 without sockets or NCCL): the receiving end's buffer — a BridgeDst's
    ``data`` in the forward pass, a BridgeSrc's ``grad`` in the backward pass — lives in a symmetric (peer-mapped)
    heap.  The sending rank copies into it with a plain device copy through the peer mapping and then release-stores
    a flag next to it; the receiving rank's stream acquire-waits for the flag (``sg_ps_flag_set`` /
    ``sg_ps_flag_wait`` with the device-resident step counter as the flag value).  Nothing touches the host, so a
    step with bridges can be captured in a CUDA graph like any other.

    Reference: src/worker/worker.cc:216-261,293-299 (blocking PUSH/PULL per bridge layer).
    ``rank_of(layer)`` maps a layer to the rank that executes it (default: its location id)."""

    capturable = True

    def __init__(self, net, rank: int, world: int, device, rank_of=None):
        from ..device import step_state
        from ..ops import native
        from .symm import make_heap
        self.cluster = None
        self.distributed = world > 1
        self.rank, self.world, self.device = rank, world, device
        self._rank_of_fn = rank_of or (lambda layer: layer.locationid % world)
        self.lib = native.require()
        self.native = native
        self.step_dev = step_state()["step"]
        self.edges = []          # (src layer, dst layer) pairs that cross ranks, in net order (same on every rank)
        regions = []
        for l in net.layers:
            if l.is_bridgesrclayer:
                d = l.dstlayers[0]
                if self._rank_of(l) != self._rank_of(d):
                    e = len(self.edges)
                    self.edges.append((l, d))
                    regions += [(f"data{e}", self._nbytes(d.data)), (f"grad{e}", self._nbytes(l.grad)),
                                (f"flags{e}", 16)]
        self.heap = make_heap(regions, rank, world, device) if regions else None
        self._edge_of = {}
        for e, (s, d) in enumerate(self.edges):
            rs, rd = self._rank_of(s), self._rank_of(d)
            # the consumer-side buffers move into the heap: on the owning rank the layer's tensor IS the heap region,
            # on the sending rank `remote` is the peer view it copies into
            d.data = self._shaped(self.heap.tensor(f"data{e}", d.data.dtype, rd), d.data)
            s.grad = self._shaped(self.heap.tensor(f"grad{e}", s.grad.dtype, rs), s.grad)
            info = {"e": e, "rs": rs, "rd": rd,
                    "fwd_flag": self.heap.ptr(f"flags{e}", rd, 0), "bwd_flag": self.heap.ptr(f"flags{e}", rs, 8)}
            self._edge_of[id(s)] = info
            self._edge_of[id(d)] = info

    @staticmethod
    def _nbytes(t: torch.Tensor) -> int:
        # span of the (possibly row-padded) activation in elements, in bytes
        span = 1 + sum((n - 1) * st for n, st in zip(t.shape, t.stride()))
        return span * t.element_size()

    @staticmethod
    def _shaped(flat: torch.Tensor, like: torch.Tensor) -> torch.Tensor:
        return flat.as_strided(like.shape, like.stride(), flat.storage_offset())

    def _rank_of(self, layer) -> int:
        return self._rank_of_fn(layer)

    def is_local(self, layer) -> bool:
        return self._rank_of(layer) == self.rank

    def _set(self, flag_ptr: int) -> None:
        import ctypes
        self.native.check(self.lib.sg_ps_flag_set(ctypes.c_void_p(flag_ptr), ctypes.c_void_p(self.step_dev.data_ptr()),
                                                  1, torch.cuda.current_stream().cuda_stream), "sg_ps_flag_set(bridge)")
        self.native.count_launch()

    def _wait(self, flag_ptr: int) -> None:
        import ctypes
        arr = (ctypes.c_void_p * 32)()
        arr[0] = flag_ptr
        self.native.check(self.lib.sg_ps_flag_wait(arr, 1, ctypes.c_void_p(self.step_dev.data_ptr()), 1, None,
                                                   torch.cuda.current_stream().cuda_stream), "sg_ps_flag_wait(bridge)")
        self.native.count_launch()

    def send_data(self, src_layer) -> None:
        dst_layer = src_layer.dstlayers[0]
        info = self._edge_of.get(id(src_layer))
        if info is None:
            dst_layer.data.copy_(src_layer.data)          # both ends on this rank
            return
        dst_layer.data.copy_(src_layer.data)              # dst_layer.data is the peer view of the consumer's buffer
        self._set(info["fwd_flag"])

    def recv_data(self, dst_layer) -> None:
        info = self._edge_of.get(id(dst_layer))
        if info is not None:
            self._wait(info["fwd_flag"])

    def send_grad(self, dst_layer) -> None:
        src_layer = dst_layer.srclayers[0]
        info = self._edge_of.get(id(dst_layer))
        src_layer.grad.copy_(dst_layer.grad)
        if info is not None:
            self._set(info["bwd_flag"])

    def recv_grad(self, src_layer) -> None:
        info = self._edge_of.get(id(src_layer))
        if info is not None:
            self._wait(info["bwd_flag"])

    def close(self) -> None:
        if self.heap is not None:
            self.heap.close()
