"""Peer-mapped ("symmetric") device buffers: every rank allocates the same-sized buffer with cudaMalloc,
exports it through CUDA IPC, and maps all peers' buffers, so kernels can ld/st any rank's copy over
NVLink/NVSwitch.  The 64-byte IPC handles travel over the control plane (``torch.distributed`` object
all-gather, or any callable the caller supplies — e.g. the ZeroMQ router).

This replaces the reference's TCP ROUTER/DEALER data plane and is the substrate of the fused
push/pull kernels in ``csrc/ps_fused.cu``.
"""
from __future__ import annotations

import ctypes
from typing import Callable, List, Optional

import torch
import torch.distributed as dist

from ..ops import native


class _RawCuda:
    """Expose a raw device pointer to torch through the CUDA array interface."""

    def __init__(self, ptr: int, nbytes: int, keepalive=None):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2}
        self._keepalive = keepalive


def wrap_pointer(ptr: int, nbytes: int, dtype: torch.dtype, device, keepalive=None) -> torch.Tensor:
    t = torch.as_tensor(_RawCuda(ptr, nbytes, keepalive), device=device)
    return t.view(dtype)


class SymmetricBuffer:
    def __init__(self, nbytes: int, rank: int, world: int, device,
                 all_gather: Optional[Callable[[bytes], List[bytes]]] = None):
        lib = native.require()
        self.lib = lib
        self.nbytes, self.rank, self.world, self.device = nbytes, rank, world, device
        p = ctypes.c_void_p()
        native.check(lib.sg_ipc_alloc(nbytes, ctypes.byref(p)), "sg_ipc_alloc")
        self.local_ptr = p.value
        self.peer_ptrs: List[int] = [0] * world
        self.peer_ptrs[rank] = self.local_ptr
        self._opened: List[int] = []
        if world > 1:
            h = ctypes.create_string_buffer(64)
            native.check(lib.sg_ipc_get_handle(ctypes.c_void_p(self.local_ptr), h), "sg_ipc_get_handle")
            mine = bytes(h.raw)
            if all_gather is None:
                gathered: List[Optional[bytes]] = [None] * world
                dist.all_gather_object(gathered, mine)
            else:
                gathered = all_gather(mine)
            for r, hb in enumerate(gathered):
                if r == rank:
                    continue
                q = ctypes.c_void_p()
                native.check(lib.sg_ipc_open_handle(ctypes.create_string_buffer(hb, 64), ctypes.byref(q)),
                             f"sg_ipc_open_handle(rank {r})")
                self.peer_ptrs[r] = q.value
                self._opened.append(q.value)

    def tensor(self, dtype: torch.dtype, offset_bytes: int = 0, nbytes: Optional[int] = None,
               rank: Optional[int] = None) -> torch.Tensor:
        r = self.rank if rank is None else rank
        nb = self.nbytes - offset_bytes if nbytes is None else nbytes
        return wrap_pointer(self.peer_ptrs[r] + offset_bytes, nb, dtype, self.device, keepalive=self)

    def close(self) -> None:
        for q in self._opened:
            self.lib.sg_ipc_close_handle(ctypes.c_void_p(q))
        self._opened = []
        if self.local_ptr:
            self.lib.sg_ipc_free(ctypes.c_void_p(self.local_ptr))
            self.local_ptr = 0


class SymmetricHeap:
    """One symmetric allocation carved into named, 256-byte aligned regions (gradient arena, bf16 mirror,
    flag block ...), so a single IPC handle per rank is exchanged."""

    def __init__(self, regions, rank: int, world: int, device, all_gather=None):
        self.offsets = {}
        off = 0
        for name, nbytes in regions:
            self.offsets[name] = (off, nbytes)
            off += (nbytes + 255) // 256 * 256
        self.buf = SymmetricBuffer(max(off, 256), rank, world, device, all_gather)
        self.rank, self.world = rank, world

    def tensor(self, name: str, dtype: torch.dtype, rank: Optional[int] = None) -> torch.Tensor:
        off, nbytes = self.offsets[name]
        return self.buf.tensor(dtype, off, nbytes, rank)

    def ptr(self, name: str, rank: int, byte_offset: int = 0) -> int:
        return self.buf.peer_ptrs[rank] + self.offsets[name][0] + byte_offset

    def close(self) -> None:
        self.buf.close()


class TorchSymmetricHeap:
    """Same interface as :class:`SymmetricHeap`, backed by ``torch.distributed._symmetric_memory`` — the
    allocation / handle exchange plumbing of PyTorch — which additionally binds the buffer to an NVLS multicast
    object when the fabric supports it.  ``mc_ptr(name)`` is the multicast address of a region (0 when
    unavailable); kernels use it with ``multimem.ld_reduce`` / ``multimem.st``."""

    def __init__(self, regions, rank: int, world: int, device):
        import torch.distributed._symmetric_memory as symm_mem
        self.offsets = {}
        off = 0
        for name, nbytes in regions:
            self.offsets[name] = (off, nbytes)
            off += (nbytes + 255) // 256 * 256
        self.rank, self.world, self.device = rank, world, device
        self.t = symm_mem.empty(max(off, 256), dtype=torch.uint8, device=device)
        self.t.zero_()
        self.hdl = symm_mem.rendezvous(self.t, dist.group.WORLD)
        self.peer_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        mc = 0
        try:
            if self.hdl.has_multicast_support(device.type if hasattr(device, "type") else "cuda", device.index or 0) \
                    if callable(getattr(self.hdl, "has_multicast_support", None)) else True:
                mc = int(self.hdl.multicast_ptr)
        except Exception:  # noqa: BLE001
            try:
                mc = int(self.hdl.multicast_ptr)
            except Exception:  # noqa: BLE001
                mc = 0
        self.multicast_base = mc
        torch.cuda.synchronize()
        dist.barrier()

    def tensor(self, name: str, dtype: torch.dtype, rank: Optional[int] = None) -> torch.Tensor:
        off, nbytes = self.offsets[name]
        r = self.rank if rank is None else rank
        if r == self.rank:
            return self.t[off:off + nbytes].view(dtype)
        return wrap_pointer(self.peer_ptrs[r] + off, nbytes, dtype, self.device, keepalive=self)

    def ptr(self, name: str, rank: int, byte_offset: int = 0) -> int:
        return self.peer_ptrs[rank] + self.offsets[name][0] + byte_offset

    def mc_ptr(self, name: str) -> int:
        return self.multicast_base + self.offsets[name][0] if self.multicast_base else 0

    def close(self) -> None:
        self.hdl = None
        self.t = None


def make_heap(regions, rank: int, world: int, device):
    """Prefer torch symmetric memory (NVLS multicast) for multi-rank jobs; CUDA-IPC heap otherwise."""
    import os
    kind = os.environ.get("SINGA_B200_HEAP", "symm" if world > 1 else "ipc")
    if kind == "symm" and world > 1:
        try:
            return TorchSymmetricHeap(regions, rank, world, device)
        except Exception as e:  # noqa: BLE001
            import logging
            logging.getLogger("singa").error("torch symmetric memory unavailable (%s); using the CUDA-IPC heap", e)
    return SymmetricHeap(regions, rank, world, device)
