"""Flat parameter arena: one contiguous fp32 buffer each for value / gradient / history (/ update /
snapshot) of ALL parameters of a net plus an operand MIRROR of the values, with every parameter starting on a
1024-element boundary.

The mirror (``data16`` / ``Param.mirror16``; the names date from the bf16-only version) is what the GEMMs read:
  * bf16 compute: a bf16 copy of the values, refreshed by the updater in the same pass;
  * fp32 (tf32) compute: no copy at all on one GPU — the GEMMs read the fp32 master values directly — and an fp32
    copy only when a sharded parameter server keeps the master elsewhere and broadcasts fresh values into it.

Why: (1) one fused multi-tensor updater launch per step (or per layer) instead of one per parameter;
(2) a "server shard" of the B200 parameter server is simply a chunk range of this arena on some GPU, so the
fused push/pull kernels address it with plain offsets through peer memory; (3) the bf16 mirror is written by
the updater in the same pass, so the GEMMs never need a separate cast kernel.
Reference counterpart: ParamManager's single flat Param buffer, src/utils/param_manager.cc:40-69.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch

from .device import ctx
from .param import Param

CHUNK = 1024


def _round_up(n: int, m: int) -> int:
    return (n + m - 1) // m * m


class ParamArena:
    def __init__(self, params: Sequence[Param], *, with_update: bool = False, with_snapshot: bool = False,
                 with_history: bool = True, device=None, allocator=None, pad_chunks_to: int = 1):
        self.device = device or ctx().device
        self.params: List[Param] = [p for p in params if p.is_owner]
        self.offsets: Dict[int, int] = {}
        self.slot: Dict[int, int] = {}      # elements occupied in the arena (>= p.size when rows are padded)
        self.use16 = self.device.type == "cuda" and ctx().is_native      # native kernels own the arena
        self.mirror_dtype = ctx().act_dtype
        off = 0
        for p in self.params:
            self.offsets[id(p)] = off
            # a 2-D GEMM weight whose row length is not a multiple of 16 bytes is stored with a padded row stride so
            # that the operand is directly TMA-legal (padding stays zero: its gradient is never written)
            p.ld = p.shape[-1] if p.shape else 1
            if self.use16 and getattr(p, "want_padded_ld", False) and len(p.shape) == 2 and p.shape[1] % 8:
                p.ld = (p.shape[1] + 7) // 8 * 8
                self.slot[id(p)] = p.shape[0] * p.ld
            else:
                self.slot[id(p)] = p.size
            # every parameter occupies a multiple of `pad_chunks_to` chunks (= the number of server shards), so any run of
            # adjacent layers splits into equal per-rank pieces: in-place NCCL reduce-scatter / all-gather and the fused
            # push both address piece r as base + r * piece
            off += _round_up(self.slot[id(p)], CHUNK * max(1, pad_chunks_to))
        self.total = _round_up(max(off, CHUNK), CHUNK * max(1, pad_chunks_to))
        self.nchunks = self.total // CHUNK
        self.pad_chunks_to = max(1, pad_chunks_to)

        def default_alloc(name, n, dt):
            return torch.zeros(n, dtype=dt, device=self.device)
        # allocator(name, nelems, dtype): lets the fused parameter server place `grad` / `data16` in
        # peer-mapped (CUDA IPC) memory; may return None to fall back to a plain torch allocation
        def make(name, dt):
            t = allocator(name, self.total, dt) if allocator is not None else None
            return t if t is not None else default_alloc(name, self.total, dt)
        self.data = make("data", torch.float32)
        self.grad = make("grad", torch.float32)
        self.history = make("history", torch.float32) if with_history else None
        self.update = make("update", torch.float32) if with_update else None
        self.snapshot = make("snapshot", torch.float32) if with_snapshot else None
        if not self.use16:
            self.data16 = None
        elif self.mirror_dtype == torch.bfloat16:
            self.data16 = make("data16", torch.bfloat16)
        else:   # fp32 compute: a mirror exists only if the exchange backend provides one (sharded master copies)
            self.data16 = allocator("data16", self.total, torch.float32) if allocator is not None else None
        chunk_param = torch.full((self.nchunks,), -1, dtype=torch.int32)
        lr_mult = torch.ones(max(len(self.params), 1), dtype=torch.float32)
        wd_mult = torch.ones(max(len(self.params), 1), dtype=torch.float32)
        for i, p in enumerate(self.params):
            o = self.offsets[id(p)]
            chunk_param[o // CHUNK:(o + _round_up(self.slot[id(p)], CHUNK)) // CHUNK] = i
            lr_mult[i] = p.learning_rate_multiplier
            wd_mult[i] = p.weight_decay_multiplier
        self.chunk_param = chunk_param.to(self.device)
        self.lr_mult = lr_mult.to(self.device)
        self.wd_mult = wd_mult.to(self.device)
        self.adopt()

    @staticmethod
    def layout_total(params: Sequence[Param], pad_chunks_to: int = 1) -> int:
        """Number of arena elements the given owner parameters will occupy (same rules as __init__)."""
        use16 = ctx().device.type == "cuda" and ctx().is_native
        off = 0
        for p in params:
            if not p.is_owner:
                continue
            n = p.size
            if use16 and getattr(p, "want_padded_ld", False) and len(p.shape) == 2 and p.shape[1] % 8:
                n = p.shape[0] * ((p.shape[1] + 7) // 8 * 8)
            off += _round_up(n, CHUNK * max(1, pad_chunks_to))
        return _round_up(max(off, CHUNK), CHUNK * max(1, pad_chunks_to))

    def adopt(self, all_params: Sequence[Param] = ()) -> None:
        """Re-point every owner parameter's blobs into the arena (existing values are preserved); parameters that
        share an owner (test nets, replicas of a data-partitioned layer) are re-aliased afterwards."""
        for p in self.params:
            o, n = self.offsets[id(p)], self.slot[id(p)]
            p.arena_offset = o
            padded = n != p.size

            def shaped(buf):
                if padded:
                    return buf[o:o + n].view(p.shape[0], p.ld)[:, :p.shape[1]]
                return buf[o:o + n].view(p.shape)

            def view(buf, old):
                v = shaped(buf)
                if old is not None and old.data_ptr() != v.data_ptr():
                    v.copy_(old.to(v.device))
                return v

            p.data = view(self.data, p.data)
            p.grad = view(self.grad, p.grad)
            if self.history is not None:
                p.history = view(self.history, p.history)
            if self.update is not None:
                p.update = view(self.update, p.update)
            if self.snapshot is not None:
                p.snapshot = view(self.snapshot, p.snapshot)
            if self.data16 is not None:
                p.mirror16 = shaped(self.data16)     # always: the parameter server broadcasts into the mirror
                if p.pack16 is None:
                    p.data16 = p.mirror16            # GEMM operand = the arena's mirror
            elif self.use16:
                p.mirror16 = None
                if p.pack16 is None:
                    p.data16 = p.data                # fp32 compute on one GPU: the GEMMs read the master values
        for q in all_params:
            if not q.is_owner:
                shared_grad = q.grad is not None and q.owner.grad is not None and getattr(q, "_shares_grad", False)
                q.data, q.data16 = q.owner.data, q.owner.data16
                if shared_grad:
                    q.grad = q.owner.grad
        self.sync_mirror()

    def view16(self, p: Param) -> torch.Tensor:
        return (p if p.is_owner else p.owner).data16

    def sync_mirror(self) -> None:
        """data16 <- data for the whole arena, then refresh every packed operand."""
        if self.data16 is not None:
            if self.data16.dtype == self.data.dtype:
                self.data16.copy_(self.data)
            else:
                from .ops import kernels as K
                K.convert(self.data, self.data16)
        for p in self.params:
            if p.pack16 is not None:
                p.pack16()

    def chunk_range(self, params: Sequence[Param]) -> Tuple[int, int]:
        """(first_chunk, nchunks) covering ``params`` (must be adjacent in the arena, e.g. one layer)."""
        owners = [p if p.is_owner else p.owner for p in params]
        owners = [p for p in owners if id(p) in self.offsets]
        if not owners:
            return 0, 0
        lo = min(self.offsets[id(p)] for p in owners)
        hi = max(self.offsets[id(p)] + _round_up(self.slot[id(p)], CHUNK * self.pad_chunks_to) for p in owners)
        return lo // CHUNK, (hi - lo) // CHUNK

    def param_index(self, p: Param) -> int:
        return self.params.index(p if p.is_owner else p.owner)

    @property
    def nbytes(self) -> int:
        return self.total * 4
