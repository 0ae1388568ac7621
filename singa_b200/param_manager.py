"""ParamManager: the worker-side owner of the parameters — arena creation, initialisation, the updater, and
the exchange with the parameter servers.

Reference: include/utils/param_manager.h:38-127, src/utils/param_manager.cc:9-235 (Stack A), whose call
points are kept: ``update_param`` right after a layer's backward, ``wait_update`` right before a layer's next
forward (the layer-granular overlap hooks of src/worker/worker.cc:249-253,290-292).

Exchange backends:
  * ``local``  — no servers: the fused arena updater runs once per step.
  * ``zmq``    — Stack A over the ZeroMQ Router (kPut / kGet / kSync / kStop): Elastic or RandomSync every
                 ``sync_frequency`` steps; CPU plumbing config and cross-host control.
  * ``nccl``   — synchronous data parallel with NCCL reduce-scatter -> shard update -> all-gather
                 (the library baseline).
  * ``async``  — asynchronous worker groups with bounded staleness: local updates + EASGD / RandomSync kernels
                 on peer-mapped server shards, device-side progress counters (parallel/async_ps.py).
  * ``fused``  — the B200 product path: reduce-scatter + updater + broadcast in ONE kernel over NVSwitch peer
                 memory with device-side version counters (singa_b200/parallel/fused_ps.py).
"""
from __future__ import annotations

import time
from typing import Dict, List, Optional

import torch

from .arena import ParamArena
from .cluster import Cluster
from .device import ctx, set_step_state, step_state
from .param import Param, kGet, kPut, kStop, kSync
from .updater import create_updater


class ParamManager:
    def __init__(self, net, updater_proto, cluster: Optional[Cluster] = None, backend: Optional[str] = None,
                 seed: int = 0):
        self.net = net
        self.cluster = cluster
        self.updater_proto = updater_proto
        self.updater = create_updater(updater_proto)
        self.seed = seed
        ngroups = cluster.ngroups if cluster else 1
        nservers = cluster.nservers if cluster else 0
        self.moving_rate = updater_proto.moving_rate / max(ngroups, 1)
        self.sync_frequency = max(1, updater_proto.sync_frequency)
        self.warmup_steps = updater_proto.warmup_steps
        self.nservers = nservers
        self.sample_ratio = 1.0
        self.param_type = updater_proto.param_type
        self.hogwild = (cluster.nthreads_per_procs == 1 if cluster else True) or updater_proto.hogwild
        owners = [p for p in net.params if p.is_owner]
        if backend is None:
            backend = "local" if nservers == 0 else "zmq"
        import torch.distributed as dist
        world = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
        heap = None
        allocator = None
        if backend == "fused":
            # gradient arena, bf16 mirror and flags live in ONE peer-mapped allocation per rank
            from .parallel.fused_ps import symmetric_regions
            from .parallel.symm import make_heap
            probe = ParamArena.layout_total(owners, pad_chunks_to=world)
            nlayers = sum(1 for l in net.layers if l.get_params())
            # operand mirror: bf16 always; fp32 only when the master copies are spread over several ranks
            mirror_bytes = 2 if ctx().is_bf16 else (4 if world > 1 else 0)
            heap = make_heap(symmetric_regions(probe, nlayers, world, mirror_bytes),
                             dist.get_rank() if world > 1 else 0, world, ctx().device)

            def allocator(name, n, dt, heap=heap):
                if name in heap.offsets and name in ("grad", "data16"):
                    return heap.tensor(name, dt)
                return None
        if backend == "async":
            from .parallel.async_ps import async_regions
            from .parallel.symm import make_heap
            probe = ParamArena.layout_total(owners, pad_chunks_to=world)
            heap = make_heap(async_regions(probe, world), dist.get_rank() if world > 1 else 0, world, ctx().device)
        self.arena = ParamArena(owners, with_update=self.updater.needs_update,
                                with_snapshot=(self.param_type == "RandomSync" and nservers > 0),
                                allocator=allocator,
                                pad_chunks_to=world if backend in ("fused", "nccl", "async") else 1)
        for p in net.params:      # replicas of data-partitioned layers: re-alias value AND gradient to the owner
            if not p.is_owner:
                p.data, p.data16, p.grad = p.owner.data, p.owner.data16, p.owner.grad
        self.paramid2version: Dict[int, int] = {p.id: 0 for p in net.params}
        self.layer_ranges = {}
        for layer in net.layers:
            ps = layer.get_params()
            if ps:
                self.layer_ranges[layer.name] = self.arena.chunk_range(ps)
        self._param_layers = [l for l in net.layers if l.get_params()]
        self._first_param_layer = self._param_layers[0] if self._param_layers else None
        self.backend_name = backend
        self.backend = None
        if backend == "zmq":
            from .parallel.zmq_ps import ZmqExchange
            self.backend = ZmqExchange(self)
        elif backend == "nccl":
            from .parallel.nccl_ps import NcclExchange
            self.backend = NcclExchange(self)
        elif backend == "fused":
            from .parallel.fused_ps import FusedExchange
            self.backend = FusedExchange(self, heap)
        elif backend == "async":
            from .parallel.async_ps import AsyncExchange
            self.backend = AsyncExchange(self, heap)
        elif backend != "local":
            raise ValueError(f"unknown exchange backend {backend!r}")
        self.use_device_lr = ctx().is_native
        import os
        self.per_layer_update = os.environ.get("SINGA_B200_PER_LAYER_UPDATE") == "1"
        if os.environ.get("SINGA_B200_BIAS_MIRROR") == "1" and self.arena.data16 is not None:
            # numerics knob for equivalence tests: read biases from the bf16 mirror like the sharded exchanges do
            for p in self.arena.params:
                p.use_mirror = True

    # -- reference API -----------------------------------------------------------------------------
    def init_params(self) -> None:
        for p in self.arena.params:
            p.init(self.seed)
        if self.arena.snapshot is not None:
            self.arena.snapshot.copy_(self.arena.data)
        self.arena.sync_mirror()
        if self.backend is not None and hasattr(self.backend, "seed_center"):
            self.backend.seed_center()

    def sync_config(self, compute_time: float) -> None:
        """RandomSync sizing: ratio = bandwidth * nservers / (modelMB * nworkers / compute_time), <= 1
        (src/utils/param_manager.cc:88-96)."""
        if self.cluster is None or self.nservers == 0:
            return
        model_mb = self.arena.nbytes / (1024.0 * 1024.0)
        nworkers = max(1, self.cluster.nworkers)
        throughput = model_mb * nworkers / max(compute_time, 1e-9)
        self.sample_ratio = min(1.0, self.cluster.bandwidth * self.nservers / max(throughput, 1e-9))

    def sync_now(self, step: int) -> bool:
        return self.nservers > 0 and (step + 1) % self.sync_frequency == 0 and step > self.warmup_steps

    def send_params_to_servers(self) -> None:
        if self.backend is not None:
            self.backend.put_all()

    def get_params_from_servers(self) -> None:
        if self.backend is not None:
            self.backend.get_all()

    def begin_step(self, step: int, lr_step: Optional[int] = None) -> None:
        """Publish LR / step to device memory (keeps a captured graph valid across steps).  ``lr_step`` (rotated
        schedule): the graph about to run holds backward + update of ``lr_step`` and the forward pass of ``step``."""
        self.lr = self.updater.learning_rate(step if lr_step is None else lr_step)
        if self.use_device_lr:
            set_step_state(step, self.lr, step if lr_step is None else lr_step)
        if self.backend is not None and hasattr(self.backend, "begin_step"):
            self.backend.begin_step(step)

    def pre_forward(self, step: int) -> None:
        """Rotated schedule: called between backward(step-1) and forward(step), on the compute stream."""
        if self.backend is not None and hasattr(self.backend, "pre_forward"):
            self.backend.pre_forward(step)

    def update_param(self, layer, step: int, side_stream=None) -> None:
        """Called right after ``layer``'s backward.  ``side_stream`` (optional) is the stream the layer's
        parameter gradient was issued on; an exchange that pushes per layer waits on it (and on the compute
        stream) itself, so the compute stream is never blocked by the push."""
        if self.backend is not None and self.backend.handles_update:
            self.backend.update_layer(layer, step, side_stream)
            return
        # local update, per layer on the layer's side stream: the update of layer L then overlaps the backward of
        # the layers below it.  The side stream first waits for the compute stream, because L's data gradient
        # (already enqueued there) still reads the bf16 weights this update overwrites.
        # (Measured on the MNIST MLP: 0.375 ms/step with per-layer updates vs 0.307 ms with one update at the end — the
        # extra stream dependencies cost more than the overlap gains — so this is opt-in.)
        if self.per_layer_update and side_stream is not None and self.arena.use16 \
                and not getattr(self.net, "param_shares", None) and self.backend is None:
            main = torch.cuda.current_stream()
            side_stream.wait_stream(main)
            first, n = self.layer_ranges[layer.name]
            with torch.cuda.stream(side_stream):
                self._update_range(first, n, step)
            if layer is self._first_param_layer:
                main.wait_stream(side_stream)
            for p in layer.get_params():
                self.paramid2version[p.id] = step + 1
            return
        # otherwise one fused launch over the whole arena once the last gradient (the first layer in forward
        # order) is ready
        if layer is self._first_param_layer:
            shares = max(self.net.param_shares.values()) if getattr(self.net, "param_shares", None) else 1
            self._update_range(0, self.arena.nchunks, step, grad_scale=1.0 / shares)
            for p in self.net.params:
                self.paramid2version[p.id] = step + 1
            if self.backend is not None and self.sync_now(step):
                self.backend.sync_all(step)

    def _update_range(self, first_chunk: int, nchunks: int, step: int, grad_scale: float = 1.0) -> None:
        if self.arena.use16:
            from .ops import kernels as K
            from .updater import _KIND_ID
            u = self.updater
            K.update_arena(self.arena.data, self.arena.grad, self.arena.history, self.arena.update,
                           self.arena.data16, self.arena.chunk_param, self.arena.lr_mult, self.arena.wd_mult,
                           first_chunk, nchunks, kind=_KIND_ID[u.kind], lr=getattr(self, "lr", 0.0),
                           momentum=u.momentum, weight_decay=u.weight_decay, delta=u.delta, rho=u.rho,
                           grad_scale=grad_scale, lr_dev=step_state()["lr"] if self.use_device_lr else None)
            lo, hi = first_chunk * 1024, (first_chunk + nchunks) * 1024
            for p in self.arena.params:
                if p.pack16 is not None and lo <= self.arena.offsets[id(p)] < hi:
                    p.pack16()
        else:
            self.updater.update_arena(step, self.arena, first_chunk, nchunks, grad_scale)

    def end_step(self, step: int) -> None:
        """End of backward: join the exchange's side streams into the compute stream (required for CUDA-graph
        capture; free in synchronous mode because the next forward needs the last push anyway)."""
        if self.backend is not None and hasattr(self.backend, "end_step"):
            self.backend.end_step(step)

    def wait_update(self, layer, step: int) -> None:
        """Called right before ``layer``'s forward: block until its parameters are fresh enough."""
        if self.backend is not None:
            self.backend.wait_layer(layer, step)

    def finish(self) -> None:
        if self.backend is not None:
            self.backend.close()
            self.backend = None
