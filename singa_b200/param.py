"""Parameters: ``Param`` (value / gradient / history / update / snapshot blobs + metadata), the two
synchronisation flavours ``ElasticParam`` (EASGD) and ``RandomSyncParam`` (random-subset delta exchange), and
the message (de)serialisation used by the ZeroMQ parameter-server stacks.

Reference: include/utils/param.h:22-167, src/utils/param.cc:24-284.  Blobs are ``torch`` tensors; after
``ParamArena.adopt`` they are views into one flat device arena (the reference's ParamManager does the same
re-pointing on the host, src/utils/param_manager.cc:40-69).
"""
from __future__ import annotations

import math
import struct
import time
from typing import List, Optional, Sequence

import numpy as np
import torch

from .device import ctx
from .ops import reference as R
from .proto import schema as S
from .utils.factory import Factory

# message types on the wire (include/utils/param.h:12-21)
kGet, kPut, kSync, kUpdate, kSyncRequest, kSyncResponse, kStop, kData = range(8)

INIT = {v: k for k, v in dict(kConstant=0, kGaussain=1, kUniform=2, kPretrained=3, kGaussainSqrtFanIn=4,
                              kUniformSqrtFanIn=5, kUniformSqrtFanInOut=6).items()}


class Param:
    # static sync-cost counters, milliseconds (include/utils/param.h:153)
    ps_handle_sync = 0.0
    worker_gen_sync = 0.0
    worker_handle_sync = 0.0

    def __init__(self):
        self.proto = S.ParamProto()
        self.name = ""
        self.id = -1
        self.shape: List[int] = []
        self.fan_in = 0
        self.owner: Optional["Param"] = None     # set by share_data
        self.data: Optional[torch.Tensor] = None
        self.grad: Optional[torch.Tensor] = None
        self.history: Optional[torch.Tensor] = None
        self.update: Optional[torch.Tensor] = None
        self.snapshot: Optional[torch.Tensor] = None
        self.data16: Optional[torch.Tensor] = None   # bf16 operand for the tcgen05 GEMMs (2-D, ld % 8 == 0)
        self.pack16 = None                           # callable refreshing data16 when it is not an arena view
        self.version = 0
        self.arena_offset = -1

    # -- setup / init ----------------------------------------------------------------------------
    def setup(self, proto, shape: Sequence[int], fan_in: int) -> None:
        self.proto.CopyFrom(proto)
        self.name = proto.name
        self.shape = [int(s) for s in shape]
        self.fan_in = int(fan_in)
        dev = ctx().device
        self.data = torch.zeros(self.shape, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(self.shape, dtype=torch.float32, device=dev)
        self.history = None
        self.update = None

    @property
    def size(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    @property
    def learning_rate_multiplier(self) -> float:
        return self.proto.learning_rate_multiplier

    @property
    def weight_decay_multiplier(self) -> float:
        return self.proto.weight_decay_multiplier

    @property
    def is_owner(self) -> bool:
        return self.owner is None

    def share_data(self, other: "Param") -> None:
        """Alias the value blob of ``other`` (test/validation nets share the train net's weights;
        include/utils/param.h:92-97)."""
        self.owner = other if other.owner is None else other.owner
        self.data = self.owner.data
        self.data16 = self.owner.data16

    def init(self, seed: int = 0) -> None:
        """Fill ``data`` per ``init_method`` (src/utils/param.cc:89-127).  Distributions match the reference;
        the stream is a seeded counter RNG (Philox on the GPU, torch.Generator on the CPU)."""
        p = self.proto
        method = INIT[p.init_method]
        v = p.value
        # value == 0 skips the WHOLE multiplier, fan-in factors included (src/utils/param.cc:97-123)
        skip = v == 0
        if method == "kPretrained":
            return
        if method == "kConstant":
            self.data.fill_(v)
            return
        s0 = self.shape[0]
        s1 = self.shape[1] if len(self.shape) > 1 else 0
        if method == "kUniform":
            kind, a, b, scale = 0, p.low, p.high, v
        elif method == "kUniformSqrtFanIn":
            kind, a, b, scale = 0, p.low, p.high, v / math.sqrt(max(self.fan_in, 1) / 3.0)
        elif method == "kUniformSqrtFanInOut":
            kind, a, b, scale = 0, p.low, p.high, v / math.sqrt(s0 + s1)
        elif method == "kGaussain":
            kind, a, b, scale = 1, p.mean, p.std, v
        elif method == "kGaussainSqrtFanIn":
            kind, a, b, scale = 1, p.mean, p.std, v / math.sqrt(s0)
        else:
            raise ValueError(f"unknown init method {p.init_method}")
        if skip:
            scale = 1.0
        if self.data.is_cuda and ctx().is_native:
            from .ops import kernels as K
            sd = seed * 1000003 + (self.id if self.id >= 0 else 0) + 12345
            if self.data.is_contiguous():
                K.fill_random(self.data.view(-1), kind, a, b, scale, sd)
            else:   # padded row stride in the arena: fill a dense temporary, copy the logical part
                tmp = torch.empty(self.shape, dtype=torch.float32, device=self.data.device)
                K.fill_random(tmp.view(-1), kind, a, b, scale, sd)
                self.data.copy_(tmp)
        else:
            g = torch.Generator(device="cpu")
            g.manual_seed(seed * 1000003 + (self.id if self.id >= 0 else 0) + 12345)
            if kind == 0:
                t = torch.rand(self.shape, generator=g) * (b - a) + a
            else:
                t = torch.randn(self.shape, generator=g) * b + a
            self.data.copy_((t * scale).to(self.data.device))

    def refresh16(self) -> None:
        """Re-derive the bf16 GEMM operand after ``data`` changed outside the fused updater."""
        if self.owner is not None:
            return self.owner.refresh16()
        if self.pack16 is not None:
            self.pack16()

    # -- (de)serialisation for the ZeroMQ stacks -----------------------------------------------
    def to_bytes(self) -> bytes:
        return self.data.detach().to("cpu", torch.float32).contiguous().numpy().tobytes()

    def from_bytes(self, payload) -> None:
        """``payload``: any buffer of ``size`` float32 values (bytes, memoryview, zmq.Frame.buffer).  One copy."""
        arr = np.frombuffer(payload, dtype=np.float32)
        if arr.size != self.size:
            raise ValueError(f"param {self.name}: payload has {arr.size} floats, expected {self.size}")
        if self.data.device.type == "cpu" and self.data.is_contiguous() and self.data.dtype == torch.float32:
            np.copyto(self.data.numpy().reshape(-1), arr)          # straight into the blob, no temporaries
        else:
            self.data.copy_(torch.from_numpy(arr.copy()).view(self.shape).to(self.data.device))

    def to_buffer(self):
        """The value as a buffer for a zero-copy send: a view of the blob itself on the CPU (the caller must send it
        before the next update of this parameter), a host copy otherwise."""
        if self.data.device.type == "cpu" and self.data.is_contiguous() and self.data.dtype == torch.float32:
            return memoryview(self.data.detach().numpy().reshape(-1)).cast("B")
        return self.to_bytes()

    # server side: kPut stores the payload, kGet returns the value, kUpdate replaces the value and echoes it
    def handle_put_msg(self, payload, shape: Optional[Sequence[int]] = None) -> None:
        n = memoryview(payload).nbytes // 4
        if self.data is None or self.size != n:
            self.shape = list(shape) if shape else [n]
            self.data = torch.zeros(self.shape, dtype=torch.float32, device="cpu")
        self.from_bytes(payload)

    def handle_get_msg(self):
        return self.to_bytes()

    def handle_update_msg(self, payload):
        """Stack-B placeholder semantics: replace and echo (src/utils/param.cc:57-61).  The echo IS the received
        buffer — identical bytes, so the reply needs no second copy of the tensor."""
        self.from_bytes(payload)
        self.version += 1
        return payload

    # sync protocol (overridden)
    def handle_sync_msg(self, control: str, payload: bytes):
        raise NotImplementedError

    def gen_sync_msg_from_worker(self, arg: float):
        raise NotImplementedError

    def parse_sync_msg_from_ps(self, control: str, payload: bytes) -> None:
        raise NotImplementedError


class ElasticParam(Param):
    """EASGD: worker sends w; server d = (w - s) alpha, s += d, replies d; worker w -= d."""

    def handle_sync_msg(self, control: str, payload: bytes):
        t0 = time.monotonic()
        alpha_s, count_s = control.rsplit("-", 1)
        alpha, count = float(alpha_s), int(count_s)
        assert count == self.size, (count, self.size)
        worker = torch.from_numpy(np.frombuffer(payload, dtype=np.float32).copy()).view(self.shape)
        server = self.data.to("cpu")
        d = (worker - server) * alpha
        self.data.add_(d.to(self.data.device))
        Param.ps_handle_sync += (time.monotonic() - t0) * 1e3
        return control, d.numpy().tobytes()

    def gen_sync_msg_from_worker(self, alpha: float):
        t0 = time.monotonic()
        out = (f"{alpha:f}-{self.size}", self.to_bytes())
        Param.worker_gen_sync += (time.monotonic() - t0) * 1e3
        return out

    def parse_sync_msg_from_ps(self, control: str, payload: bytes) -> None:
        t0 = time.monotonic()
        diff = torch.from_numpy(np.frombuffer(payload, dtype=np.float32).copy()).view(self.shape)
        self.data.sub_(diff.to(self.data.device))
        self.refresh16()
        Param.worker_handle_sync += (time.monotonic() - t0) * 1e3


class RandomSyncParam(Param):
    """Random-subset delta exchange: worker sends w - snap on a shared-seed subset, server adds it and
    replies its old values, worker folds them in and re-snapshots."""

    def setup(self, proto, shape, fan_in):
        super().setup(proto, shape, fan_in)
        self.snapshot = torch.zeros(self.shape, dtype=torch.float32, device=self.data.device)

    def init(self, seed: int = 0) -> None:
        super().init(seed)
        if self.snapshot is not None:
            self.snapshot.copy_(self.data)

    @staticmethod
    def random_sample(seed: int, ratio: float, n: int) -> torch.Tensor:
        if ratio >= 1.0:
            return torch.arange(n)
        return torch.nonzero(R.philox_uniform(seed, n) < ratio).squeeze(1)

    def handle_sync_msg(self, control: str, payload: bytes):
        t0 = time.monotonic()
        seed_s, ratio_s, count_s = control.split("-")
        seed, ratio, count = int(seed_s), float(ratio_s), int(count_s)
        idx = self.random_sample(seed, ratio, self.size)
        assert idx.numel() == count, (idx.numel(), count)
        delta = torch.from_numpy(np.frombuffer(payload, dtype=np.float32).copy())
        flat = self.data.view(-1)
        old = flat[idx.to(flat.device)].to("cpu").clone()
        flat[idx.to(flat.device)] += delta.to(flat.device)
        Param.ps_handle_sync += (time.monotonic() - t0) * 1e3
        return control, old.numpy().tobytes()

    def gen_sync_msg_from_worker(self, sample_ratio: float):
        t0 = time.monotonic()
        seed = time.time_ns() & 0x7FFFFFFF
        idx = self.random_sample(seed, sample_ratio, self.size)
        w, sn = self.data.view(-1), self.snapshot.view(-1)
        delta = (w[idx.to(w.device)] - sn[idx.to(w.device)]).to("cpu")
        Param.worker_gen_sync += (time.monotonic() - t0) * 1e3
        return f"{seed}-{sample_ratio:.9g}-{idx.numel()}", delta.numpy().tobytes()

    def parse_sync_msg_from_ps(self, control: str, payload: bytes) -> None:
        t0 = time.monotonic()
        seed_s, ratio_s, count_s = control.split("-")
        idx = self.random_sample(int(seed_s), float(ratio_s), self.size)
        old = torch.from_numpy(np.frombuffer(payload, dtype=np.float32).copy())
        w, sn = self.data.view(-1), self.snapshot.view(-1)
        di = idx.to(w.device)
        w[di] += old.to(w.device) - sn[di]
        sn[di] = w[di]
        self.refresh16()
        Param.worker_handle_sync += (time.monotonic() - t0) * 1e3


param_factory: Factory[Param] = Factory("Param")


def register_param_type(param_type: str) -> None:
    """``NeuralNet::RegistryParam``: bind the factory id "Param" to the configured sync flavour
    (src/worker/neuralnet.cc:35-44)."""
    if param_type == "Elastic":
        param_factory.register("Param", ElasticParam)
    elif param_type == "RandomSync":
        param_factory.register("Param", RandomSyncParam)
    elif param_type in ("", "Plain", "Param"):
        param_factory.register("Param", Param)
    else:
        raise ValueError(f"unknown param_type {param_type!r}")
    for k, cls in (("Elastic", ElasticParam), ("RandomSync", RandomSyncParam), ("Plain", Param)):
        param_factory.register(k, cls)


register_param_type("Elastic")
