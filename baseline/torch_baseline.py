"""Library baseline for bench.py (``--impl torch``): the five benchmark nets written the way a PyTorch user would
write them — torch.nn modules (cuBLAS / cuDNN kernels, channels-last convolutions), autograd, fused-foreach SGD,
gradients all-reduced through NCCL in buckets on a side stream while backward runs, and the whole step (input parse,
forward, backward, all-reduce, update) captured in ONE CUDA graph.  None of this repository's kernels, layers or engine
is on this path; only the synthetic records (``singa_b200.layers.data.SyntheticSource``) are shared so that both arms
read identical uint8 batches.

Precision: ``fp32`` = fp32 storage with TF32 tensor-core products allowed (the equal-precision arm for the product's
tf32 path), ``fp32-strict`` = TF32 off, ``bf16`` = torch.autocast(bfloat16) over fp32 master weights.

Architectures, initial distributions, LR / momentum / weight decay follow singa_b200/models/configs.py (which restates
the reference's example configs); the scaled tanh is the reference's 1.7159047 tanh(0.66666667 x).
"""
from __future__ import annotations

import math
import os
import time
from typing import List

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class STanh(nn.Module):
    def forward(self, x):
        return 1.7159047 * torch.tanh(0.66666667 * x)


def _uniform(m: nn.Module, low, high):
    with torch.no_grad():
        m.weight.uniform_(low, high)
        if m.bias is not None:
            m.bias.uniform_(low, high)
    return m


def _gauss(m: nn.Module, std, bias=0.0):
    with torch.no_grad():
        m.weight.normal_(0.0, std)
        if m.bias is not None:
            m.bias.fill_(bias)
    return m


def _lrn(size=3, alpha=5e-5, beta=0.75):
    return nn.LocalResponseNorm(size, alpha=alpha, beta=beta, k=1.0)


def build_model(config: str, classes_override=None):
    """-> (module, input shape (C,H,W) or (H,W), crop, scale, norm_a, norm_b, lr, momentum, wd, classes, batch)."""
    if config == "mnist_mlp":
        dims = [784, 2500, 2000, 1500, 1000, 500]
        layers: List[nn.Module] = [nn.Flatten()]
        for a, b in zip(dims, dims[1:]):
            layers += [_uniform(nn.Linear(a, b), -0.05, 0.05), STanh()]
        layers.append(_uniform(nn.Linear(500, 10), -0.05, 0.05))
        return nn.Sequential(*layers), dict(shape=(28, 28), norm=(127.5, 1.0), lr=0.001, mom=0.0, wd=0.0, classes=10,
                                            batch=1000)
    if config == "wide_mlp":
        h = 16384
        s = 0.5 / math.sqrt(h)
        layers = [nn.Flatten()]
        prev = 128 * 128
        for _ in range(4):
            layers += [_uniform(nn.Linear(prev, h), -s, s), nn.ReLU()]
            prev = h
        layers.append(_uniform(nn.Linear(h, 1000), -s, s))
        return nn.Sequential(*layers), dict(shape=(128, 128), norm=(127.5, 1.0), lr=0.001, mom=0.9, wd=0.0,
                                            classes=1000, batch=4096)
    if config == "lenet":
        def fan(m, fan_in):
            b = 1.0 / math.sqrt(fan_in / 3.0)
            return _uniform(m, -b, b)
        net = nn.Sequential(
            fan(nn.Conv2d(1, 20, 5), 25), nn.MaxPool2d(2, 2, ceil_mode=True),
            fan(nn.Conv2d(20, 50, 5), 500), nn.MaxPool2d(2, 2, ceil_mode=True),
            nn.Flatten(), fan(nn.Linear(800, 500), 800), nn.ReLU(), fan(nn.Linear(500, 10), 500))
        return net, dict(shape=(1, 28, 28), norm=(255.0, 0.0), lr=0.01, mom=0.9, wd=0.0005, classes=10, batch=64)
    if config == "cifar10":
        net = nn.Sequential(
            _gauss(nn.Conv2d(3, 32, 5, 1, 2), 1e-4), nn.MaxPool2d(3, 2, ceil_mode=True), nn.ReLU(), _lrn(),
            _gauss(nn.Conv2d(32, 32, 5, 1, 2), 0.01), nn.ReLU(), nn.AvgPool2d(3, 2, ceil_mode=True), _lrn(),
            _gauss(nn.Conv2d(32, 64, 5, 1, 2), 0.01), nn.ReLU(), nn.AvgPool2d(3, 2, ceil_mode=True), _lrn(),
            nn.Flatten(), _gauss(nn.Linear(64 * 4 * 4, 10), 0.01))
        return net, dict(shape=(3, 32, 32), scale=1.0 / 128, lr=0.001, mom=0.9, wd=0.004, classes=10, batch=128)
    if config == "alexnet":
        net = nn.Sequential(
            _gauss(nn.Conv2d(3, 96, 11, 4, 0), 0.01), nn.ReLU(), nn.MaxPool2d(3, 2, ceil_mode=True), _lrn(5, 1e-4, 0.75),
            _gauss(nn.Conv2d(96, 256, 5, 1, 2), 0.01, 1.0), nn.ReLU(), nn.MaxPool2d(3, 2, ceil_mode=True),
            _lrn(5, 1e-4, 0.75),
            _gauss(nn.Conv2d(256, 384, 3, 1, 1), 0.01), nn.ReLU(),
            _gauss(nn.Conv2d(384, 384, 3, 1, 1), 0.01, 1.0), nn.ReLU(),
            _gauss(nn.Conv2d(384, 256, 3, 1, 1), 0.01, 1.0), nn.ReLU(), nn.MaxPool2d(3, 2, ceil_mode=True),
            nn.Flatten(),
            _gauss(nn.Linear(256 * 6 * 6, 4096), 0.005, 1.0), nn.ReLU(), nn.Dropout(0.5),
            _gauss(nn.Linear(4096, 4096), 0.005, 1.0), nn.ReLU(), nn.Dropout(0.5),
            _gauss(nn.Linear(4096, 1000), 0.01))
        return net, dict(shape=(3, 256, 256), crop=227, scale=1.0 / 128, lr=0.01, mom=0.9, wd=0.0005, classes=1000,
                         batch=256)
    raise ValueError(f"unknown config {config}")


class Trainer:
    """One rank of the baseline job: static device buffers, one captured step."""

    def __init__(self, config: str, dtype: str, batch: int, rank: int, world: int, device, use_graph=True,
                 bucket_elems=6_000_000):
        import numpy as np  # noqa: F401
        from singa_b200.layers.data import SyntheticSource
        tf32 = dtype == "fp32"
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cudnn.benchmark = True
        self.autocast = dtype == "bf16"
        self.rank, self.world, self.device = rank, world, device
        torch.manual_seed(1234)                                   # same initial weights on every rank
        self.model, cfg = build_model(config)
        self.cfg = cfg
        self.batch = batch or cfg["batch"]
        self.model.to(device)
        self.conv = len(cfg["shape"]) == 3
        if self.conv:
            self.model.to(memory_format=torch.channels_last)
        self.params = [p for p in self.model.parameters()]
        # gradients are views of one flat buffer (bucketed all-reduce without copies)
        n = sum(p.numel() for p in self.params)
        self.flat_grad = torch.zeros(n, device=device)
        off = 0
        self.ranges = []
        for p in self.params:
            p.grad = self.flat_grad[off:off + p.numel()].view_as(p)
            self.ranges.append((off, off + p.numel()))
            off += p.numel()
        self.opt = torch.optim.SGD(self.params, lr=cfg["lr"], momentum=cfg["mom"], weight_decay=cfg["wd"], foreach=True)
        # buckets in backward order (last parameters first), >= bucket_elems each; the first layer's parameters alone
        self.buckets = []
        hi = n
        cur_lo = n
        first_hi = self.ranges[1][1] if len(self.ranges) > 1 else n      # weight + bias of the first layer
        for lo, _ in reversed(self.ranges):
            if lo < first_hi:
                break
            cur_lo = lo
            if hi - cur_lo >= bucket_elems:
                self.buckets.append((cur_lo, hi))
                hi = cur_lo
        if hi > first_hi:
            self.buckets.append((first_hi, hi))
        self.buckets.append((0, first_hi))
        self.comm = torch.cuda.Stream()
        self._hooks()
        # data: the same synthetic records as the product arm; every rank reads its own slice of each global batch
        shape = list(cfg["shape"])
        self.source = SyntheticSource(shape, cfg["classes"], seed=0)
        d = int(math.prod(shape))
        self.h_pixels = torch.zeros(self.batch, d, dtype=torch.uint8, pin_memory=True)
        self.h_labels = torch.zeros(self.batch, dtype=torch.int32, pin_memory=True)
        self.d_pixels = torch.zeros(self.batch, d, dtype=torch.uint8, device=device)
        self.d_labels = torch.zeros(self.batch, dtype=torch.int64, device=device)
        self.d_labels32 = torch.zeros(self.batch, dtype=torch.int32, device=device)
        self.loss = torch.zeros(1, device=device)
        self.h_loss = torch.zeros(1, pin_memory=True)
        self.crop = cfg.get("crop", 0)
        self.off = torch.zeros(2, dtype=torch.int64, device=device)
        if rank:
            self.source.skip(rank * self.batch)
        self.use_graph = use_graph
        self.graph = None
        self.h2d_bytes = self.h_pixels.numel() + self.h_labels.numel() * 4
        self.next_host_batch()
        self.stage()

    # -- gradient all-reduce, overlapped with backward ------------------------------------------------------
    def _hooks(self):
        if self.world == 1:
            return
        # a bucket is reduced when the gradients of ALL its parameters have been accumulated (autograd does not
        # promise an order between a layer's weight and bias gradient)
        self._pending = [0] * len(self.buckets)
        self._need = [0] * len(self.buckets)
        for p, (a, b) in zip(self.params, self.ranges):
            for k, (lo, hi) in enumerate(self.buckets):
                if lo <= a and b <= hi:
                    self._need[k] += 1

                    def hook(_p, k=k, lo=lo, hi=hi):
                        self._pending[k] += 1
                        if self._pending[k] == self._need[k]:
                            self._pending[k] = 0
                            self.comm.wait_stream(torch.cuda.current_stream())
                            with torch.cuda.stream(self.comm):
                                dist.all_reduce(self.flat_grad[lo:hi], op=dist.ReduceOp.AVG)
                    p.register_post_accumulate_grad_hook(hook)
                    break

    # -- input ------------------------------------------------------------------------------------------------
    def next_host_batch(self):
        self.source.next_batch(self.h_pixels.numpy(), self.h_labels.numpy())
        if self.world > 1:
            self.source.skip((self.world - 1) * self.batch)

    def stage(self):
        self.d_pixels.copy_(self.h_pixels, non_blocking=True)
        self.d_labels32.copy_(self.h_labels, non_blocking=True)

    def parse(self):
        cfg = self.cfg
        x = self.d_pixels.view(self.batch, *cfg["shape"])
        if self.conv and len(cfg["shape"]) == 3 and cfg["shape"][0] == 1:
            x = x.float() / cfg["norm"][0] - cfg["norm"][1]
        elif self.conv:
            if self.crop:
                c = (cfg["shape"][1] - self.crop) // 2
                x = x[:, :, c:c + self.crop, c:c + self.crop]
            x = x.float() * cfg["scale"]
        else:
            x = x.float() / cfg["norm"][0] - cfg["norm"][1]
        if self.conv:
            x = x.contiguous(memory_format=torch.channels_last)
        return x

    # -- one step ---------------------------------------------------------------------------------------------
    def _body(self):
        x = self.parse()
        self.d_labels.copy_(self.d_labels32)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.autocast):
            logits = self.model(x)
        loss = F.cross_entropy(logits.float(), self.d_labels)
        self.flat_grad.zero_()
        loss.backward()
        if self.world > 1:
            torch.cuda.current_stream().wait_stream(self.comm)
        self.opt.step()
        self.loss.copy_(loss.detach().reshape(1))

    def step(self):
        if self.use_graph and self.graph is None and getattr(self, "_warm", 0) >= 3:
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(g):
                    self._body()
                self.graph = g
            except Exception as e:  # noqa: BLE001
                print(f"[baseline] CUDA-graph capture failed ({str(e).splitlines()[0]}); running eagerly", flush=True)
                torch.cuda.synchronize()
                self.use_graph = False
        if self.graph is not None:
            self.graph.replay()
        else:
            self._body()
            self._warm = getattr(self, "_warm", 0) + 1


def run(args, rank: int, local_rank: int, world: int, sampler_cls) -> dict:
    """Timing per the bench contract; returns the JSON fields of the library arm."""
    dev = torch.device("cuda", local_rank)
    t = Trainer(args.config, args.dtype, args.batch, rank, world, dev, use_graph=not args.no_graph)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 5)):
        t.step()
    barrier()
    sampler = sampler_cls(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        t.step()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    # end to end: every step a new host batch is assembled in pinned memory, copied to the device, and the loss is
    # copied back and read by the host (one step late, as in the product arm)
    copy_stream = torch.cuda.Stream()
    ev = [torch.cuda.Event() for _ in range(2)]
    barrier()
    t0 = time.perf_counter()
    last = 0.0
    for i in range(args.steps):
        t.next_host_batch()
        torch.cuda.current_stream().wait_stream(copy_stream)
        t.stage()
        t.step()
        if i:
            ev[(i - 1) % 2].synchronize()
            last = float(t.h_loss[0])
        t.h_loss.copy_(t.loss, non_blocking=True)
        ev[i % 2].record()
    ev[(args.steps - 1) % 2].synchronize()
    last = float(t.h_loss[0])
    barrier()
    e2e_s = time.perf_counter() - t0
    tt = torch.tensor([ms_total, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    return {"ms_total": float(tt[0]), "e2e_ms": float(tt[1]), "clocks": clocks, "per_gpu_batch": t.batch,
            "graph": t.graph is not None, "h2d": t.h2d_bytes, "d2h": 4, "last_loss": last,
            "buckets": len(t.buckets)}
