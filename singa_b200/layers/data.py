"""Data layers (record sources) and parser layers (records -> device activations).

Reference: DataLayer / ParserLayer (include/worker/base_layer.h:335-560), ShardDataLayer, LMDBDataLayer,
MnistImageLayer, RGBImageLayer, LabelLayer (src/worker/layer.cc:217-328, 382-473, 573-673).

B200 design: a data layer assembles each batch as two *pinned host* arrays (uint8 pixels ``[N, C*H*W]`` and
int32 labels ``[N]``) instead of a vector of protobuf records; a parser layer copies them to a device staging
buffer (async, from pinned memory) and runs one tiny kernel (normalise / crop / mirror) that writes the
activation the first GEMM/conv consumes.  Staging is double-buffered so the copy of batch i+1 overlaps the
compute of batch i.
"""
from __future__ import annotations

import zlib

import os
import threading
from typing import List, Optional

import numpy as np
import torch

from ..device import ctx
from ..proto import schema as S
from .base import Layer, alloc_act


# ================================================================================================
# record sources
# ================================================================================================
class BatchSource:
    """Yields (pixels uint8 [N, D], labels int32 [N]) batches; ``sample_shape`` is the per-record shape."""
    sample_shape: List[int]

    def next_batch(self, pixels: np.ndarray, labels: np.ndarray) -> None:
        raise NotImplementedError

    def skip(self, n: int) -> None:
        pass

    def count(self) -> int:
        return 0


class SyntheticSource(BatchSource):
    """Deterministic pseudo-data of the configured shape (no network / datasets on the box).  A pool of
    pre-generated records lives in pinned host memory and is cycled; a batch that does not wrap around the pool is
    handed out as a VIEW (``next_view``) so the host->device copy reads the records in place, like DMA from a
    page-cached file, instead of assembling them into a second pinned buffer first."""

    def __init__(self, shape, classes=10, pool=4096, seed=0):
        self.sample_shape = [int(s) for s in shape]
        d = int(np.prod(self.sample_shape))
        rng = np.random.default_rng(seed)
        pool = max(min(pool, (512 << 20) // max(d, 1)), 1)      # at most 512 MB of pinned records
        labels = rng.integers(0, classes, size=pool, dtype=np.int32)
        # class-dependent mean so that the loss can actually go down on synthetic data
        base = rng.integers(0, 256, size=(classes, d), dtype=np.int32)
        pin = torch.cuda.is_available()
        self.pixels_t = torch.empty(pool, d, dtype=torch.uint8, pin_memory=pin)
        self.labels_t = torch.empty(pool, dtype=torch.int32, pin_memory=pin)
        self.pixels = self.pixels_t.numpy()
        self.labels = self.labels_t.numpy()
        self.labels[:] = labels
        for lo in range(0, pool, 256):                           # chunked: the int32 temporaries stay small
            hi = min(pool, lo + 256)
            noise = rng.integers(-64, 64, size=(hi - lo, d), dtype=np.int32)
            np.clip(base[labels[lo:hi]] + noise, 0, 255, out=noise)
            self.pixels[lo:hi] = noise
        self.pos = 0

    def next_view(self, n):
        """(pixels, labels) pinned tensor views of the next n records, or None when the batch wraps around."""
        pool = self.pixels.shape[0]
        if self.pos + n > pool:
            return None
        v = (self.pixels_t[self.pos:self.pos + n], self.labels_t[self.pos:self.pos + n])
        self.pos = (self.pos + n) % pool
        return v

    def next_batch(self, pixels, labels):
        n = pixels.shape[0]
        pool = self.pixels.shape[0]
        done = 0
        while done < n:                                          # contiguous slices (np.take is ~4x slower)
            m = min(n - done, pool - self.pos)
            pixels[done:done + m] = self.pixels[self.pos:self.pos + m]
            labels[done:done + m] = self.labels[self.pos:self.pos + m]
            self.pos = (self.pos + m) % pool
            done += m

    def skip(self, n):
        self.pos = (self.pos + n) % self.pixels.shape[0]

    def count(self):
        return self.pixels.shape[0]


class ShardSource(BatchSource):
    """Records from a Shard folder (``shard.dat``), wrapping around at the end."""

    def __init__(self, folder: str):
        from ..io.shard import Shard
        self.shard = Shard(folder, Shard.kRead)
        rec = self._read_one()
        if rec is None:
            raise RuntimeError(f"shard {folder} is empty")
        self.sample_shape = list(rec.image.shape)
        # float-valued records (Record.image.data: e.g. mean-subtracted images written by the loader tool) are staged
        # as float32 and reach the parser unchanged, as in the reference (src/worker/layer.cc:595-598)
        self.is_float = len(rec.image.data) > 0
        self.shard.seek_to_first()

    def _read_one(self):
        kv = self.shard.next()
        if kv is None:
            return None
        rec = S.Record()
        rec.ParseFromString(kv[1])
        return rec

    def next_batch(self, pixels, labels):
        from ..io.shard import read_image_batch
        read_image_batch(self.shard, pixels, labels)

    def skip(self, n):
        for _ in range(n):
            if self.shard.next() is None:
                self.shard.seek_to_first()

    def count(self):
        return self.shard.count()


def make_source(data_param, type_id: str, seed: int = 0) -> BatchSource:
    path = data_param.path
    src = data_param.source
    force_syn = os.environ.get("SINGA_B200_SYNTHETIC", "") == "1"
    if src == "synthetic" or (force_syn and not os.path.exists(path)):
        shape = list(data_param.synthetic_shape) or [28, 28]
        return SyntheticSource(shape, data_param.synthetic_classes, seed=seed)
    if type_id == "kLMDBData":
        from ..io.lmdb_reader import LMDBSource
        return LMDBSource(path)
    return ShardSource(path)


# ================================================================================================
# data layers
# ================================================================================================
class DataLayer(Layer):
    is_datalayer = True

    def __init__(self):
        super().__init__()
        self.batchsize = 0
        self.source: Optional[BatchSource] = None
        self.sample_shape: List[int] = []
        self.random_skip = 0
        self.pixels: Optional[torch.Tensor] = None   # pinned host uint8 [N, D]
        self.labels: Optional[torch.Tensor] = None   # pinned host int32 [N]
        self._np_pixels = None
        self._np_labels = None

    @property
    def partition_type(self):
        return S.kNone

    def setup(self):
        dp = self.proto.data_param
        self.batchsize = int(dp.batchsize)
        assert self.batchsize > 0, f"data layer {self.name}: batchsize must be set"
        self.random_skip = int(dp.random_skip)
        self.source = make_source(dp, self.type_id, seed=zlib.crc32(self.name.encode()) & 0xFFFF)
        self.sample_shape = list(self.source.sample_shape)
        d = int(np.prod(self.sample_shape))
        pin = torch.cuda.is_available()
        px_dtype = torch.float32 if getattr(self.source, "is_float", False) else torch.uint8
        self.pixels = torch.zeros(self.batchsize, d, dtype=px_dtype, pin_memory=pin)
        self.labels = torch.zeros(self.batchsize, dtype=torch.int32, pin_memory=pin)
        self._own_pixels, self._own_labels = self.pixels, self.labels
        self._h2d_pending = []
        self._np_pixels = self.pixels.numpy()
        self._np_labels = self.labels.numpy()
        self.shape_ = [self.batchsize] + self.sample_shape
        if self.random_skip:
            n = int(np.random.default_rng(1234).integers(0, self.random_skip + 1))
            self.source.skip(n)
            self.random_skip = 0
        # data parallel worker groups read disjoint records: group g takes batch g of every window of
        # `ngroups` batches (the reference partitions the record list per worker group offline,
        # script/load_data.py; here the stride is applied at read time)
        self.group_id, self.ngroups = 0, 1
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                self.group_id, self.ngroups = dist.get_rank(), dist.get_world_size()
            else:
                # ZeroMQ jobs have no process group: the Cluster singleton knows this worker's group
                from ..cluster import Cluster
                c = Cluster._instance
                if c is not None and c.am_i_worker() and c.ngroups > 1:
                    self.group_id, self.ngroups = c.groupid, c.ngroups
        except Exception:  # noqa: BLE001
            pass
        if self.group_id:
            self.source.skip(self.group_id * self.batchsize)

    def compute_feature(self, training):
        """Read the next batch of records into the pinned host buffers (or take a pinned view of them when the
        source can hand the records out in place)."""
        view = self.source.next_view(self.batchsize) if hasattr(self.source, "next_view") else None
        if view is not None:
            self.pixels, self.labels = view
        else:
            # the pinned batch buffers are about to be overwritten: the host->device copies that read the previous
            # batch out of them (issued by the parser layers) must have finished
            for ev in self._h2d_pending:
                ev.synchronize()
            self._h2d_pending.clear()
            self.pixels, self.labels = self._own_pixels, self._own_labels
            self.source.next_batch(self._np_pixels, self._np_labels)
        if self.ngroups > 1:
            self.source.skip((self.ngroups - 1) * self.batchsize)

    prefetching = compute_feature

    def compute_gradient(self):
        pass


class ShardDataLayer(DataLayer):
    type_id = "kShardData"


class LMDBDataLayer(DataLayer):
    type_id = "kLMDBData"


# ================================================================================================
# parser layers
# ================================================================================================
class ParserLayer(Layer):
    """Turns the data layer's host batch into a device activation.  ``prefetching`` (any thread / side
    stream) stages batch i+1 while batch i is being consumed; ``compute_feature`` swaps the staged batch in
    (the mutex/condvar double buffer of base_layer.h:510-537, expressed with CUDA events)."""
    is_parserlayer = True

    def __init__(self):
        super().__init__()
        self.staging: List[Optional[torch.Tensor]] = [None, None]
        self.stage_events: List[Optional[torch.cuda.Event]] = [None, None]
        self.parse_events: List[Optional[torch.cuda.Event]] = [None, None]
        self.stage_fill = 0          # staging slot the next prefetch writes
        self.stage_ready = -1        # slot holding a staged-but-unconsumed batch
        self._lock = threading.Lock()

    @property
    def partition_type(self):
        return S.kNone

    def _data_layer(self) -> DataLayer:
        d = self.srclayers[0]
        assert d.is_datalayer, f"{self.name}: parser layers read from a data layer"
        return d

    def compute_gradient(self):
        pass

    # host batch -> device staging (async when pinned + CUDA)
    def _stage(self, host: torch.Tensor, slot: int, stream=None) -> torch.Tensor:
        dev = ctx().device
        if self.staging[slot] is None or self.staging[slot].shape != host.shape:
            self.staging[slot] = torch.empty_like(host, device=dev)
        if dev.type == "cuda":
            st = stream or torch.cuda.current_stream()
            with torch.cuda.stream(st):
                if self.parse_events[slot] is not None:
                    st.wait_event(self.parse_events[slot])   # the parse kernel that last read this slot is done
                self.staging[slot].copy_(host, non_blocking=True)
                ev = self.stage_events[slot] or torch.cuda.Event()
                ev.record(st)
                self.stage_events[slot] = ev
                pend = getattr(self._data_layer(), "_h2d_pending", None)
                if pend is not None and ev not in pend:
                    pend.append(ev)
        else:
            self.staging[slot].copy_(host)
        return self.staging[slot]

    def host_tensor(self) -> torch.Tensor:
        raise NotImplementedError

    def parse(self, staged: torch.Tensor, training: bool) -> None:
        raise NotImplementedError

    def prefetching(self, training: bool, stream=None) -> None:
        with self._lock:
            slot = self.stage_fill
            self._stage(self.host_tensor(), slot, stream)
            self.stage_ready = slot
            self.stage_fill ^= 1

    def compute_feature(self, training):
        with self._lock:
            if self.stage_ready < 0:
                slot = self.stage_fill
                self._stage(self.host_tensor(), slot)
                self.stage_fill ^= 1
            else:
                slot = self.stage_ready
            self.stage_ready = -1
        self.last_slot = slot
        ev = self.stage_events[slot]
        if ev is not None and ctx().device.type == "cuda":
            torch.cuda.current_stream().wait_event(ev)
        self.parse(self.staging[slot], training)
        if ctx().device.type == "cuda":
            pe = self.parse_events[slot] or torch.cuda.Event()
            pe.record(torch.cuda.current_stream())
            self.parse_events[slot] = pe

    # -- rotated schedule (Executor): the output is double buffered.  Batch t+1 is parsed into the buffer that is NOT
    #    ``data`` while the graph of step t still reads batch t (weight gradient of the first layer); ``flip`` makes it
    #    current between backward(t) and forward(t+1).  No copy: consumers fetch ``data`` at call time and keep the
    #    tensor they read in forward for their backward pass.
    def _other(self) -> torch.Tensor:
        if getattr(self, "_bufs", None) is None:
            self._bufs = [self.data, torch.empty_like(self.data)]
        return self._bufs[1] if self.data is self._bufs[0] else self._bufs[0]

    def flip(self) -> None:
        self.data = self._other()

    def compute_feature_next(self, training: bool) -> None:
        cur, self.data = self.data, self._other()
        try:
            self.compute_feature(training)
        finally:
            self.data = cur

    def reparse_next(self, training: bool) -> None:
        cur, self.data = self.data, self._other()
        try:
            self.parse(self.staging[self.last_slot], training)
        finally:
            self.data = cur

    def reparse(self, training: bool) -> None:
        """Parse the most recently consumed staged batch again (device-resident input, no host traffic)."""
        self.parse(self.staging[self.last_slot], training)

    @property
    def h2d_bytes(self) -> int:
        h = self.host_tensor()
        return h.numel() * h.element_size()


class MnistImageLayer(ParserLayer):
    """uint8 pixel -> x / norm_a - norm_b, output {N, s, s} (src/worker/layer.cc:382-473).  The distortion fields of
    MnistProto (gamma: random scaling in %, beta: random rotation in degrees or shear, kernel / sigma / alpha: elastic
    displacement field) are applied to TRAINING batches — the reference parses them and then skips the block
    (commented out, layer.cc:406-438); the intended augmentation is implemented here (``sg_mnist_distort``)."""
    type_id = "kMnistImage"

    def setup(self):
        d = self._data_layer()
        mp = self.proto.mnist_param
        self.norm_a, self.norm_b = mp.norm_a, mp.norm_b
        sh = d.sample_shape
        assert len(sh) >= 2 and sh[-1] == sh[-2], f"{self.name}: MNIST records must be square, got {sh}"
        s = mp.resize if mp.resize else sh[-1]
        self.in_size = sh[-1]
        self.gamma, self.beta = float(mp.gamma), float(mp.beta)
        self.ekernel, self.esigma, self.ealpha = int(mp.kernel), float(mp.sigma), float(mp.alpha)
        self.distort = bool(self.gamma or self.beta or (self.ekernel and self.esigma and self.ealpha))
        if s != sh[-1] and not self.distort:
            raise NotImplementedError(f"{self.name}: MnistProto.resize without a distortion is not supported")
        if self.distort:
            self._gen = torch.Generator(device=ctx().device)
            self._gen.manual_seed(97531)
        self._alloc([d.batchsize, s, s], need_grad=False)

    def host_tensor(self):
        return self._data_layer().pixels

    def parse(self, staged, training):
        if self.distort and training:
            from ..ops import reference as R
            n, out = self.shape_[0], self.shape_[1]
            labels = self._data_layer().labels
            A, disp = R.mnist_distortion_params(n, self.in_size, out, self.gamma, self.beta, self.ekernel, self.esigma,
                                                self.ealpha, labels, self._gen, ctx().device)
            self._last_distortion = (A, disp)
            if self.native:
                from ..ops import kernels as K
                K.mnist_distort(staged, self.data, n, self.in_size, out, A, disp, self.norm_a, self.norm_b)
            else:
                src = staged.view(n, self.in_size, self.in_size)
                self.data.copy_(R.mnist_distort(src, A, disp, out, self.norm_a, self.norm_b))
            return
        if self.native:
            from ..ops import kernels as K
            K.mnist_parse(staged, self.data, self.norm_a, self.norm_b)
        else:
            self.data.copy_((staged.to(torch.float32) / self.norm_a - self.norm_b).view(self.shape_))


class RGBImageLayer(ParserLayer):
    """Random crop (training) / centre crop (test), random mirror, scale; output {N, 3, c, c}
    (intent of src/worker/layer.cc:573-643; its non-mirror path never writes the image — fixed here)."""
    type_id = "kRGBImage"

    def setup(self):
        d = self._data_layer()
        rp = self.proto.rgbimage_param
        self.scale, self.cropsize, self.mirror = rp.scale, rp.cropsize, rp.mirror
        sh = d.sample_shape
        assert len(sh) == 3, f"{self.name}: RGB records must be C,H,W, got {sh}"
        self.C, self.H, self.W = sh
        out = self.cropsize if self.cropsize else None
        self._alloc([d.batchsize, self.C, out or self.H, out or self.W], need_grad=False)
        n = d.batchsize
        dev = ctx().device
        self.off_h = torch.zeros(n, dtype=torch.int32, device=dev)
        self.off_w = torch.zeros(n, dtype=torch.int32, device=dev)
        self.mir = torch.zeros(n, dtype=torch.int32, device=dev)
        self._gen = torch.Generator(device=dev)
        self._gen.manual_seed(4321)

    def host_tensor(self):
        return self._data_layer().pixels

    def parse(self, staged, training):
        n = self.shape_[0]
        crop = self.cropsize
        if crop:
            if training:
                self.off_h.copy_(torch.randint(0, self.H - crop + 1, (n,), generator=self._gen,
                                               device=self.off_h.device, dtype=torch.int32))
                self.off_w.copy_(torch.randint(0, self.W - crop + 1, (n,), generator=self._gen,
                                               device=self.off_w.device, dtype=torch.int32))
            else:
                self.off_h.fill_((self.H - crop) // 2)
                self.off_w.fill_((self.W - crop) // 2)
        if self.mirror and training:
            self.mir.copy_(torch.randint(0, 2, (n,), generator=self._gen, device=self.mir.device,
                                         dtype=torch.int32))
        else:
            self.mir.zero_()
        if self.native:
            from ..ops import kernels as K
            K.rgb_parse(staged, None, self.data, n, self.C, self.H, self.W, crop, self.off_h, self.off_w,
                        self.mir, self.scale)
        else:
            img = staged.view(n, self.C, self.H, self.W).to(torch.float32)
            oh = crop or self.H
            ow = crop or self.W
            ar = torch.arange(n, device=img.device)
            ys = self.off_h.long()[:, None] + torch.arange(oh, device=img.device)[None, :]
            xs_base = torch.arange(ow, device=img.device)[None, :].expand(n, ow)
            xs_m = torch.where(self.mir.bool()[:, None], ow - 1 - xs_base, xs_base) + self.off_w.long()[:, None]
            out = img[ar[:, None, None, None], torch.arange(self.C, device=img.device)[None, :, None, None],
                      ys[:, None, :, None], xs_m[:, None, None, :]]
            self.data.copy_(out * self.scale)


class LabelLayer(ParserLayer):
    """label -> {N} (src/worker/layer.cc:217-233; the reference's ``label < 10`` CHECK is not kept)."""
    type_id = "kLabel"

    def setup(self):
        d = self._data_layer()
        self.shape_ = [d.batchsize]
        self.data = torch.zeros(d.batchsize, dtype=torch.int32, device=ctx().device)
        self.grad = None

    def host_tensor(self):
        return self._data_layer().labels

    def parse(self, staged, training):
        self.data.copy_(staged)
