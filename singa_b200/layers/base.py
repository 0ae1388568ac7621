"""Layer base class and the connector layers the partitioner inserts (Bridge, Concate, Slice, Split).

Reference: include/worker/base_layer.h:38-563, src/worker/base_layer.cc.  The reference leaves the
Concate/Slice/Split bodies empty (base_layer.cc:108-110,172-173,186-191); here they move data (views where
possible, one copy otherwise) so partitioned nets actually execute.

Storage conventions (native backend): activations are fp32 or bf16 (``ctx().act_dtype``); 2-D activations
``[batch, dim]`` live in a buffer whose row stride is a multiple of 16 bytes (TMA-legal), 4-D activations
``[N,C,H,W]`` are stored channels-last.  On the torch backend everything is plain fp32.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..device import ctx, pad8
from ..proto import schema as S
from ..utils.factory import Factory


def alloc_act(shape: Sequence[int], dtype: Optional[torch.dtype] = None, zero: bool = True) -> torch.Tensor:
    """Allocate an activation / gradient tensor of logical ``shape`` in the backend's storage convention."""
    c = ctx()
    dtype = dtype or c.act_dtype
    shape = [int(s) for s in shape]
    mk = torch.zeros if zero else torch.empty
    if c.is_native and len(shape) == 2:
        ld = pad8(shape[1]) if dtype == torch.bfloat16 else (shape[1] + 3) // 4 * 4
        return mk(shape[0], ld, dtype=dtype, device=c.device)[:, :shape[1]]
    if c.is_native and len(shape) == 4:
        return mk(shape, dtype=dtype, device=c.device).contiguous(memory_format=torch.channels_last)
    return mk(shape, dtype=dtype, device=c.device)


class Layer:
    type_id = "kLayer"

    def __init__(self):
        self.proto = S.LayerProto()
        self.name = ""
        self.srclayers: List["Layer"] = []
        self.dstlayers: List["Layer"] = []
        self.data: Optional[torch.Tensor] = None
        self.grad: Optional[torch.Tensor] = None
        self.shape_: List[int] = []
        self.forced_shape: Optional[List[int]] = None   # set when instantiated as one partition of a layer
        self.net = None                                  # back-pointer set by NeuralNet
        self.native = ctx().is_native                    # backend is fixed when the layer is created

    # -- construction ----------------------------------------------------------------------------
    def init(self, proto) -> None:
        self.proto.CopyFrom(proto)
        self.name = proto.name

    def init_from(self, other: "Layer", shape: Sequence[int]) -> None:
        """``Init(const Layer& other, const vector<int>& shape)``: a partition of ``other``."""
        self.proto.CopyFrom(other.proto)
        self.name = other.name
        self.forced_shape = [int(s) for s in shape]

    def set_name(self, name: str) -> None:
        self.name = name
        self.proto.name = name

    def setup(self) -> None:
        raise NotImplementedError

    def setup_after_partition(self) -> None:
        self.setup()

    def compute_feature(self, training: bool) -> None:
        raise NotImplementedError

    def compute_gradient(self) -> None:
        raise NotImplementedError

    def get_params(self) -> list:
        return []

    def to_proto(self, copy_data: bool = False):
        """Layer state for snapshots: the (possibly partition-adjusted) LayerProto; parameters are written
        separately by the checkpoint module (include/worker/base_layer.h:52-59)."""
        p = S.LayerProto()
        p.CopyFrom(self.proto)
        p.name = self.name
        return p

    # -- partitioning ----------------------------------------------------------------------------
    @property
    def partition_type(self) -> int:
        return self.proto.partition_type if self.proto.HasField("partition_type") else S.kNone

    def partition_dimension(self) -> int:
        pt = self.partition_type
        if pt == S.kLayerPartition:
            return 1
        if pt == S.kNone:
            return -1
        return 0

    def connection_type(self, k: int) -> int:
        assert k < len(self.srclayers)
        return S.kOneToOne

    @property
    def locationid(self) -> int:
        return self.proto.locationid

    @locationid.setter
    def locationid(self, v: int) -> None:
        self.proto.locationid = v

    @property
    def partitionid(self) -> int:
        return self.proto.partitionid

    @partitionid.setter
    def partitionid(self, v: int) -> None:
        self.proto.partitionid = v

    @property
    def type(self) -> str:
        return self.proto.type

    # -- data access -------------------------------------------------------------------------------
    def shape(self, dst: Optional["Layer"] = None) -> List[int]:
        return list(self.shape_)

    def data_for(self, dst: Optional["Layer"] = None) -> torch.Tensor:
        return self.data

    def grad_for(self, dst: Optional["Layer"] = None) -> torch.Tensor:
        return self.grad

    def _alloc(self, shape: Sequence[int], need_grad: bool = True) -> None:
        self.shape_ = [int(s) for s in shape]
        self.data = alloc_act(self.shape_)
        self.grad = alloc_act(self.shape_) if need_grad else None

    # -- kind flags ----------------------------------------------------------------------------------
    is_datalayer = False
    is_parserlayer = False
    is_losslayer = False
    is_bridgesrclayer = False
    is_bridgedstlayer = False
    is_connector = False

    def __repr__(self):
        return f"<{type(self).__name__} {self.name} {self.shape_}>"


# ================================================================================================
# Connector layers
# ================================================================================================
class BridgeLayer(Layer):
    is_connector = True

    def __init__(self):
        super().__init__()
        self.ready = False

    def set_ready(self, a: bool) -> None:
        self.ready = a


class BridgeSrcLayer(BridgeLayer):
    """Sending end of an edge that crosses a location boundary.  Holds no storage of its own: data/grad are
    those of the source layer; the executor moves the bytes (peer copy / send-recv)."""
    type_id = "kBridgeSrc"
    is_bridgesrclayer = True

    def setup(self):
        assert len(self.srclayers) == 1
        src = self.srclayers[0]
        self.shape_ = src.shape(self)
        self.data = src.data_for(self)
        self.grad = alloc_act(self.shape_)   # gradient arrives here from the far side

    def compute_feature(self, training):
        self.data = self.srclayers[0].data_for(self)
        self.ready = True

    def compute_gradient(self):
        src = self.srclayers[0]
        g = src.grad_for(self)
        if g is not None:
            g.copy_(self.grad)
        self.ready = False


class BridgeDstLayer(BridgeLayer):
    type_id = "kBridgeDst"
    is_bridgedstlayer = True

    def setup(self):
        assert len(self.srclayers) == 1
        self.shape_ = self.srclayers[0].shape(self)
        self.data = alloc_act(self.shape_)
        self.grad = alloc_act(self.shape_)

    def compute_feature(self, training):
        self.ready = False

    def compute_gradient(self):
        self.ready = True


class ConcateLayer(Layer):
    """Concatenate the sources along ``concate_dimension`` (base_layer.cc:85-101)."""
    type_id = "kConcate"
    is_connector = True

    def setup(self):
        dim = self.proto.concate_param.concate_dimension
        assert len(self.srclayers) > 1 and dim >= 0
        shape = self.srclayers[0].shape(self)
        for s in self.srclayers[1:]:
            sh = s.shape(self)
            assert len(sh) == len(shape)
            for i, (a, b) in enumerate(zip(shape, sh)):
                if i != dim:
                    assert a == b, f"concate {self.name}: shape mismatch {shape} vs {sh}"
            shape[dim] += sh[dim]
        self.dim = dim
        self._alloc(shape)

    def compute_feature(self, training):
        off = 0
        for s in self.srclayers:
            d = s.data_for(self)
            n = d.shape[self.dim]
            self.data.narrow(self.dim, off, n).copy_(d)
            off += n

    def compute_gradient(self):
        off = 0
        for s in self.srclayers:
            g = s.grad_for(self)
            n = s.shape(self)[self.dim]
            if g is not None:
                g.copy_(self.grad.narrow(self.dim, off, n))
            off += n


class SliceLayer(Layer):
    """Slice the source along ``slice_dimension`` into one piece per destination; the remainder goes to the
    last slice (base_layer.cc:114-171).  ``data_for(dst)`` / ``grad_for(dst)`` return that piece."""
    type_id = "kSlice"
    is_connector = True

    def setup(self):
        dim = self.proto.slice_param.slice_dimension
        num = self.proto.slice_param.slice_num or max(len(self.dstlayers), 1)
        assert len(self.srclayers) == 1 and dim >= 0 and num > 0
        shape = self.srclayers[0].shape(self)
        self.dim, self.num = dim, num
        base = shape[dim] // num
        self.sizes = [base] * num
        self.sizes[-1] += shape[dim] % num
        self.shape_ = shape
        self.data = alloc_act(shape)
        self.grad = alloc_act(shape)
        self._views()

    def _views(self):
        self.datavec, self.gradvec, off = [], [], 0
        for n in self.sizes:
            self.datavec.append(self.data.narrow(self.dim, off, n))
            self.gradvec.append(self.grad.narrow(self.dim, off, n))
            off += n

    def _slice_id(self, dst) -> int:
        if dst is None:
            return -1
        for i, d in enumerate(self.dstlayers):
            if d is dst or d.name == dst.name:
                return i if i < self.num else self.num - 1
        raise KeyError(f"{dst.name} is not a destination of {self.name}")

    def shape(self, dst=None):
        k = self._slice_id(dst)
        if k < 0:
            return list(self.shape_)
        s = list(self.shape_)
        s[self.dim] = self.sizes[k]
        return s

    def data_for(self, dst=None):
        k = self._slice_id(dst)
        return self.data if k < 0 else self.datavec[k]

    def grad_for(self, dst=None):
        k = self._slice_id(dst)
        return self.grad if k < 0 else self.gradvec[k]

    def compute_feature(self, training):
        self.data.copy_(self.srclayers[0].data_for(self))

    def compute_gradient(self):
        g = self.srclayers[0].grad_for(self)
        if g is not None:
            g.copy_(self.grad)


class SplitLayer(Layer):
    """Replicate the source to several destinations; gradients from the destinations are summed
    (base_layer.cc:175-191)."""
    type_id = "kSplit"
    is_connector = True

    def setup(self):
        assert len(self.srclayers) == 1
        self.shape_ = self.srclayers[0].shape(self)
        self.data = self.srclayers[0].data_for(self)
        n = self.proto.split_param.num_splits or max(len(self.dstlayers), 1)
        self.grads = [alloc_act(self.shape_) for _ in range(n)]
        self.grad = self.grads[0]

    def _dst_id(self, dst) -> int:
        for i, d in enumerate(self.dstlayers):
            if d is dst or (dst is not None and d.name == dst.name):
                return min(i, len(self.grads) - 1)
        return 0

    def grad_for(self, dst=None):
        return self.grads[self._dst_id(dst)]

    def compute_feature(self, training):
        self.data = self.srclayers[0].data_for(self)

    def compute_gradient(self):
        g = self.srclayers[0].grad_for(self)
        if g is None:
            return
        acc = self.grads[0].float()
        for t in self.grads[1:]:
            acc = acc + t.float()
        g.copy_(acc)


layer_factory: Factory[Layer] = Factory("Layer")
