"""NeuralNet: layer-graph construction from ``NetProto``, shape setup, parameter numbering, weight sharing,
the graph partitioner (Slice / Concate / Split / Bridge insertion) and the activation-fusion pass.

Reference: include/worker/neuralnet.h:27-148, src/worker/neuralnet.cc:13-393.  Construction and partition
rules follow the reference so its golden layer counts hold (8 layers unpartitioned, 28 with kDataPartition over
a group of 3; src/test/test_neuralnet.cc:109-125); the fusion pass and ``to_proto`` are additions.
"""
from __future__ import annotations

from typing import Dict, List, Optional

from . import layers as L
from .graph import Graph, LayerInfo
from .layers.base import Layer, layer_factory
from .ops.gemm import ACT_NONE
from .param import Param, register_param_type
from .proto import schema as S
from .utils.common import int_vec_to_string


def registry_layers() -> None:
    """``NeuralNet::RegistryLayers`` — the 18 built-in layer types (src/worker/neuralnet.cc:13-33)."""
    for cls in L.ALL_LAYERS:
        layer_factory.register(cls.type_id, cls)


def registry_param(param_type: str) -> None:
    register_param_type(param_type)


registry_layers()


class NeuralNet:
    def __init__(self, net_proto, group_size: int = 1, fuse: bool = True):
        proto = S.NetProto()
        proto.CopyFrom(net_proto)
        for lp in proto.layer:
            if not lp.HasField("partition_type"):
                lp.partition_type = proto.partition_type
        self.group_size = group_size
        self.graph = Graph()
        self.param_shares: Dict[int, int] = {}
        self.layers: List[Layer] = []
        self.name2layer_: Dict[str, Layer] = {}
        self.params: List[Param] = []
        self._construct(proto)
        if group_size > 1:
            self._partition()
        pid = 0
        for layer in self.layers:
            layer.net = self
            for p in layer.get_params():
                self.params.append(p)
                p.id = pid
                pid += 1
        self.datalayers = [l for l in self.layers if l.is_datalayer]
        self.parserlayers = [l for l in self.layers if l.is_parserlayer]
        self.losslayers = [l for l in self.layers if l.is_losslayer]
        if fuse:
            self.fuse_activations()

    # ------------------------------------------------------------------------------------------
    def _construct(self, net_proto) -> None:
        protos = {}
        for lp in net_proto.layer:
            self.graph.add_node(lp.name)
            protos[lp.name] = lp
        for lp in net_proto.layer:
            for src in lp.srclayers:
                if src not in protos:
                    raise ValueError(f"layer {lp.name!r} names unknown source layer {src!r}")
                self.graph.add_edge(src, lp.name)
        self.graph.sort()
        for node in self.graph.nodes:
            lp = protos[node.name]
            layer = layer_factory.create(lp.type)
            layer.init(lp)
            self.name2layer_[node.name] = layer
            self.layers.append(layer)
        self._connect()
        for layer in self.layers:
            layer.setup()

    def _connect(self) -> None:
        for node in self.graph.nodes:
            layer = self.name2layer_[node.name]
            layer.dstlayers = [self.name2layer_[d.name] for d in node.dstnodes]
            layer.srclayers = [self.name2layer_[s.name] for s in node.srcnodes]

    # ------------------------------------------------------------------------------------------
    def _partition(self) -> None:
        old_name2layer = dict(self.name2layer_)
        self.graph = self._create_partitioned_graph(self.layers, old_name2layer)
        self.name2layer_ = {}
        self.layers = []
        gsize = self.group_size
        for node in self.graph.nodes:
            origin = node.val.origin
            proto = S.LayerProto()
            proto.name = node.name
            proto.locationid = node.val.locationid
            proto.partitionid = node.val.partitionid
            if origin == "kSlice":
                proto.type = origin
                proto.slice_param.slice_dimension = node.val.slice_dimension
                proto.slice_param.slice_num = len(node.dstnodes)
            elif origin == "kConcate":
                proto.type = origin
                proto.concate_param.concate_dimension = node.val.concate_dimension
                proto.concate_param.concate_num = len(node.srcnodes)
            elif origin == "kSplit":
                proto.type = origin
                proto.split_param.num_splits = len(node.dstnodes)
            elif origin in ("kBridgeSrc", "kBridgeDst"):
                proto.type = origin
            elif origin not in old_name2layer:
                raise ValueError(f"unknown origin {origin!r} for node {node.name!r}")
            if proto.HasField("type"):
                layer = layer_factory.create(proto.type)
                layer.init(proto)
            else:
                old = old_name2layer[origin]
                if old.partition_type == S.kNone:
                    layer = old
                else:
                    shape = old.shape(None)
                    pdim = old.partition_dimension()
                    shape[pdim] = shape[pdim] // gsize + \
                        (shape[pdim] % gsize if node.val.partitionid == gsize - 1 else 0)
                    layer = layer_factory.create(old.type)
                    layer.init_from(old, shape)
                    layer.set_name(node.name)
                    layer.locationid = node.val.locationid
                    layer.partitionid = node.val.partitionid
            self.layers.append(layer)
            self.name2layer_[node.name] = layer
        self._connect()
        for layer in self.layers:
            expect = layer.forced_shape
            layer.setup_after_partition()
            if expect is not None:
                got = layer.shape(None)
                assert list(expect) == list(got), \
                    f"{layer.name}: shape after partition {got} != expected {expect}"
        # replicas of a DATA-partitioned layer are the same layer on different slices of the batch: they share
        # one set of parameters and accumulate into one gradient (the reference aggregates the gradients of the
        # threads sharing a Param and applies ONE update with grad_scale = 1/shares,
        # src/utils/param_manager.cc:169-190)
        first: Dict[str, Layer] = {}
        for node in self.graph.nodes:
            layer = self.name2layer_[node.name]
            origin = node.val.origin
            old = old_name2layer.get(origin)
            if old is None or old.partition_type != S.kDataPartition or not layer.get_params():
                continue
            layer._accumulate_wgrad = True        # every replica adds into the shared (zeroed) gradient
            if origin not in first:
                first[origin] = layer
                continue
            lead = first[origin]
            for p, q in zip(layer.get_params(), lead.get_params()):
                p.share_data(q)
                p.grad = q.grad
                self.param_shares[id(q)] = self.param_shares.get(id(q), 1) + 1
            layer._accumulate_wgrad = True

    def _create_partitioned_graph(self, layers, name2layer) -> Graph:
        graph = Graph()
        gsize = self.group_size
        layer2nodes: Dict[str, list] = {}
        for layer in layers:
            nodes = []
            pt = layer.partition_type
            if pt in (S.kDataPartition, S.kLayerPartition):
                for i in range(gsize):
                    nodes.append(graph.add_node(f"{layer.name}-{i:02d}", LayerInfo(layer.name, i, i, -1, -1)))
            elif pt == S.kNone:
                nodes.append(graph.add_node(layer.name, LayerInfo(layer.name, layer.locationid, 0, -1, -1)))
            else:
                raise ValueError(f"unknown partition type {pt}")
            layer2nodes[layer.name] = nodes

        for layer in layers:
            name, typ = layer.name, layer.partition_type
            nodes = layer2nodes[name]
            for srcid, srclayer in enumerate(layer.srclayers):
                srcname = srclayer.name
                srcnodes = layer2nodes[srcname]
                srctype = srclayer.partition_type
                connection = layer.connection_type(srcid)
                partitioned = (S.kDataPartition, S.kLayerPartition)
                if srctype == S.kNone:
                    assert len(srcnodes) == 1, f"local layer {srcname} should not be partitioned"
                    srcnode = srcnodes[0]
                    if typ == S.kDataPartition or (typ == S.kLayerPartition and connection == S.kOneToOne):
                        info = srcnode.val.copy(slice_dimension=name2layer[name].partition_dimension())
                        graph.insert_slice_node(srcnode, nodes, info)
                    elif typ == S.kNone:
                        assert len(nodes) == 1
                        graph.add_edge(srcnode, nodes[0])
                    else:   # kLayerPartition with kOneToAll: every partition needs the whole source
                        graph.insert_split_node(srcnode, nodes)
                elif (typ == S.kNone and srctype in partitioned) or \
                        (typ == S.kLayerPartition and connection == S.kOneToAll and srctype in partitioned):
                    for node in nodes:
                        info = node.val.copy(concate_dimension=name2layer[srcname].partition_dimension())
                        assert info.concate_dimension >= 0
                        graph.insert_concate_node(srcnodes, node, info)
                elif (srctype == S.kLayerPartition and typ == S.kDataPartition) or \
                        (srctype == S.kDataPartition and typ == S.kLayerPartition):
                    slicenodes = []
                    for srcnode in srcnodes:
                        info = srcnode.val.copy(slice_dimension=name2layer[name].partition_dimension())
                        slicenodes.append(graph.insert_slice_node(srcnode, nodes, info, connect_dst=False))
                    for node in nodes:
                        info = node.val.copy(concate_dimension=name2layer[srcname].partition_dimension())
                        assert info.concate_dimension >= 0
                        graph.insert_concate_node(slicenodes, node, info)
                elif (srctype == S.kDataPartition and typ == S.kDataPartition) or \
                        (srctype == S.kLayerPartition and typ == S.kLayerPartition
                         and connection == S.kOneToOne):
                    assert len(srcnodes) == len(nodes)
                    for a, b in zip(srcnodes, nodes):
                        graph.add_edge(a, b)
        graph.sort()

        # a node feeding several consumers gets an explicit Split (the first node — the data layer — excepted)
        data_node = True
        for node in list(graph.nodes):
            if len(node.dstnodes) > 1 and node.val.origin not in ("kSlice", "kSplit") and not data_node:
                dsts = list(node.dstnodes)
                for d in dsts:
                    graph.remove_edge(node, d)
                graph.insert_split_node(node, dsts)
            data_node = False

        # edges that cross a location boundary become BridgeSrc -> BridgeDst pairs
        for node in list(graph.nodes):
            for dst in list(node.dstnodes):
                if node.val.locationid != dst.val.locationid:
                    graph.insert_bridge_node(node, dst)      # replaces the edge in place (source order kept)
        graph.sort()
        return graph

    def zero_shared_grads(self) -> None:
        """Gradients shared by the replicas of a data-partitioned layer are accumulated: clear them before a
        backward pass."""
        for layer in self.layers:
            if getattr(layer, "_accumulate_wgrad", False):
                for p in layer.get_params():
                    if p.is_owner:
                        p.grad.zero_()

    # ------------------------------------------------------------------------------------------
    def fuse_activations(self) -> None:
        """Fold ReLU/Tanh layers into the GEMM epilogues of their producer (forward) and consumer (backward)."""
        from .layers.neuron import ActivationLayer, ConvolutionLayer, InnerProductLayer
        gemm_layers = (ConvolutionLayer, InnerProductLayer)
        for act in self.layers:
            if not isinstance(act, ActivationLayer) or len(act.srclayers) != 1:
                continue
            prod = act.srclayers[0]
            if isinstance(prod, gemm_layers) and len(prod.dstlayers) == 1 and prod.fused_act == ACT_NONE \
                    and prod.data.dtype == act.data.dtype:
                prod.fused_act = act.act
                act.alias_to_source()
            if len(act.dstlayers) == 1 and isinstance(act.dstlayers[0], gemm_layers):
                cons = act.dstlayers[0]
                if cons.srclayers[0] is act and prod.grad_for(act) is not None \
                        and not (prod.is_datalayer or prod.is_parserlayer):
                    cons.src_grad_act = act.act
                    cons.src_grad_target = prod
                    act.bwd_fused = True

    # ------------------------------------------------------------------------------------------
    def name2layer(self, name: str) -> Optional[Layer]:
        return self.name2layer_.get(name)

    def share_weights(self, other: "NeuralNet") -> None:
        """Alias this net's parameters to ``other``'s, layer by layer name (neuralnet.cc:379-391)."""
        for layer in self.layers:
            ol = other.name2layer(layer.name)
            if ol is None:
                continue
            mine, theirs = layer.get_params(), ol.get_params()
            assert len(mine) == len(theirs), f"layer {layer.name}: parameter count differs between nets"
            for a, b in zip(mine, theirs):
                a.share_data(b)

    def to_string(self) -> str:
        info = {l.name: int_vec_to_string(l.shape(None)) for l in self.layers}
        return self.graph.to_string(info)

    def to_adjacency(self) -> str:
        return "".join(f"{l.name}: " + "".join(f"{d.name}, " for d in l.dstlayers) + "\n" for l in self.layers)

    def to_proto(self, copy_data: bool = False):
        """NetProto of the (possibly partitioned) net — a stub in the reference (neuralnet.cc:346-348)."""
        proto = S.NetProto()
        for l in self.layers:
            lp = proto.layer.add()
            lp.CopyFrom(l.to_proto(copy_data))
            del lp.srclayers[:]
            lp.srclayers.extend(s.name for s in l.srclayers)
        return proto

    def debug_info(self) -> str:
        """Per-layer L1 norms of data / grad / params (neuralnet.cc:350-378)."""
        from .utils.norms import asum
        out = []
        for l in self.layers:
            if not l.is_datalayer and l.data is not None and l.data.is_floating_point():
                out.append("Forward layer  %10s data norm1 %13.9f\n" % (l.name, asum(l.data)))
        for l in reversed(self.layers):
            if not (l.is_datalayer or l.is_losslayer or l.is_parserlayer) and l.grad is not None:
                out.append("Backward layer %10s grad norm1 %13.9f\n" % (l.name, asum(l.grad)))
        for l in self.layers:
            for p in l.get_params():
                out.append("Layer %10s, param id %2d, name %10s, value norm1 %13.9f, grad norm1 %13.9f\n"
                           % (l.name, p.id, p.name, asum(p.data), asum(p.grad)))
        return "".join(out)
