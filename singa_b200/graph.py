"""Layer graph: nodes, topological sort, partition-time node insertion and the node-link JSON dump.

Reference behaviour: include/utils/graph.h:18-150, src/utils/graph.cc:8-146.  The JSON produced by
``to_string`` has the same structure as ``script/node-link.json`` (``directed``, ``nodes`` with
id/color/shape, ``links`` with source/target indices) so ``script/graph.py``-style renderers keep working.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

CONNECTOR_ORIGINS = ("kSlice", "kConcate", "kSplit", "kBridgeSrc", "kBridgeDst")


@dataclass
class LayerInfo:
    origin: str = ""            # name of the original layer, or the connector type inserted by the partitioner
    locationid: int = 0
    partitionid: int = 0
    slice_dimension: int = -1
    concate_dimension: int = -1

    def copy(self, **kw) -> "LayerInfo":
        d = dict(origin=self.origin, locationid=self.locationid, partitionid=self.partitionid,
                 slice_dimension=self.slice_dimension, concate_dimension=self.concate_dimension)
        d.update(kw)
        return LayerInfo(**d)


@dataclass
class Node:
    name: str
    val: LayerInfo = field(default_factory=LayerInfo)
    srcnodes: List["Node"] = field(default_factory=list)
    dstnodes: List["Node"] = field(default_factory=list)

    def __hash__(self):
        return hash(self.name)

    def __eq__(self, other):
        return isinstance(other, Node) and other.name == self.name


class Graph:
    def __init__(self):
        self.nodes: List[Node] = []
        self.name2node: Dict[str, Node] = {}

    # -- construction --------------------------------------------------------------------------
    def add_node(self, name: str, info: Optional[LayerInfo] = None) -> Node:
        if name in self.name2node:
            raise ValueError(f"duplicate node {name!r}")
        node = Node(name, info if info is not None else LayerInfo(origin=name))
        self.nodes.append(node)
        self.name2node[name] = node
        return node

    def node(self, name: str) -> Node:
        return self.name2node[name]

    def _resolve(self, n) -> Node:
        return self.name2node[n] if isinstance(n, str) else n

    def add_edge(self, src, dst) -> None:
        s, d = self._resolve(src), self._resolve(dst)
        s.dstnodes.append(d)
        d.srcnodes.append(s)

    def remove_edge(self, src, dst) -> None:
        s, d = self._resolve(src), self._resolve(dst)
        s.dstnodes.remove(d)
        d.srcnodes.remove(s)

    # -- ordering ------------------------------------------------------------------------------
    def sort(self) -> None:
        """Topological order with sources first.  Iterative DFS that visits successors in reverse order
        and emits a node after all of its successors, then reverses — the traversal of graph.cc:66-101,
        so ties come out in the same order as the reference (insertion order of the layers)."""
        visited = {n.name: False for n in self.nodes}
        order: List[Node] = []
        for root in self.nodes:
            if visited[root.name]:
                continue
            stack: List[Tuple[Node, int]] = [(root, 0)]
            visited[root.name] = True
            while stack:
                node, idx = stack[-1]
                succ = node.dstnodes[::-1]
                while idx < len(succ) and visited[succ[idx].name]:
                    idx += 1
                if idx < len(succ):
                    stack[-1] = (node, idx + 1)
                    nxt = succ[idx]
                    visited[nxt.name] = True
                    stack.append((nxt, 0))
                else:
                    order.append(node)
                    stack.pop()
        self.nodes = order[::-1]

    def check_acyclic(self) -> bool:
        pos = {n.name: i for i, n in enumerate(self.nodes)}
        return all(pos[s.name] < pos[d.name] for s in self.nodes for d in s.dstnodes)

    # -- partition-time insertions (names: slice-X, concate-X, split-X, s-A-B / d-A-B) ------------
    def insert_slice_node(self, srcnode: Node, dstnodes: List[Node], info: LayerInfo,
                          connect_dst: bool = True) -> Node:
        node = self.add_node(f"slice-{srcnode.name}", info.copy(origin="kSlice"))
        self.add_edge(srcnode, node)
        if connect_dst:
            for d in dstnodes:
                self.add_edge(node, d)
        return node

    def insert_concate_node(self, srcnodes: List[Node], dstnode: Node, info: LayerInfo) -> Node:
        node = self.add_node(f"concate-{dstnode.name}", info.copy(origin="kConcate"))
        self.add_edge(node, dstnode)
        for s in srcnodes:
            self.add_edge(s, node)
        return node

    def insert_split_node(self, srcnode: Node, dstnodes: List[Node]) -> Node:
        node = self.add_node(f"split-{srcnode.name}", srcnode.val.copy(origin="kSplit"))
        self.add_edge(srcnode, node)
        for d in dstnodes:
            self.add_edge(node, d)
        return node

    def insert_bridge_node(self, srcnode: Node, dstnode: Node) -> Tuple[Node, Node]:
        """Replace the edge srcnode -> dstnode by srcnode -> s-.. -> d-.. -> dstnode IN PLACE (the position of the
        edge in dstnode's source list is kept: Concate / Slice consumers depend on the order of their sources).
        Works whether or not the direct edge still exists."""
        src = self.add_node(f"s-{srcnode.name}-{dstnode.name}", srcnode.val.copy(origin="kBridgeSrc"))
        dst = self.add_node(f"d-{srcnode.name}-{dstnode.name}", dstnode.val.copy(origin="kBridgeDst"))
        if dstnode in srcnode.dstnodes:
            i = srcnode.dstnodes.index(dstnode)
            j = dstnode.srcnodes.index(srcnode)
            srcnode.dstnodes[i] = src
            src.srcnodes.append(srcnode)
            dstnode.srcnodes[j] = dst
            dst.dstnodes.append(dstnode)
        else:
            self.add_edge(srcnode, src)
            self.add_edge(dst, dstnode)
        self.add_edge(src, dst)
        return src, dst

    # -- dump ----------------------------------------------------------------------------------
    def to_dict(self, info: Optional[Dict[str, str]] = None) -> dict:
        info = info or {}
        colors = ["red", "blue", "black", "green"]
        nodeid = {}
        nodes = []
        for i, n in enumerate(self.nodes):
            shape = "ellipse" if n.val.origin in CONNECTOR_ORIGINS else "box"
            nodes.append({"id": n.name + info.get(n.name, ""), "color": colors[n.val.locationid % len(colors)],
                          "shape": shape})
            nodeid[n.name] = i
        links = [{"source": nodeid[s.name], "target": nodeid[d.name], "color": "black"}
                 for s in self.nodes for d in s.dstnodes]
        return {"directed": 1, "nodes": nodes, "links": links}

    def to_string(self, info: Optional[Dict[str, str]] = None) -> str:
        return json.dumps(self.to_dict(info), indent=None, separators=(",", ":"))
