#!/usr/bin/env python
"""Render the node-link JSON that NeuralNet.to_string() / Graph.to_string() produce (the reference's
script/graph.py used pygraphviz; this writes Graphviz DOT, which needs no extra package)."""
import json
import sys


def to_dot(d):
    out = ["digraph net {", "  rankdir=TB;"]
    for i, n in enumerate(d["nodes"]):
        out.append(f'  n{i} [label="{n["id"]}", color={n["color"]}, shape={n["shape"]}];')
    for l in d["links"]:
        out.append(f'  n{l["source"]} -> n{l["target"]};')
    out.append("}")
    return "\n".join(out)


if __name__ == "__main__":
    src = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
    print(to_dot(json.loads(src)))
