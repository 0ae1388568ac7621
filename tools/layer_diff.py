"""Per-layer comparison of the native path against the plain-PyTorch fp32 path of the same net on the same batch:
prints, for every layer, the relative difference of its output after one forward pass and of its parameter gradients
after one backward pass.  Usage: python tools/layer_diff.py <config> [--dtype fp32|bf16] [--batch N]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from singa_b200 import device  # noqa: E402
from singa_b200.models import configs  # noqa: E402
from singa_b200.worker import Worker  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("config")
ap.add_argument("--dtype", default="fp32")
ap.add_argument("--batch", type=int, default=8)
args = ap.parse_args()


def build(backend, dtype):
    device.set_context("cuda", backend, dtype)
    w = Worker(use_graph=False)
    w.build(configs.CONFIGS[args.config](batch=args.batch))
    w.executor.model.prefetch = False
    return w


ref, nat = build("torch", "fp32"), build("native", args.dtype)
for a, b in zip(nat.train_net.params, ref.train_net.params):
    a.data.copy_(b.data)
nat.pm.arena.sync_mirror()
for name, w in (("ref", ref), ("nat", nat)):
    device.set_context("cuda", "torch" if name == "ref" else "native", "fp32" if name == "ref" else args.dtype)
    ex = w.executor
    w.pm.begin_step(0)
    ex._feed(w.train_net, True)
    ex.forward(w.train_net, 0, True, skip_inputs=True)
    for layer in reversed(w.train_net.layers):
        if not (layer.is_datalayer or layer.is_parserlayer):
            layer.compute_gradient()
    torch.cuda.synchronize()
for ln, lr in zip(nat.train_net.layers, ref.train_net.layers):
    if ln.data is None or lr.data is None or ln.data.dtype not in (torch.float32, torch.bfloat16):
        continue
    a, b = ln.data.float(), lr.data.float()
    if a.shape != b.shape:
        print(f"{ln.name:14s} shapes differ {tuple(a.shape)} {tuple(b.shape)}")
        continue
    rel = float((a - b).norm() / (b.norm() + 1e-20))
    print(f"{ln.name:14s} {ln.type_id:16s} out rel diff {rel:.3e}   |ref| {float(b.norm()):.3e}")
for ln, lr in zip(reversed(nat.train_net.layers), reversed(ref.train_net.layers)):
    if ln.grad is None or lr.grad is None or ln.grad.shape != lr.grad.shape:
        continue
    a, b = ln.grad.float(), lr.grad.float()
    print(f"{ln.name:14s} {ln.type_id:16s} grad rel diff {float((a - b).norm() / (b.norm() + 1e-20)):.3e}   |ref| {float(b.norm()):.3e}")
for a, b in zip(nat.train_net.params, ref.train_net.params):
    ga, gb = a.grad.float().flatten(), b.grad.float().flatten()
    print(f"grad {a.name:10s} {str(tuple(a.shape)):16s} rel diff {float((ga - gb).norm() / (gb.norm() + 1e-20)):.3e}")
print("metric", nat.train_net.losslayers[0].metric.tolist(), ref.train_net.losslayers[0].metric.tolist())
