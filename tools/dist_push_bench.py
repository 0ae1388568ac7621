"""Fused push kernel in isolation on N GPUs: per bucket of the MNIST MLP (or --config), time back-to-back pushes with
CUDA events on the comm stream (max over ranks) and report the NVLink bytes per direction against the measured
770 GB/s peer-copy rate.  Launch with torch.distributed.run; knobs through the usual SINGA_B200_* variables."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="mnist_mlp")
ap.add_argument("--dtype", default="fp32")
ap.add_argument("--iters", type=int, default=30)
args = ap.parse_args()

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))

from singa_b200 import device  # noqa: E402
from singa_b200.cluster import Cluster  # noqa: E402
from singa_b200.models import configs  # noqa: E402
from singa_b200.proto import schema as S  # noqa: E402
from singa_b200.worker import Worker  # noqa: E402

device.set_context(f"cuda:{local}", "native", args.dtype)
model = configs.CONFIGS[args.config]()
model.display_frequency = 0
cp = S.ClusterProto()
cp.nworkers, cp.nservers, cp.synchronous, cp.push_pull = world, world, True, "fused"
cp.workspace = "/tmp/singa_b200_pushbench"
Cluster.reset()
w = Worker(Cluster.get(cp, "", rank), backend="fused", use_graph=False)
w.build(model)
be = w.pm.backend
w.pm.begin_step(0)
mirror_bytes = 0 if w.pm.arena.data16 is None else w.pm.arena.data16.element_size()
rows = []
for b, layers in enumerate(be.buckets):
    trig = layers[-1]
    first, n = be.bucket_mine[b]
    params = n * 1024 * world
    for _ in range(5):
        be.update_layer(trig, 0, None)
    be.end_step(0)
    torch.cuda.synchronize(); dist.barrier()
    comm = be.comms[b % len(be.comms)]
    # the pushes are captured in a CUDA graph (as in the training step) so that host-side launch cost is excluded
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(args.iters):
            be.update_layer(trig, 0, None)
        be.end_step(0)
    g.replay()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / args.iters * 1e3], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    us = float(t)
    # per rank and direction: gradients of my piece pulled from the other ranks + my fresh piece written to them
    shard = params / world
    inbound = shard * 4 * (world - 1) if not be.use_nvls else shard * 4
    outbound = shard * mirror_bytes * (world - 1) if not be.use_nvls else shard * mirror_bytes
    served = shard * 4 * (world - 1)          # my gradients read by the other owners (or by the switch)
    per_dir = max(inbound + shard * mirror_bytes * (world - 1) * (1 if be.use_nvls else 0), outbound + served)
    # phase stamps of one more push (eager, after a barrier so both ranks start together)
    be.trace_buf = torch.zeros(8, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize(); dist.barrier()
    be.update_layer(trig, 0, None)
    be.end_step(0)
    torch.cuda.synchronize()
    st = be.trace_buf.tolist()
    be.trace_buf = None
    phases = [round((st[i] - st[0]) / 1e3, 1) for i in range(1, 7)]
    rows.append({"phases_us": phases, "bucket": [l.name for l in layers], "params": int(params), "us_per_push": round(us, 1),
                 "link_bytes_per_direction": int(per_dir), "GBps_per_direction": round(per_dir / us / 1e3, 1),
                 "link_floor_us_at_770GBps": round(per_dir / 770e3, 1)})
if rank == 0:
    print(json.dumps({"world": world, "dtype": args.dtype, "nvls": be.use_nvls, "ctas": be.max_ctas,
                      "pushes": [[r["bucket"][0], r["params"], r["us_per_push"], r["GBps_per_direction"],
                                  r["link_floor_us_at_770GBps"], r["phases_us"]] for r in rows]}))
dist.barrier()
torch.cuda.synchronize()
os._exit(0)
