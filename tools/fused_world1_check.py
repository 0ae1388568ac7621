"""One-process check of the fused parameter-server path (world = 1: the rank is its own server shard): a few eager
steps, then CUDA-graph capture of forward + backward + push, then replays.  Catches capture-time stream errors of the
exchange on ONE GPU (e.g. joining a comm stream that took no work).  usage: fused_world1_check.py [config]"""
import os, sys, traceback
os.environ.setdefault("MASTER_ADDR","127.0.0.1"); os.environ.setdefault("MASTER_PORT", str(29533 + os.getpid() % 100))
os.environ.setdefault("RANK","0"); os.environ.setdefault("WORLD_SIZE","1")
import torch, torch.distributed as dist
torch.cuda.set_device(0)
dist.init_process_group("nccl", device_id=torch.device("cuda",0))
from singa_b200 import device
from singa_b200.cluster import Cluster
from singa_b200.models import configs
from singa_b200.proto import schema as S
from singa_b200.worker import Worker
device.set_context("cuda:0","native")
cfg = sys.argv[1] if len(sys.argv)>1 else "cifar10"
model = configs.CONFIGS[cfg]()
model.display_frequency = 0; model.test_frequency = 0; model.train_steps = 1<<30; model.prefetch=False
cp = S.ClusterProto(); cp.nworkers, cp.nservers = 1, 1; cp.workspace="/tmp/ws_dbg"; cp.synchronous=True; cp.push_pull="fused"
Cluster.reset(); cluster = Cluster.get(cp, "", 0)
w = Worker(cluster, backend="fused", use_graph=True); w.build(model)
ex, net = w.executor, w.train_net
net.datalayers[0].compute_feature(True)
for p in net.parserlayers:
    p.prefetching(True); p.compute_feature(True)
import singa_b200.worker as W
orig = W.log.error
for step in range(4):
    ex.ticks += 1; w.pm.begin_step(step)
    for p in net.parserlayers: p.reparse(True)
    ex.forward(net, step, True, skip_inputs=True); ex.backward(net, step)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(g):
        ex.forward(net, 4, True, skip_inputs=True)
        ex.backward(net, 4)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    loss = float(net.losslayers[0].metric[0])
    assert loss == loss, "loss is NaN"
    print(f"PASS fused world=1 [{cfg}]: eager steps, graph capture and 3 replays; buckets={len(w.pm.backend.buckets)}")
    rc = 0
except Exception:
    traceback.print_exc()
    rc = 1
sys.stdout.flush()
os._exit(rc)
