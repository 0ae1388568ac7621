"""Multi-GPU check of asynchronous worker groups with bounded staleness (run under torch.distributed.run).

Every rank trains with local updates + EASGD (or RandomSync) against the peer-memory server shards; rank 1 is
artificially slowed down (fault injection: extra host sleep per step).  Checks: (1) the fast group never runs more
than `staleness` steps ahead of the slow one (sampled through the device progress counters), (2) the loss falls,
(3) the groups' weights stay close to the center (elastic force works).
"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="mnist_mlp")
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--groups", type=int, default=1, help="server groups (each keeps its own sharded center)")
    ap.add_argument("--sync-interval", type=int, default=2, help="worker exchanges between neighbour-group averagings")
    ap.add_argument("--staleness", type=int, default=3)
    ap.add_argument("--steps", type=int, default=60)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from singa_b200 import device
    from singa_b200.cluster import Cluster
    from singa_b200.models import configs
    from singa_b200.proto import schema as S
    from singa_b200.worker import Worker
    device.set_context(f"cuda:{local}", "native", args.dtype)
    ok = True
    for param_type in ("Elastic", "RandomSync"):
        model = configs.mnist_mlp(batch=128, hidden=(256, 128)) if args.config == "mnist_mlp" \
            else configs.CONFIGS[args.config](batch=32)
        model.display_frequency = 0
        model.prefetch = False
        u = model.updater
        u.base_learning_rate, u.sync_frequency, u.warmup_steps, u.moving_rate = 0.05, 2, 2, 0.5
        u.param_type = param_type
        cp = S.ClusterProto()
        cp.nworkers = cp.nservers = world
        cp.workspace = "/tmp/singa_b200_async"
        cp.staleness = args.staleness
        cp.nserver_groups, cp.server_sync_interval = args.groups, args.sync_interval
        Cluster.reset()
        cluster = Cluster.get(cp, "", rank)
        w = Worker(cluster, backend="async")
        w.build(model)
        w.pm.sample_ratio = 0.5
        ex, be = w.executor, w.pm.backend
        steps, worst_lead = args.steps, 0
        first = last = None
        for step in range(steps):
            ex.ticks += 1
            ex.train_one_batch(step)
            if rank == 1:
                time.sleep(0.003)                       # the slow group
            if step % 5 == 4:
                prog = be.progress()
                worst_lead = max(worst_lead, max(prog) - min(prog))
            if step == 9 or step == steps - 1:
                torch.cuda.synchronize()
                m = w.train_net.losslayers[0].metric
                val = float(m[0]) / (10 if step == 9 else steps - 10)
                m.zero_()
                if step == 9:
                    first = val
                else:
                    last = val
        torch.cuda.synchronize()
        dist.barrier()
        dev_from_center = 0.0
        for o, (a, b) in enumerate(be.shards):
            if b > a:
                c = be.center_peer[o][a:b]
                dev_from_center = max(dev_from_center,
                                      float((w.pm.arena.data[a:b] - c).norm() / (c.norm() + 1e-12)))
        t = torch.tensor([worst_lead, dev_from_center, first, last], device="cuda", dtype=torch.float64)
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        if rank == 0:
            lead = max(float(x[0]) for x in allt)
            devc = max(float(x[1]) for x in allt)
            f0, l0 = float(allt[0][2]), float(allt[0][3])
            # the MLP must learn in these few steps; the conv nets start from ~1e-4 filters and barely move in 60 steps,
            # so for them the check is that the exchange keeps the loss where it is (no blow-up, no NaN)
            learns = l0 < f0 if args.config == "mnist_mlp" else (l0 == l0 and l0 < f0 + 0.01)
            good = lead <= cp.staleness + 1 and learns and devc < 0.5 and be.nsync > 0 and \
                (args.groups == 1 or be.ngroup_sync > 0) and (ex._variant_graphs or not ex.use_graph)
            ok = ok and good
            print(f"{'PASS' if good else 'FAIL'} async {param_type} [{args.config} {args.dtype}]: world={world} "
                  f"server groups={args.groups} staleness={cp.staleness} max observed lead={lead:.0f} steps, "
                  f"loss {f0:.4f} -> {l0:.4f}, max relative distance to the group's center {devc:.3f}, "
                  f"worker exchanges={be.nsync}, neighbour-group averagings={be.ngroup_sync}, "
                  f"graph variants captured={sorted(ex._variant_graphs)}")
        w.pm.finish()
        dist.barrier()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    sys.stdout.flush()
    os._exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
