"""Multi-GPU checkpoint check (run under torch.distributed.run): with a sharded parameter server the fp32 master copy
and the optimizer state are spread over the ranks; every owner writes its own shard file (no gather), and a fresh job
— with the same or a different exchange backend — resumes from the set of files.  Reports the save / load times.

    torchrun ... tools/dist_ckpt_check.py [config] [--dtype fp32|bf16] [--batch N]
"""
import argparse
import os
import shutil
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def build(push_pull, world, rank, folder, config, batch):
    from singa_b200.cluster import Cluster
    from singa_b200.models import configs
    from singa_b200.proto import schema as S
    from singa_b200.worker import Worker
    if config == "mnist_mlp":
        model = configs.mnist_mlp(batch=batch or 64, hidden=(256, 136))
    else:
        model = configs.CONFIGS[config](**({"batch": batch} if batch else {}))
    model.display_frequency = 0
    model.prefetch = False
    model.updater.momentum = 0.9
    cp = S.ClusterProto()
    cp.nworkers = cp.nservers = world
    cp.workspace = folder
    cp.synchronous = True
    Cluster.reset()
    w = Worker(Cluster.get(cp, "", rank), backend=push_pull, use_graph=False)
    w.build(model)
    return w, model


def weights_everyone_sees(p):
    return (p.mirror16 if getattr(p, "mirror16", None) is not None else p.data).float().clone()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", nargs="?", default="mnist_mlp")
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--batch", type=int, default=0)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from singa_b200 import checkpoint, device
    device.set_context(f"cuda:{local}", "native", args.dtype)
    folder = "/tmp/singa_b200_ckpt_check"
    if rank == 0:
        shutil.rmtree(folder, ignore_errors=True)
    dist.barrier()
    ok = True
    for pp, pp_resume in (("fused", "fused"), ("nccl", "fused")):
        w, model = build(pp, world, rank, folder, args.config, args.batch)
        for step in range(3):
            w.executor.ticks += 1
            w.executor.train_one_batch(step)
        torch.cuda.synchronize()
        dist.barrier()
        sub = os.path.join(folder, pp)
        t0 = time.monotonic()
        path = checkpoint.save(sub, w.train_net, w.pm, 3, model.name)        # every owner writes its shard
        t_save = time.monotonic() - t0
        seen = {p.id: weights_everyone_sees(p) for p in w.pm.arena.params}
        nparams = sum(p.size for p in w.pm.arena.params)
        dist.barrier()
        nbytes = sum(os.path.getsize(f) for f in checkpoint.step_files(path))
        w.pm.finish()
        del w
        torch.cuda.empty_cache()
        w2, _ = build(pp_resume, world, rank, folder, args.config, args.batch)
        t0 = time.monotonic()
        step = checkpoint.load(checkpoint.latest(sub), w2.train_net, w2.pm)
        t_load = time.monotonic() - t0
        worst = 0.0
        for p in w2.pm.arena.params:
            ref = seen[p.id]
            got = p.data.float() if ref.dtype == p.data.dtype and getattr(p, "mirror16", None) is None else \
                weights_everyone_sees(p)
            worst = max(worst, float((got - ref).abs().max()))
        h2 = sum(float(p.history.norm()) for p in w2.pm.arena.params)
        good = step == 3 and worst == 0.0 and h2 > 0 and os.path.basename(path).startswith("step3.shard")
        t = torch.tensor([1.0 if good else 0.0, t_save, t_load], device="cuda")
        dist.all_reduce(t[:1], op=dist.ReduceOp.MIN)
        dist.all_reduce(t[1:], op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"{'PASS' if int(t[0]) else 'FAIL'} checkpoint {pp} -> resume {pp_resume} [{args.config} {args.dtype}]: "
                  f"world={world}, {nparams / 1e6:.1f} M parameters, {world} shard files, {nbytes / 2**20:.0f} MiB total; "
                  f"save {float(t[1]):.2f} s, load {float(t[2]):.2f} s (max over ranks); "
                  f"max |resumed weights - weights at save| = {worst:.3e}")
        ok = ok and bool(int(t[0]))
        w2.pm.finish()
        del w2
        torch.cuda.empty_cache()
        dist.barrier()
    sys.stdout.flush()
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
