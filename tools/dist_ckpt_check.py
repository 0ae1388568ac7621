"""Multi-GPU checkpoint check (run under torch.distributed.run): with a sharded parameter server the fp32 master copy
is spread over the ranks; `checkpoint.save` must gather it, and the saved values must round to the bf16 mirrors every
rank computes with.  A fresh job then resumes from the file."""
import os
import shutil
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def build(push_pull, world, rank, folder):
    from singa_b200.cluster import Cluster
    from singa_b200.models import configs
    from singa_b200.proto import schema as S
    from singa_b200.worker import Worker
    model = configs.mnist_mlp(batch=64, hidden=(256, 136))
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


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from singa_b200 import checkpoint, device
    device.set_context(f"cuda:{local}", "native")
    folder = "/tmp/singa_b200_ckpt_check"
    if rank == 0:
        shutil.rmtree(folder, ignore_errors=True)
    dist.barrier()
    ok = True
    for pp in ("fused", "nccl"):
        w, model = build(pp, world, rank, folder)
        for step in range(5):
            w.executor.ticks += 1
            w.executor.train_one_batch(step)
        torch.cuda.synchronize()
        dist.barrier()
        path = checkpoint.save(os.path.join(folder, pp), w.train_net, w.pm, 5, model.name)   # collective
        mirrors = {p.id: p.mirror16.float().clone() for p in w.pm.arena.params}
        hist_norm = sum(float(p.history.norm()) for p in w.pm.arena.params)
        dist.barrier()
        w.pm.finish()
        w2, _ = build(pp, world, rank, folder)
        step = checkpoint.load(os.path.join(folder, pp, "step5.bin"), w2.train_net, w2.pm)
        worst = 0.0
        for p in w2.pm.arena.params:
            worst = max(worst, float((p.data.to(torch.bfloat16).float() - mirrors[p.id]).abs().max()))
        h2 = sum(float(p.history.norm()) for p in w2.pm.arena.params)
        good = step == 5 and worst == 0.0 and h2 > 0 and (rank != 0 or path is not None)
        flag = torch.tensor([1 if good else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(f"{'PASS' if int(flag) else 'FAIL'} checkpoint {pp}: world={world} resumed step {step}, "
                  f"max |bf16(saved master) - mirror| = {worst:.3e}, history norm saved {hist_norm:.3f} / loaded {h2:.3f}")
        ok = ok and bool(int(flag))
        w2.pm.finish()
        dist.barrier()
    sys.stdout.flush()
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
