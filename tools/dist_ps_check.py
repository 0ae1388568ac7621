"""Multi-GPU check of the parameter-server data plane (run under torch.distributed.run, 1 process per GPU):

  W ranks x batch B with push_pull in {fused, nccl}  ==  1 process x batch W*B   (same records, same init)

Every rank trains a small net for a few steps (the last ones replayed from the captured CUDA graph, rotated
schedule included); rank 0 then trains the equivalent single-process job and compares the weights every rank ended up
with (the operand mirrors, or the fp32 values themselves where no mirror exists).  Prints PASS/FAIL lines; exit code
0 only if all pass.

    torchrun ... tools/dist_ps_check.py [config] [--dtype fp32|bf16] [--steps N] [--no-graph]
                                        [--delay-rank R --delay-ms T]   (stress: one rank always late)
Environment: SINGA_B200_NVLS=0/1 selects peer loads / in-switch reduction for the fused push.
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(push_pull, world, rank, steps, batch, hidden, config="mnist_mlp", graph=True, delay=(-1, 0.0)):
    from singa_b200 import device
    from singa_b200.cluster import Cluster
    from singa_b200.models import configs
    from singa_b200.proto import schema as S
    from singa_b200.worker import Worker
    model = configs.mnist_mlp(batch=batch, hidden=hidden) if config == "mnist_mlp" else configs.CONFIGS[config](batch=batch)
    model.display_frequency = 0
    model.prefetch = False
    model.updater.base_learning_rate = 0.02
    model.updater.momentum = 0.9
    if config == "alexnet":
        # W ranks x B equals 1 process x W*B only for deterministic layers: the dropout mask is a function of the
        # element index inside a rank's own batch, so the two jobs would draw different masks; and 0.02 with momentum
        # diverges on this net within a few steps, which turns rounding differences into O(1) weight differences
        model.updater.base_learning_rate = 0.002
        for l in model.neuralnet.layer:
            if l.type == "kDropout":
                l.dropout_param.dropout_ratio = 0.0
    cluster = None
    Cluster.reset()      # the single-process reference must not inherit the worker-group layout of the previous run
    if push_pull is not None:
        cp = S.ClusterProto()
        cp.nworkers = cp.nservers = world
        cp.workspace = "/tmp/singa_b200_check"
        cp.synchronous = True
        Cluster.reset()
        cluster = Cluster.get(cp, "", rank)
    w = Worker(cluster, backend=push_pull, use_graph=graph)
    w.build(model)
    ex = w.executor
    for step in range(steps):
        ex.ticks += 1
        if push_pull is not None and rank == delay[0] and delay[1] > 0:
            # stress: this rank's whole step (its gradients, ready flags and pushes) starts late, so every other rank's
            # push kernel sits in the rendezvous and every consumer of its shard in the version-flag acquire
            torch.cuda._sleep(int(delay[1] * 1.9e6))
        ex.train_one_batch(step)
    torch.cuda.synchronize()
    if push_pull is not None:
        dist.barrier()
    # compare what every rank computes with: the parameter server's mirrors, or the fp32 values where there is none
    # (repacked conv operands are re-derived from them at the next forward)
    out = {p.name + str(p.id): (p.mirror16 if getattr(p, "mirror16", None) is not None else p.data).float().clone()
           for p in w.pm.arena.params}
    loss = float(w.train_net.losslayers[0].metric[0]) / steps
    w.pm.finish()
    return out, loss


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ap = argparse.ArgumentParser()
    ap.add_argument("config", nargs="?", default="mnist_mlp")
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--delay-rank", type=int, default=-1, help="stress: delay every step of this rank ...")
    ap.add_argument("--delay-ms", type=float, default=0.0, help="... by this many milliseconds of device time")
    args = ap.parse_args()
    from singa_b200 import device
    device.set_context(f"cuda:{local}", "native", args.dtype)
    steps, batch, hidden = args.steps, 64, (256, 136)
    config = args.config
    graph = not args.no_graph
    if config != "mnist_mlp":
        batch = 16
    ok = True
    results = {}
    for pp in ("nccl", "fused"):
        results[pp] = run(pp, world, rank, steps, batch, hidden, config, graph, (args.delay_rank, args.delay_ms))
        dist.barrier()
    # every rank must hold identical mirrors
    for pp, (mir, loss) in results.items():
        for k, v in mir.items():
            ref = v.clone()
            dist.broadcast(ref, 0)
            same = bool(torch.equal(ref, v))
            if not same:
                ok = False
                print(f"[rank {rank}] FAIL {pp}: mirror {k} differs from rank 0 (max {float((ref - v).abs().max())})")
    if rank == 0:
        # single-process reference with batch world*B reading the same records
        import torch.distributed as d2
        saved = d2.is_initialized
        d2.is_initialized = lambda: False     # make the data layer / arena behave as a single process
        os.environ["SINGA_B200_BIAS_MIRROR"] = "1"   # same bias precision as the sharded exchanges (bf16 mirror)
        try:
            ref, ref_loss = run(None, 1, 0, steps, batch * world, hidden, config, graph)
        finally:
            d2.is_initialized = saved
        for pp, (mir, loss) in results.items():
            worst, worst_k = 0.0, ""
            for k, v in mir.items():
                # relative to the parameter's norm, floored at an RMS of 1e-2 so a bias vector that is still ~0
                # after a few steps does not turn bf16 rounding into a large "relative" error
                denom = max(float(ref[k].norm()), 1e-2 * ref[k].numel() ** 0.5)
                rel = float((v - ref[k]).norm()) / denom
                if os.environ.get("SINGA_B200_CHECK_VERBOSE"):
                    print(f"   {pp} {k}: rel {rel:.3e}  |ref| {float(ref[k].norm()):.3e}  n={ref[k].numel()}")
                if rel > worst:
                    worst, worst_k = rel, k
            # conv nets: split-K weight gradients are summed in a different order per run and per rank count, and the
            # small bias vectors of the first layers carry the difference of every layer above them
            lim = (2e-2 if config == "mnist_mlp" else 2e-1) if args.dtype == "bf16" else \
                (2e-3 if config == "mnist_mlp" else 5e-2)
            status = "PASS" if worst < lim else "FAIL"
            ok = ok and status == "PASS"
            stress = f" rank {args.delay_rank} late by {args.delay_ms} ms" if args.delay_rank >= 0 else ""
            print(f"{status} {pp} [{config} {args.dtype} graph={graph} nvls={os.environ.get('SINGA_B200_NVLS', 'auto')}{stress}]: world={world} worst relative weight difference vs 1 process x batch "
                  f"{batch * world}: {worst:.3e} ({worst_k}); mean loss {loss:.4f} vs {ref_loss:.4f}")
        a, b = results["nccl"][0], results["fused"][0]
        worst = max(float((a[k] - b[k]).abs().max()) for k in a)
        print(f"INFO fused vs nccl max abs mirror difference: {worst:.3e}")
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
