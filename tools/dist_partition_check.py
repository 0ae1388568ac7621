"""Cross-rank execution of a partitioned net over peer-memory bridges (run under torch.distributed.run, 2 ranks).

A small conv net is partitioned for a group of 2 with a HYBRID plan — convolution layer-partitioned (filters split),
inner product and loss data-partitioned (batch split) — so the graph holds Slice, Concate and Bridge layers
(src/worker/neuralnet.cc:198-323).  Location i runs on rank i; activations and gradients cross ranks through
``PeerBridgeTransport`` (peer copies + device flags).  Checks, against the unpartitioned net on the same records with
the same parameters: the loss, and every local partition's parameter gradient; then the same step replayed from a CUDA
graph."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))


def walk(net, bridge, is_local):
    for l in net.layers:
        if not is_local(l) or l.is_datalayer:
            continue
        if l.is_bridgedstlayer:
            bridge.recv_data(l)
        l.compute_feature(True)
        if l.is_bridgesrclayer:
            bridge.send_data(l)
    net.zero_shared_grads()
    for l in reversed(net.layers):
        if not is_local(l) or l.is_datalayer or l.is_parserlayer:
            continue
        if l.is_bridgesrclayer:
            bridge.recv_grad(l)
        l.compute_gradient()
        if l.is_bridgedstlayer:
            bridge.send_grad(l)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    assert world == 2, "this check is written for a group of 2 ranks"
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from singa_b200 import device
    from singa_b200.device import set_step_state
    from singa_b200.neuralnet import NeuralNet
    from singa_b200.parallel.bridge import PeerBridgeTransport
    from singa_b200.proto import schema as S
    from test_config_and_graph import convnet
    dtype = sys.argv[1] if len(sys.argv) > 1 else "fp32"
    device.set_context(f"cuda:{local}", "native", dtype)
    import torch.distributed as d2
    real_init = d2.is_initialized
    d2.is_initialized = lambda: False        # both nets read the SAME records on both ranks (no per-rank data split)
    try:
        base = NeuralNet(convnet(), fuse=False)
        p2 = convnet()
        p2.partition_type = S.kLayerPartition                 # conv1 (and everything by default): filters split
        p2.layer[-2].partition_type = S.kDataPartition        # fc1: batch split
        p2.layer[-1].partition_type = S.kDataPartition        # loss: batch split
        part = NeuralNet(p2, 2, fuse=False)
    finally:
        d2.is_initialized = real_init
    from singa_b200.arena import ParamArena
    arenas = []
    for net in (base, part):        # GEMM operands are views of a parameter arena (normally made by the ParamManager)
        ar = ParamArena([p for p in net.params if p.is_owner])
        ar.adopt(net.params)
        arenas.append(ar)
    g = torch.Generator().manual_seed(5)
    for p in base.params:
        p.data.copy_((torch.randn(p.shape, generator=g) * 0.1).to(p.data.device))
    # parameters of the partitions: conv1 sliced along the filters, fc1 shared (data partition)
    cw, cb = base.name2layer("conv1").get_params()
    off = 0
    for i in range(2):
        w, b = part.name2layer(f"conv1-{i:02d}").get_params()
        n = b.shape[0]
        w.data.copy_(cw.data[off:off + n]); b.data.copy_(cb.data[off:off + n])
        off += n
    fw, fb = base.name2layer("fc1").get_params()
    w, b = part.name2layer("fc1-00").get_params()
    w.data.copy_(fw.data); b.data.copy_(fb.data)
    for ar in arenas:
        ar.sync_mirror()            # refresh mirrors / packed convolution filters from the values just written
    bridge = PeerBridgeTransport(part, rank, world, torch.device("cuda", local))
    is_local = bridge.is_local
    kinds = sorted({l.type_id for l in part.layers if l.is_connector})
    # the unpartitioned reference, entirely on this rank
    set_step_state(0, 0.0)
    for d in base.datalayers:
        d.compute_feature(True)
    for d in part.datalayers:
        d.compute_feature(True)
    walk(base, BridgeNull(), lambda l: True)
    torch.cuda.synchronize()
    ok = True
    results = []
    for mode in ("eager", "graph"):
        for p in part.params:
            if p.grad is not None:
                p.grad.zero_()
        for l in part.losslayers:
            l.metric.zero_()
        dist.barrier()
        if mode == "eager":
            set_step_state(1, 0.0)
            walk(part, bridge, is_local)
        else:
            gr = torch.cuda.CUDAGraph()
            set_step_state(2, 0.0)
            torch.cuda.synchronize(); dist.barrier()
            with torch.cuda.graph(gr):
                walk(part, bridge, is_local)
            for p in part.params:
                if p.grad is not None:
                    p.grad.zero_()
            for l in part.losslayers:
                l.metric.zero_()
            set_step_state(3, 0.0)
            torch.cuda.synchronize(); dist.barrier()
            gr.replay()
        torch.cuda.synchronize()
        dist.barrier()
        # loss: mean over the data partitions (each on its rank)
        loss = torch.zeros(1, device="cuda")
        for l in part.losslayers:
            if is_local(l):
                loss += l.metric[0]
        dist.all_reduce(loss)
        loss_p = float(loss) / len(part.losslayers)
        loss_b = float(base.losslayers[0].metric[0])
        worst = 0.0
        off = 0
        for i in range(2):
            lw = part.name2layer(f"conv1-{i:02d}")
            w, b = lw.get_params()
            n = b.shape[0]
            if is_local(lw):
                ref = cw.grad[off:off + n]
                # each loss partition normalises by ITS batch (N / 2), so every gradient below the data-partitioned
                # layers is 2 x the full-batch mean gradient — the same factor the shared fc1 gradient carries
                rel = float((w.grad / 2 - ref).norm() / (ref.norm() + 1e-12))
                if os.environ.get("SINGA_B200_CHECK_VERBOSE"):
                    print(f"[rank {rank}] {mode} conv1-{i:02d} weight grad rel {rel:.3e} |g| {float(w.grad.norm()):.3e} |ref| {float(ref.norm()):.3e}")
                worst = max(worst, rel)
            off += n
        # fc1 is data-partitioned with a shared parameter: the two ranks hold partial sums of the gradient
        wg = torch.zeros_like(fw.grad)
        for i in range(2):
            lw = part.name2layer(f"fc1-{i:02d}")
            if is_local(lw):
                wg += lw.get_params()[0].grad
        dist.all_reduce(wg)
        rel = float((wg / 2 - fw.grad).norm() / (fw.grad.norm() + 1e-12))
        if os.environ.get("SINGA_B200_CHECK_VERBOSE") and rank == 0:
            print(f"{mode} fc1 weight grad rel {rel:.3e} |sum/2| {float(wg.norm()) / 2:.3e} |ref| {float(fw.grad.norm()):.3e}")
        worst = max(worst, rel)
        t = torch.tensor([worst], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tol = 3e-2 if dtype == "bf16" else 3e-3
        good = abs(loss_p - loss_b) <= tol * max(1.0, abs(loss_b)) and float(t) < 10 * tol
        ok = ok and good
        results.append(f"{'PASS' if good else 'FAIL'} hybrid partition over peer-memory bridges [{mode}, {dtype}]: "
                       f"{len(bridge.edges)} cross-rank bridges, connectors {kinds}; loss {loss_p:.5f} vs unpartitioned "
                       f"{loss_b:.5f}; worst relative gradient difference {float(t):.2e}")
    if rank == 0:
        print("\n".join(results))
    dist.barrier()
    sys.stdout.flush()
    os._exit(0 if ok else 1)


class BridgeNull:
    def recv_data(self, l): pass
    def send_data(self, l): pass
    def recv_grad(self, l): pass
    def send_grad(self, l): pass


if __name__ == "__main__":
    main()
