#!/usr/bin/env python
"""Headline benchmark: training throughput (images/sec) of the reference's flagship job — the examples/mnist
MLP (784-2500-2000-1500-1000-500-10, tanh, batch 1000 per worker group, SGD) — on N B200 GPUs of one node,
one worker group (= one process) per GPU, synchronous push/pull through the sharded parameter server.

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 200 --warmup 20
    python bench.py --impl reference ...     # the unmodified reference (unavailable offline: see DESIGN.md)

Prints ONE JSON line on rank 0 (contract in the task statement).  Two measurements:
  * ``value``      device-timed (CUDA events, max over ranks): K steps of parse + forward + backward + fused
                   update (+ push/pull for N > 1) with the uint8 batches already staged on the device;
  * ``e2e.value``  the same K steps through the public API (Executor.train_one_batch) with, EVERY step, the
                   batch assembled in pinned host memory, copied host->device, and the loss read back.
Other configs of BASELINE.json: ``--config {mnist_mlp,lenet,cifar10,alexnet,wide_mlp}``.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

L2_BYTES = 126 * 1024 * 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch", "torch-layers"],
                    help="ours = this framework's native path; reference = the unmodified reference (unavailable offline); "
                         "torch = library baseline: torch.nn + autograd + cuBLAS/cuDNN + bucketed NCCL all-reduce under one "
                         "CUDA graph (baseline/torch_baseline.py); torch-layers = this framework's layer classes on their "
                         "plain-PyTorch backend")
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16", "fp32-strict"],
                    help="compute precision. fp32 (default; the reference computes in fp32): fp32 storage, tf32 tensor-core "
                         "products, fp32 accumulation - ours = tcgen05 kind::tf32, torch = cuBLAS/cuDNN with TF32 allowed. "
                         "bf16: bf16 operands, fp32 accumulation - ours = kind::f16, torch = autocast(bf16). "
                         "fp32-strict (torch only): TF32 off, plain fp32 FMA kernels")
    ap.add_argument("--config", default="mnist_mlp")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (0 = the config's own)")
    ap.add_argument("--push-pull", default=os.environ.get("SINGA_B200_PUSH_PULL", "fused"),
                    choices=["fused", "nccl", "async"],
                    help="parameter exchange between the ranks: fused = one kernel per bucket over peer memory (product), "
                         "nccl = bucketed NCCL reduce-scatter / all-gather (library baseline), async = asynchronous worker "
                         "groups (local updates + EASGD exchange with peer-memory server shards, bounded staleness)")
    ap.add_argument("--staleness", type=int, default=2, help="--push-pull async: steps a group may run ahead")
    ap.add_argument("--sync-frequency", type=int, default=0, help="--push-pull async: steps between exchanges (0 = config's)")
    ap.add_argument("--server-groups", type=int, default=1, help="--push-pull async: server groups (neighbour averaging)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--expect-loss-window", default="",
                    help="'loss,metric' of another arm's e2e.final_loss_sum_window for the same config / steps / GPUs "
                         "(e.g. --push-pull nccl): the run fails (exit 3) when its own window differs by more than "
                         "--loss-tol — tools/run_matrix.sh checks every fused run against the NCCL exchange this way")
    ap.add_argument("--loss-tol", type=float, default=0.0, help="relative tolerance (0 = 2e-2 for bf16, 5e-3 otherwise)")
    return ap.parse_args()


class ClockSampler:
    """Samples SM clocks and throttle reasons of one GPU DURING the timed region: an in-process NVML poll at ~1 kHz
    (the timed region of the small configs lasts tens of milliseconds - an `nvidia-smi -lms` child process delivers
    one or two samples in that time, or none), with `nvidia-smi` as the fallback when NVML cannot be loaded."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []
        self.samples = []          # (sm_mhz, reasons bitmask) from NVML
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None

    def _physical_index(self) -> int:
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[self.gpu])
            except (ValueError, IndexError):
                pass
        return self.gpu

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index())
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            self._nvml = pynvml

            def poll():
                while not self._stop.is_set():
                    try:
                        self.samples.append((float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)),
                                             int(get_reasons(h))))
                    except Exception:  # noqa: BLE001
                        pass
                    time.sleep(0.001)
            self._thread = threading.Thread(target=poll, daemon=True)
            self._thread.start()
            return
        except Exception:  # noqa: BLE001 - no NVML binding / library: fall back to the nvidia-smi child process
            self._nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:  # noqa: BLE001 - nvidia-smi missing: report empty clocks
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self._nvml is not None:
            self._stop.set()
            self._thread.join(timeout=1)
            sm = sorted(s for s, _ in self.samples)
            mask = 0
            for _, r in self.samples:
                mask |= r
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(n for b, n in self.REASONS.items() if mask & b), "samples": len(sm),
                    "source": "nvml"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


def reference_arm(args):
    """The reference cannot be installed or built here: it has no setup.py/pyproject.toml (pip: "Neither
    'setup.py' nor 'pyproject.toml' found") and `make singa` needs protoc, glog, gflags, czmq, lmdb, OpenCV,
    cblas, none of which exist on the box (no network).  Recorded in DESIGN.md."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    why = "reference is a C++/Makefile project (no setup.py/pyproject.toml; make needs protoc, glog, gflags, " \
          "czmq, lmdb, OpenCV, cblas - absent offline); baseline/_ref is empty"
    if os.path.isdir(ref) and os.listdir(ref):
        why = "baseline/_ref exists but holds no runnable reference entry point"
    print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


def torch_arm(args, rank, local_rank, world):
    """Library baseline (see baseline/torch_baseline.py): same metric, config, timing rules and JSON line."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import torch_baseline as TB
    r = TB.run(args, rank, local_rank, world, ClockSampler)
    global_batch = r["per_gpu_batch"] * world
    if rank == 0:
        out = {
            "metric": "images/sec (device-timed, max over ranks)",
            "value": round(global_batch * args.steps / (r["ms_total"] / 1e3), 1), "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 5), "ms_per_step": round(r["ms_total"] / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp32": "fp32 (tf32 tensor-core products, fp32 accumulate)", "bf16": "bf16 (autocast)",
                      "fp32-strict": "fp32 (no tensor cores)"}[args.dtype],
            "data": "synthetic", "impl": "torch",
            "config": {"model": args.config, "global_batch": global_batch, "per_gpu_batch": r["per_gpu_batch"],
                       "seq_len": None,
                       "parallelism": f"dp{world}: torch.nn + autograd, {r['buckets']} gradient buckets all-reduced by NCCL on "
                                      "a side stream during backward, foreach SGD",
                       "cuda_graph": r["graph"], "l2": "inputs + activations + parameters exceed L2 or are re-read per step; "
                                                       "no explicit flush (same as the product arm for this config)"},
            "clocks": r["clocks"],
            "e2e": {"value": round(global_batch * args.steps / (r["e2e_ms"] / 1e3), 1), "unit": "images/s",
                    "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                    "ms_per_step": round(r["e2e_ms"] / args.steps, 4), "final_loss": round(r["last_loss"], 4)},
            "gpu_launches": 0,
        }
        print(json.dumps(out))
    sys.stdout.flush()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} needs WORLD_SIZE={args.gpus} (launch with torch.distributed.run); "
                         f"got WORLD_SIZE={world}")
    assert torch.cuda.is_available(), "bench.py needs a GPU"
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if args.impl == "torch":
        return torch_arm(args, rank, local_rank, world)

    from singa_b200 import device
    from singa_b200.cluster import Cluster
    from singa_b200.models import configs
    from singa_b200.ops import native
    from singa_b200.proto import schema as S
    from singa_b200.worker import Worker

    backend_kind = "torch" if args.impl == "torch-layers" else "native"
    if args.impl == "ours" and args.dtype == "fp32-strict":
        raise SystemExit("--dtype fp32-strict is a library-arm setting (tensor cores off); ours runs fp32 (tf32) or bf16")
    device.set_context(f"cuda:{local_rank}", backend_kind, "bf16" if args.dtype == "bf16" else "fp32")
    if args.impl == "torch-layers":
        from singa_b200.ops import reference as R
        tf32 = args.dtype == "fp32"
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cudnn.benchmark = True
        R.AUTOCAST_BF16 = args.dtype == "bf16"

    kw = {}
    if args.batch:
        kw["batch"] = args.batch
    model = configs.CONFIGS[args.config](**kw)
    model.display_frequency = 0
    model.test_frequency = 0
    model.train_steps = 1 << 30
    model.prefetch = False          # bench drives the staging explicitly (see the two timed regions)

    cluster = None
    exchange = None
    if world > 1:
        cp = S.ClusterProto()
        cp.nworkers, cp.nservers = world, world        # one server shard co-located with every worker group
        cp.workspace = os.environ.get("SINGA_B200_WORKSPACE", "/tmp/singa_b200_bench")
        cp.synchronous = args.push_pull != "async"
        cp.push_pull = args.push_pull if args.push_pull != "async" else "fused"
        if args.push_pull == "async":
            cp.staleness = args.staleness
            cp.nserver_groups = args.server_groups
            model.updater.param_type = "Elastic"
            model.updater.warmup_steps = 0
            model.updater.sync_frequency = args.sync_frequency or max(1, model.updater.sync_frequency)
            if model.updater.moving_rate <= 0:
                model.updater.moving_rate = 0.5
        Cluster.reset()
        cluster = Cluster.get(cp, "", rank)
        exchange = args.push_pull
    w = Worker(cluster, backend=exchange, use_graph=not args.no_graph)
    w.build(model)
    ex, net = w.executor, w.train_net
    data_layer = net.datalayers[0]
    per_gpu_batch = data_layer.batchsize
    loss_layer = net.losslayers[0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------------------------------------------------------- device-timed region
    # batches staged on the device up front; each step: parse kernels + fwd + bwd + update (+ push/pull)
    def stage_all():
        data_layer.compute_feature(True)
        for p in net.parserlayers:
            p.prefetching(True)
            p.compute_feature(True)

    def step_device(step):
        ex.ticks += 1
        if ex.rotate:                   # rotated schedule: backward(step) + push || forward(step + 1) per replay
            for p in net.parserlayers:
                p.reparse_next(True)
            ex._run_rotated(step, feed=False)
            return
        w.pm.begin_step(step)
        for p in net.parserlayers:      # re-parse the device-resident staged batch (no host traffic)
            p.reparse(True)
        ex._run_compute(step)

    stage_all()
    step = 0
    for _ in range(max(args.warmup, 3)):
        step_device(step); step += 1
    barrier()
    ws_bytes = w.pm.arena.total * (4 + 4 + (2 if args.dtype == "bf16" else 0)) + sum(
        l.data.numel() * l.data.element_size() for l in net.layers if l.data is not None and l.data.is_cuda)
    flush = None
    if ws_bytes < L2_BYTES:
        flush = torch.empty(2 * L2_BYTES, dtype=torch.uint8, device="cuda")
    sampler = ClockSampler(local_rank)
    launches0 = native.launch_count
    if rank == 0:
        sampler.start()
    barrier()
    if flush is None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step_device(step); step += 1
        e1.record()
        barrier()
        ms_total = e0.elapsed_time(e1)
    else:
        evs = []
        for _ in range(args.steps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); step_device(step); step += 1; b.record()
            evs.append((a, b))
        barrier()
        ms_total = sum(a.elapsed_time(b) for a, b in evs)
    clocks = sampler.stop() if rank == 0 else None
    eager_launches = native.launch_count - launches0
    graph_launches = ex.graph_kernel_count * args.steps if ex._graph is not None else 0
    gpu_launches = eager_launches + graph_launches

    # optional kernel timeline of a few extra (untimed) steps: SINGA_B200_TRACE_OUT=<prefix> writes
    # <prefix>.rank<r>.txt with (start us, duration us, stream, kernel) per launch -- attribution only
    trace_out = os.environ.get("SINGA_B200_TRACE_OUT")
    if trace_out:
        from torch.profiler import ProfilerActivity, profile
        barrier()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                step_device(step); step += 1
            torch.cuda.synchronize()
        barrier()
        evs_ = sorted((e for e in prof.events() if e.device_type is not None and "cuda" in str(e.device_type).lower()),
                      key=lambda e: e.time_range.start)
        t0_ = evs_[0].time_range.start if evs_ else 0
        with open(f"{trace_out}.rank{rank}.txt", "w") as fh:
            for e in evs_:
                fh.write(f"{e.time_range.start - t0_:10.1f} {e.time_range.end - e.time_range.start:8.1f} "
                         f"s{getattr(e, 'device_resource_id', -1)} {e.name[:110]}\n")

    # ---------------------------------------------------------------- end-to-end region (public API)
    h2d = sum(p.h2d_bytes for p in net.parserlayers)
    d2h = loss_layer.metric.numel() * loss_layer.metric.element_size()
    model.prefetch = True
    ex.setup()                         # stage batch 0, start the prefetch thread
    for _ in range(3):
        ex.ticks += 1
        ex.train_one_batch_async(step).result(); step += 1
    barrier()
    t0 = time.perf_counter()
    last, pending = None, None
    for _ in range(args.steps):
        ex.ticks += 1
        # pinned host batch -> H2D -> parse -> fwd/bwd/update, + D2H copy of this step's metrics into a pinned slot
        h = ex.train_one_batch_async(step); step += 1
        if pending is not None:
            last = pending.result()              # host read of the PREVIOUS step's result: the GPU already runs step i
        pending = h
    last = pending.result()                      # ... and of the final step, inside the timed region
    barrier()
    e2e_s = time.perf_counter() - t0
    ex._join_prefetch()

    # ---------------------------------------------------------------- reduce over ranks, report
    t = torch.tensor([ms_total, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = float(t[0]), float(t[1])
    global_batch = per_gpu_batch * world
    value = global_batch * args.steps / (ms_total / 1e3)
    e2e_value = global_batch * args.steps / (e2e_ms / 1e3)
    if rank == 0:
        out = {
            "metric": "images/sec (device-timed, max over ranks)",
            "value": round(value, 1),
            "unit": "images/s",
            "n_gpus": world,
            "steps": args.steps,
            "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms_total / args.steps, 4),
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": None,
            "dtype": {"fp32": "fp32 (tf32 tensor-core products, fp32 accumulate)", "bf16": "bf16",
                      "fp32-strict": "fp32 (no tensor cores)"}[args.dtype],
            "data": "synthetic",
            "impl": args.impl,
            "config": {
                "model": {"mnist_mlp": "examples/mnist MLP 784-2500-2000-1500-1000-500-10 tanh"}.get(
                    args.config, args.config),
                "global_batch": global_batch,
                "per_gpu_batch": per_gpu_batch,
                "seq_len": None,
                "parallelism": (f"dp{world} (1 worker group + 1 server shard per GPU, sync push/pull: "
                                f"{exchange or 'local'})") if exchange != "async" else
                               (f"dp{world} asynchronous worker groups, EASGD exchange every "
                                f"{model.updater.sync_frequency} steps with {world} peer-memory server shards in "
                                f"{args.server_groups} group(s), staleness {args.staleness}"),
                "precision": {"fp32": "fp32 activations/gradients/weights, tf32 tensor-core products, fp32 accumulate",
                              "bf16": "bf16 operands, fp32 accumulate, fp32 master weights + optimizer state",
                              "fp32-strict": "fp32 FMA, TF32 off"}[args.dtype],
                "cuda_graph": ex._graph is not None,
                "schedule": "rotated (backward(t) + push(t) || forward(t+1) per graph replay)" if ex.rotate
                            else "forward(t), backward(t) + push(t) per graph replay",
                "l2": ("per-step working set %.0f MB > 126 MB L2 (no flush needed)" % (ws_bytes / 2**20))
                      if flush is None else "L2 flushed (252 MB write) between timed steps, per-step events",
            },
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 1), "unit": "images/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": round(e2e_ms / args.steps, 4),
                    "result_read": "every step's metrics are copied to pinned memory and read by the host one step "
                                   "later (Executor.train_one_batch_async); the last one inside the timed region",
                    "final_loss_sum_window": [round(float(x), 4) for x in last.tolist()] if last is not None else None},
            "gpu_launches": gpu_launches,
        }
        loss_ok = True
        if args.expect_loss_window and last is not None:
            want = [float(v) for v in args.expect_loss_window.split(",")]
            tol = args.loss_tol or (2e-2 if args.dtype == "bf16" else 5e-3)
            got = [float(x) for x in last.tolist()]
            errs = [abs(g - w_) / max(abs(w_), 1e-6) for g, w_ in zip(got, want)]
            loss_ok = errs[0] <= tol      # the loss decides; the accuracy sum of a few steps at chance level is noise
            out["loss_check"] = {"expected": want, "got": [round(g, 4) for g in got], "rel_err": [round(e, 5) for e in errs],
                                 "tol": tol, "ok": loss_ok}
        print(json.dumps(out))
        if not loss_ok:
            sys.stderr.write(f"bench.py: loss window {out['loss_check']['got']} differs from the expected "
                             f"{out['loss_check']['expected']} by more than {out['loss_check']['tol']}\n")
            sys.stdout.flush()
            os._exit(3)
    sys.stdout.flush()
    if world > 1:
        # NCCL work captured in CUDA graphs can make destroy_process_group hang; every rank has passed the final
        # barrier, so leave without the collective teardown
        dist.barrier()
        torch.cuda.synchronize()
        os._exit(0)
    w.pm.finish()
    return 0


if __name__ == "__main__":
    sys.exit(main())
