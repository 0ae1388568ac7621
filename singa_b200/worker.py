"""Worker / Executor / Performance — the training runtime.

Reference: include/worker/worker.h:37-213, src/worker/worker.cc:14-388.  The step structure is the
reference's (optional validation/test -> TrainOneBatch -> metrics -> display), with the per-layer
``wait_update`` / ``update_param`` hooks around forward / backward.  B200 specifics:

* data is staged from pinned host memory by the parser layers while the previous step computes;
* forward + backward + the fused arena update of one step are captured ONCE in a CUDA graph and replayed
  (the MNIST-scale models are launch-latency bound); LR and step live in device memory so the graph follows
  the schedule;
* phase timers use CUDA events on the compute stream, metrics stay on the device until a display step.
"""
from __future__ import annotations

import logging
import os
import threading
import time
from typing import List, Optional

import torch

from .cluster import Cluster
from .device import ctx
from .neuralnet import NeuralNet, registry_param
from .param import Param
from .param_manager import ParamManager
from .proto import schema as S

log = logging.getLogger("singa")


def _sync():
    if ctx().device.type == "cuda":
        torch.cuda.synchronize()


class Performance:
    """Averages every loss layer's metric blob over the display window and prints
    ``Output from <layer> layer 0 : <loss>\\t1 : <precision>`` (src/worker/worker.cc:350-386).  The sums are
    accumulated on the device by the loss kernel itself; the host reads them only in ``to_string``."""

    def __init__(self, net: NeuralNet):
        self.net = net
        self.names = [l.name for l in net.losslayers]
        self.counter = 0
        self.reset()

    def update(self) -> None:
        self.counter += 1

    def reset(self) -> None:
        for l in self.net.losslayers:
            l.metric.zero_()
        self.counter = 0

    def values(self) -> List[List[float]]:
        c = max(self.counter, 1)
        return [[float(v) / c for v in l.metric.detach().cpu().tolist()] for l in self.net.losslayers]

    def to_string(self) -> str:
        out = ""
        for name, m in zip(self.names, self.values()):
            out += f"Output from {name} layer " + "".join(f"{j} : {v:.6f}\t" for j, v in enumerate(m)) + "\n"
        return out


class Executor:
    def __init__(self, local_threadid: int, model, cluster: Optional[Cluster], pm: ParamManager,
                 train_net: NeuralNet, test_net: Optional[NeuralNet] = None,
                 validation_net: Optional[NeuralNet] = None, use_graph: Optional[bool] = None):
        self.local_threadid = local_threadid
        self.model = model
        self.cluster = cluster
        self.pm = pm
        self.train_net, self.test_net, self.validation_net = train_net, test_net, validation_net
        self.step = 0
        self.ticks = 0
        self.t_forward = self.t_backward = self.t_sync_data = self.t_sync_param = 0.0
        self._events: List[tuple] = []
        self._prefetch_thread: Optional[threading.Thread] = None
        self._copy_stream = torch.cuda.Stream() if ctx().device.type == "cuda" else None
        env = os.environ.get("SINGA_B200_CUDA_GRAPH")
        if use_graph is None:
            use_graph = ctx().is_native and (env != "0") and pm.backend_name in ("local", "fused", "nccl", "async")
        if pm.backend_name == "zmq":
            use_graph = False        # host messages inside the step: run eagerly
        # asynchronous groups: a step that exchanges with the servers launches more kernels than a plain one, so one
        # graph is captured per variant (plain / worker exchange / + server-group averaging)
        self._variant_graphs = {}
        self.use_graph = use_graph
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._graph_warmup = 3
        self._prefed_step = -1
        self.progress_counter = 0          # bumped once per training step (watchdog / heartbeat, utils/watchdog.py)
        self._nvtx = os.environ.get("SINGA_B200_NVTX") == "1" and ctx().device.type == "cuda"
        self._pipeline_feed = ctx().device.type == "cuda" and os.environ.get("SINGA_B200_PIPELINE_FEED", "1") != "0" \
            and pm.backend_name in ("local", "fused", "nccl")
        self._prefetch_q = None
        nside = int(os.environ.get("SINGA_B200_SIDE_STREAMS", "2"))
        self._side_streams = [torch.cuda.Stream() for _ in range(nside)] \
            if (ctx().is_native and nside > 0) else None
        self.graph_kernel_count = 0
        # Rotated schedule: one replayed unit = backward(t) + pushes(t) || forward(t+1).  The parameter pushes of step t
        # then overlap the forward pass of step t+1 (each layer's forward waits for ITS weights only) instead of being
        # joined at the end of backward, where nothing can hide the last bucket.  Used when parameters are exchanged
        # between ranks (SINGA_B200_ROTATE=1/0 forces it on/off).
        rot_env = os.environ.get("SINGA_B200_ROTATE", "auto")
        multi = cluster is not None and pm.backend_name in ("fused", "nccl") and getattr(pm.backend, "world", 1) > 1
        self.rotate = ctx().is_native and pm.backend_name in ("local", "fused", "nccl") and \
            (rot_env == "1" or (rot_env == "auto" and multi))
        self._primed_step = -1            # step whose forward pass has already run (rotated schedule)
        self._rot_parity = 0              # which of the parsers' two output buffers holds the current batch
        self._rot_graphs = [None, None]
        if self.rotate and hasattr(pm.backend, "use_rotated_schedule"):
            pm.backend.use_rotated_schedule()
        self._has_bridges = any(l.is_bridgesrclayer or l.is_bridgedstlayer for l in train_net.layers)
        from .parallel.bridge import BridgeTransport, PeerBridgeTransport
        self.bridge = None
        if self._has_bridges:
            import torch.distributed as dist
            one_group = cluster is not None and cluster.nprocs_per_group > 1 and \
                cluster.nworkers == cluster.nprocs_per_group
            if ctx().is_native and dist.is_available() and dist.is_initialized() and one_group:
                # a group spread over several GPU ranks: bridges are peer-memory copies + device flags (capturable)
                base = cluster.global_procsid - cluster.group_procsid
                self.bridge = PeerBridgeTransport(train_net, dist.get_rank(), dist.get_world_size(), ctx().device,
                                                  rank_of=lambda l: base + cluster.group_procsid_of(l.locationid))
            else:
                self.bridge = BridgeTransport(cluster, local_threadid)

    # -- locality ------------------------------------------------------------------------------------
    def _is_local(self, layer) -> bool:
        if self.cluster is None or not self._has_bridges:
            return True
        return self.bridge.is_local(layer)

    # -- schedule predicates (include/worker/worker.h:118-158) -----------------------------------------
    def _every(self, step, freq, after) -> bool:
        return freq > 0 and step >= after and (step - after) % freq == 0

    def display_now(self, step) -> bool:
        return self._every(step, self.model.display_frequency, self.model.display_after_steps)

    def display_debug_info(self, step) -> bool:
        return self.display_now(step) and self.model.debug and self._group0()

    def stop_now(self, step) -> bool:
        return step >= self.model.train_steps

    def _group0(self) -> bool:
        return self.cluster is None or self.cluster.groupid == 0

    def test_now(self, step) -> bool:
        return self._group0() and self.test_net is not None and \
            self._every(step, self.model.test_frequency, self.model.test_after_steps)

    def validate_now(self, step) -> bool:
        return self._group0() and self.validation_net is not None and \
            self._every(step, self.model.validation_frequency, self.model.validation_after_steps)

    def checkpoint_now(self, step) -> bool:
        # with sharded servers every rank owns a piece of the master copy and writes its own shard file
        sharded = getattr(self.pm.backend, "world", 1) > 1 and hasattr(self.pm.backend, "owned_ranges")
        return (self._group0() or sharded) and step > 0 and \
            self._every(step, self.model.checkpoint_frequency, self.model.checkpoint_after_steps)

    # -- data --------------------------------------------------------------------------------------------
    def setup(self) -> None:
        """Stage the first batch (and start prefetching) — Executor::Setup, src/worker/worker.cc:123-156."""
        self._prefetch(self.train_net, True)

    def _prefetch(self, net: NeuralNet, training: bool) -> None:
        for d in net.datalayers:
            if self._is_local(d):
                d.prefetching(training)
        for p in net.parserlayers:
            if self._is_local(p):
                p.prefetching(training, self._copy_stream)

    def _prefetch_loop(self) -> None:
        while True:
            job = self._prefetch_q.get()
            if job is None:
                return
            try:
                self._prefetch(*job)
            finally:
                self._prefetch_done.set()

    def _start_prefetch(self, net, training) -> None:
        """Hand the next batch to the (persistent) prefetch thread — the reference starts a new std::thread per
        batch (src/worker/worker.cc:306-310); a queue avoids the per-step thread start."""
        if not self.model.prefetch:
            return
        if self._prefetch_q is None:
            import queue
            self._prefetch_q = queue.SimpleQueue()
            self._prefetch_done = threading.Event()
            self._prefetch_done.set()
            self._prefetch_thread = threading.Thread(target=self._prefetch_loop, daemon=True)
            self._prefetch_thread.start()
        self._prefetch_done.clear()
        self._prefetch_q.put((net, training))

    def _join_prefetch(self) -> None:
        if self._prefetch_q is not None:
            self._prefetch_done.wait()

    def _feed(self, net: NeuralNet, training: bool) -> None:
        """Make the staged batch current (parse kernels), then start staging the next one."""
        self._join_prefetch()
        if not self.model.prefetch:
            for d in net.datalayers:
                if self._is_local(d):
                    d.compute_feature(training)
        for p in net.parserlayers:
            if self._is_local(p):
                p.compute_feature(training)
        self._start_prefetch(net, training)

    # -- forward / backward (src/worker/worker.cc:240-302) ---------------------------------------------
    def forward(self, net: NeuralNet, step: int, training: bool, skip_inputs: bool = False) -> None:
        for layer in net.layers:
            if not self._is_local(layer):
                continue
            if layer.is_datalayer or layer.is_parserlayer:
                if not skip_inputs and layer.is_parserlayer:
                    layer.compute_feature(training)
                continue
            if layer.is_bridgedstlayer:
                self.bridge.recv_data(layer)
            if training:
                for _p in layer.get_params():
                    self.pm.wait_update(layer, step)
                    break
            if self._nvtx:
                torch.cuda.nvtx.range_push(f"fwd/{layer.name}")
            layer.compute_feature(training)
            if self._nvtx:
                torch.cuda.nvtx.range_pop()
            if layer.is_bridgesrclayer:
                self.bridge.send_data(layer)
        if training and self.display_debug_info(step):
            log.error("\n" + net.debug_info())

    def backward(self, net: NeuralNet, step: int, join: bool = True) -> None:
        """Reverse-topological walk.  On the native path a layer's parameter gradient (wgrad GEMM + bias
        reduction) is independent of its data gradient, so it is issued on a side stream and the data
        gradient — the critical path to the layers below — continues on the main stream; the parameter
        update joins the side streams.  Small GEMMs that cannot fill 148 SMs alone then overlap."""
        if net.param_shares:
            net.zero_shared_grads()
        split = self._side_streams is not None and self.pm.backend_name in ("local", "fused", "nccl")
        main = torch.cuda.current_stream() if split else None
        k = 0
        for layer in reversed(net.layers):
            if not self._is_local(layer) or layer.is_datalayer or layer.is_parserlayer:
                continue
            if layer.is_bridgesrclayer:
                self.bridge.recv_grad(layer)
            has_params = bool(layer.get_params())
            if self._nvtx:
                torch.cuda.nvtx.range_push(f"bwd/{layer.name}")
            if split and has_params and getattr(layer, "split_backward", False):
                side = self._side_streams[k % len(self._side_streams)]
                k += 1
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    layer.compute_param_gradient()
                layer.compute_data_gradient()
                last = layer is self.pm._first_param_layer
                if last and (self.pm.backend_name == "nccl" or
                             (self.pm.backend_name == "local" and not self.pm.per_layer_update)):
                    for s_ in self._side_streams:      # the update / collective on the compute stream needs
                        main.wait_stream(s_)           # every weight gradient issued on the side streams
                self.pm.update_param(layer, step, side)
                if last:
                    for s_ in self._side_streams:
                        main.wait_stream(s_)
            else:
                layer.compute_gradient()
                if has_params:
                    self.pm.update_param(layer, step)
            if self._nvtx:
                torch.cuda.nvtx.range_pop()
            if layer.is_bridgedstlayer:
                self.bridge.send_grad(layer)
        if join:
            self.pm.end_step(step)

    # -- one training step --------------------------------------------------------------------------------
    def train_one_batch(self, step: int, _fetch: bool = False):
        """One training step.  On a GPU the host-side preparation of step+1 (LR/step scalars, parse kernels of the
        already staged batch, start of the next staging copy) is ENQUEUED right after this step's compute has been
        launched, so it overlaps the GPU work; stream order keeps it behind this step's kernels."""
        if self.rotate and not self._has_bridges:
            self._run_rotated(step)
            return self._fetch_metrics_async(step) if _fetch else None
        if self._prefed_step != step:
            self.pm.begin_step(step)
            self._feed(self.train_net, True)
        self._prefed_step = -1
        self._run_compute(step)
        res = self._fetch_metrics_async(step) if _fetch else None
        self._prefeed_next(step)
        return res

    def train_one_batch_async(self, step: int) -> "StepResult":
        """``train_one_batch`` that also starts a device->host copy of the loss layers' metrics (into a ring of
        pinned slots, right behind the step's last kernel) and returns a handle; ``handle.result()`` blocks until
        that copy has landed.  Calling it for step i+1 BEFORE asking for step i's result keeps the GPU busy while the
        host reads: every step's result is still read, one step late (used by bench.py's end-to-end region)."""
        return self.train_one_batch(step, _fetch=True)

    def _fetch_metrics_async(self, step: int) -> "StepResult":
        layers = self.train_net.losslayers
        if not hasattr(self, "_metric_ring"):
            n = sum(l.metric.numel() for l in layers)
            pin = ctx().device.type == "cuda"
            self._metric_ring = [torch.zeros(max(n, 1), dtype=torch.float32, pin_memory=pin) for _ in range(8)]
            self._metric_events = [torch.cuda.Event() if pin else None for _ in range(8)]
            self._metric_slot = 0
        k = self._metric_slot
        self._metric_slot = (k + 1) % len(self._metric_ring)
        buf, off = self._metric_ring[k], 0
        for l in layers:
            m = l.metric.reshape(-1)
            buf[off:off + m.numel()].copy_(m, non_blocking=True)
            off += m.numel()
        ev = self._metric_events[k]
        if ev is not None:
            ev.record()
        return StepResult(step, buf, ev)

    def _prefeed_next(self, step: int) -> None:
        if self._pipeline_feed and not self.stop_now(step + 1) and not self.test_now(step + 1) \
                and not self.validate_now(step + 1):
            self.pm.begin_step(step + 1)
            self._feed(self.train_net, True)
            self._prefed_step = step + 1

    # -- rotated schedule ---------------------------------------------------------------------------------
    def _feed_next(self, net: NeuralNet, training: bool) -> None:
        """Parse the staged batch into the parsers' spare output buffers (the current batch is still needed by the
        backward pass of the running step), then start staging the batch after it."""
        self._join_prefetch()
        if not self.model.prefetch:
            for d in net.datalayers:
                if self._is_local(d):
                    d.compute_feature(training)
        for p in net.parserlayers:
            if self._is_local(p):
                p.compute_feature_next(training)
        self._start_prefetch(net, training)

    def _rotated_body(self, step: int) -> None:
        """backward(step) with the pushes left running, then forward(step + 1) on the batch in the parsers' spare buffers; the
        pushes are joined at the very end."""
        net = self.train_net
        self.pm.pre_forward(step + 1)      # nothing in backward reads what this advances (off the critical path)
        self.backward(net, step, join=False)
        for p in net.parserlayers:
            if self._is_local(p):
                p.flip()                   # batch step+1 (parsed into the other buffer) becomes current: no copy
        self.forward(net, step + 1, True, skip_inputs=True)
        self.pm.end_step(step)

    def _run_rotated(self, step: int, feed: bool = True) -> None:
        net = self.train_net
        if self._primed_step != step:
            # first step (or after a test / resume): run this step's forward pass on its own
            self.pm.begin_step(step)
            self._feed(net, True)
            self.forward(net, step, True, skip_inputs=True)
        if feed:
            self._feed_next(net, True)
        self.pm.begin_step(step + 1, lr_step=step)     # dropout seeds of forward(step+1), LR of the pushes of `step`
        if self.use_graph:
            # two graphs: the parsers' output buffers alternate, so consecutive steps capture different pointers
            par = self._rot_parity
            if self._rot_graphs[par] is None and self.ticks > self._graph_warmup:
                self._graph = None
                self._capture(step, rotated=True)      # runs nothing: the replay below performs this step
                self._rot_graphs[par] = self._graph
                if self._graph is not None:            # capturing walked the body once in Python: undo its flip
                    for p in net.parserlayers:
                        if self._is_local(p):
                            p.flip()
            if self._rot_graphs[par] is not None:
                self._rot_graphs[par].replay()
                for p in net.parserlayers:             # the capture flipped the Python-side pointers once; keep them
                    if self._is_local(p):              # in step with what the replayed graph just did
                        p.flip()
                self._rot_parity ^= 1
                self._primed_step = step + 1
                return
        self._rotated_body(step)
        self._rot_parity ^= 1
        self._primed_step = step + 1

    def _run_compute(self, step: int) -> None:
        """forward + backward (+ per-layer parameter update / push) of one step on the already-parsed batch;
        replayed from a CUDA graph once captured."""
        cuda = ctx().device.type == "cuda"
        if self.use_graph and (not self._has_bridges or getattr(self.bridge, "capturable", False)):
            be = self.pm.backend
            if be is not None and hasattr(be, "graph_variant"):
                v = be.graph_variant(step)
                fresh = False
                if v not in self._variant_graphs and self.ticks > self._graph_warmup:
                    self._graph = None
                    self._capture(step)                      # walks the step once in Python (host counters included)
                    if self._graph is None:
                        return self._run_compute(step)       # capture failed: use_graph is off now
                    self._variant_graphs[v] = self._graph
                    fresh = True
                if v in self._variant_graphs:
                    self._variant_graphs[v].replay()
                    if not fresh:
                        be.after_replay(step)
                    return
            else:
                if self._graph is None and self.ticks > self._graph_warmup:
                    self._capture(step)
                if self._graph is not None:
                    self._graph.replay()
                    return
        if cuda:
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
        t0 = time.monotonic()
        self.forward(self.train_net, step, True, skip_inputs=True)
        if cuda:
            e1.record()
        t1 = time.monotonic()
        self.backward(self.train_net, step)
        if cuda:
            e2.record()
            self._events.append((e0, e1, e2))
            if len(self._events) > 4096:
                self._events = self._events[-1024:]
        else:
            self.t_forward += (t1 - t0) * 1e3
            self.t_backward += (time.monotonic() - t1) * 1e3

    def _capture(self, step: int, rotated: bool = False) -> None:
        from .ops import native
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        n0 = native.launch_count
        try:
            with torch.cuda.graph(g):
                if rotated:
                    self._rotated_body(step)
                else:
                    self.forward(self.train_net, step, True, skip_inputs=True)
                    self.backward(self.train_net, step)
        except Exception as e:  # noqa: BLE001 - e.g. a library op that cannot be captured: run eagerly instead
            log.error("CUDA-graph capture failed (%s); continuing without graphs", str(e).splitlines()[0])
            torch.cuda.synchronize()
            self.use_graph = False
            native.launch_count = n0
            return
        self.graph_kernel_count = native.launch_count - n0
        native.launch_count = n0          # capture records launches, it does not execute them
        self._graph = g
        # the replay in _run_compute performs this step's work

    def timer_info(self) -> str:
        """``Total / forward / backward`` milliseconds per iteration (include/worker/worker.h:91-114)."""
        if self._events:
            _sync()
            for e0, e1, e2 in self._events:
                self.t_forward += e0.elapsed_time(e1)
                self.t_backward += e1.elapsed_time(e2)
            self._events.clear()
        ticks = max(self.ticks, 1)
        tf, tb = self.t_forward / ticks, self.t_backward / ticks
        s = "Total\t%6.2f\tforward\t%6.2f\tbackward\t%6.2f\t" % (tf + tb, tf, tb)
        s += "worker_gen_sync\t%6.2f\tworker_handle_sync\t%6.2f\n" % (Param.worker_gen_sync / ticks,
                                                                      Param.worker_handle_sync / ticks)
        Param.worker_gen_sync = Param.worker_handle_sync = 0.0
        self.t_forward = self.t_backward = self.t_sync_data = self.t_sync_param = 0.0
        self.ticks = 0
        return s

    def run_one_batch(self, step: int, perf: Optional[Performance] = None) -> None:
        self.ticks += 1
        if self.validate_now(step):
            log.error("Validation at step %d", step)
            self.test(self.validation_net, self.model.validation_steps, perf is not None)
        if self.test_now(step):
            log.error("Test at step %d", step)
            self.test(self.test_net, self.model.test_steps, perf is not None)
        self.train_one_batch(step)
        self.progress_counter += 1
        if perf is not None:
            perf.update()
            if self.display_now(step):
                log.error("Training at step %d", step)
                log.error("\t%s", perf.to_string())
                perf.reset()
                log.error("\t%s", self.timer_info())

    def test(self, net: NeuralNet, nsteps: int, disperf: bool = True) -> Optional[List[List[float]]]:
        """Forward-only passes over ``nsteps`` batches with their own staging (worker.cc:318-348)."""
        if net is None or nsteps <= 0:
            return None
        self._join_prefetch()
        perf = Performance(net)
        for b in range(nsteps):
            for d in net.datalayers:
                d.compute_feature(False)
            self.forward(net, b, False)
            perf.update()
        vals = perf.values()
        if disperf:
            log.error("\t%s", perf.to_string())
        # training prefetch may have been consumed by nothing: restage for the train net
        return vals

    def run(self, perf: Optional[Performance], start_step: int = 0, on_step=None) -> int:
        step = start_step
        from .utils import watchdog as _wd
        guards = _wd.start_for(self, float(os.environ.get("SINGA_B200_WATCHDOG", "0") or 0))
        try:
            while not self.stop_now(step):
                self.run_one_batch(step, perf)
                if on_step is not None:
                    on_step(step)
                step += 1
        finally:
            for g in guards:
                g.stop()
        self.step = step
        self._join_prefetch()
        return step


class StepResult:
    """Handle for the metrics of one step (``Executor.train_one_batch_async``).  At most 8 may be outstanding."""

    def __init__(self, step: int, buf: torch.Tensor, event):
        self.step, self._buf, self._event = step, buf, event

    def ready(self) -> bool:
        return self._event is None or self._event.query()

    def result(self) -> torch.Tensor:
        """Host tensor with the loss layers' metrics of this step (loss, accuracy, ...), blocking until it arrived."""
        if self._event is not None:
            self._event.synchronize()
        return self._buf


class HogwildGroup:
    """Several executor threads inside one process training ONE shared set of parameters.  On a GPU every thread
    issues its replica's kernels on its own CUDA stream (hogwild = unordered updates of the shared master weights
    across streams; the bf16 operand mirror is refreshed right after each update).

    Reference: ``Worker::Start`` spawns ``nthreads_per_procs - 1`` extra ``Executor`` threads
    (src/worker/worker.cc:33-39) and ``ParamManager::UpdateParam`` either lets every thread update the shared
    Param without locks (hogwild, src/utils/param_manager.cc:165-167) or aggregates the threads' gradients under
    a mutex and applies one update with ``grad_scale = 1/shares`` (:169-190).  Each thread owns a replica of the
    net (its own activations, gradients and data stream) whose parameters alias the master net's values.  The
    aggregated mode never blocks the host on the device: the mutex fixes the order in which the threads add their
    gradients, a CUDA event per parameter makes each thread's stream wait for the previous add, and the thread that
    arrives last launches the one update.
    """

    def __init__(self, worker: "Worker", nthreads: int, hogwild: bool):
        self.worker, self.nthreads, self.hogwild = worker, nthreads, hogwild
        self.replicas: List[NeuralNet] = [worker.train_net]
        for t in range(1, nthreads):
            net = worker.setup_neuralnet(worker.model, S.kTrain)
            net.share_weights(worker.train_net)
            for d in net.datalayers:                      # every thread reads its own part of the stream
                d.source.skip(t * d.batchsize)
            self.replicas.append(net)
        self._lock = threading.Lock()
        self._agg = {}
        self._agg_evt = {}
        self._arrived = {}
        self.updates = 0
        self._cuda = ctx().device.type == "cuda"
        self._streams = [torch.cuda.Stream() for _ in range(nthreads)] if self._cuda else None
        self.errors: List[BaseException] = []

    def _apply(self, upd, step: int, o, grad_scale: float = 1.0) -> None:
        upd.update(step, o, grad_scale=grad_scale)
        if getattr(o, "data16", None) is not None:        # native path: GEMM operands are read from the bf16 mirror
            o.data16.copy_(o.data)
            if getattr(o, "pack16", None) is not None:
                o.pack16()

    def _thread(self, tid: int, steps: int, start: int) -> None:
        try:
            if self._cuda:
                torch.cuda.set_device(ctx().device)
                with torch.cuda.stream(self._streams[tid]):
                    self._thread_body(tid, steps, start)
                    self._streams[tid].synchronize()
            else:
                self._thread_body(tid, steps, start)
        except BaseException as e:  # noqa: BLE001 - surfaced by run()
            self.errors.append(e)

    def _thread_body(self, tid: int, steps: int, start: int) -> None:
        net = self.replicas[tid]
        upd = self.worker.pm.updater
        owners = [p if p.is_owner else p.owner for p in net.params]
        for step in range(start, start + steps):
            for d in net.datalayers:
                d.compute_feature(True)
                if self.nthreads > 1:
                    d.source.skip((self.nthreads - 1) * d.batchsize)
            for l in net.layers:
                if not l.is_datalayer:
                    l.compute_feature(True)
            for l in reversed(net.layers):
                if l.is_datalayer or l.is_parserlayer:
                    continue
                l.compute_gradient()
                for p, o in zip(l.get_params(), [q if q.is_owner else q.owner for q in l.get_params()]):
                    if self.hogwild:
                        o.grad = p.grad                   # lock-free: this thread's gradient, shared value
                        self._apply(upd, step, o)
                        self.updates += 1
                    else:
                        # aggregation without host synchronisation: the lock orders the threads on the host, and an
                        # event per parameter carries that order to the device — every thread adds its gradient on its
                        # OWN stream behind the previous thread's add (or update), into a buffer kept across steps
                        with self._lock:
                            key = id(o)
                            st = torch.cuda.current_stream() if self._cuda else None
                            if self._cuda and key in self._agg_evt:
                                st.wait_event(self._agg_evt[key])
                            acc = self._agg.get(key)
                            if acc is None:
                                acc = self._agg[key] = torch.empty_like(p.grad)
                            n = self._arrived.get(key, 0)
                            if n == 0:
                                acc.copy_(p.grad)
                            else:
                                acc.add_(p.grad)
                            self._arrived[key] = n + 1
                            if n + 1 == self.nthreads:
                                o.grad = acc
                                self._arrived[key] = 0
                                self._apply(upd, step, o, grad_scale=1.0 / self.nthreads)
                                self.updates += 1
                            if self._cuda:
                                ev = torch.cuda.Event()
                                ev.record(st)
                                self._agg_evt[key] = ev

    def run(self, steps: int, start: int = 0) -> None:
        ths = [threading.Thread(target=self._thread, args=(t, steps, start)) for t in range(self.nthreads)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        if self.errors:
            raise self.errors[0]


class Worker:
    """``Worker(cluster).start(model)`` — build the phase nets, parameters and executors, warm up, exchange
    the initial parameters with the servers and train (src/worker/worker.cc:14-106)."""

    def __init__(self, cluster: Optional[Cluster] = None, backend: Optional[str] = None, seed: int = 0,
                 use_graph: Optional[bool] = None):
        self.cluster = cluster
        self.backend = backend
        self.seed = seed
        self.use_graph = use_graph
        self.train_net = self.test_net = self.validation_net = None
        self.pm: Optional[ParamManager] = None
        self.executor: Optional[Executor] = None
        self.perf: Optional[Performance] = None
        self.model = None
        self.start_step = 0

    # phase filtering by LayerProto.exclude (worker.cc:69-95)
    def setup_neuralnet(self, model, phase: int) -> Optional[NeuralNet]:
        proto = S.NetProto()
        proto.partition_type = model.neuralnet.partition_type
        for lp in model.neuralnet.layer:
            if phase in list(lp.exclude):
                continue
            proto.layer.add().CopyFrom(lp)
        if not any(l.type in ("kShardData", "kLMDBData") for l in proto.layer):
            return None
        gsize = self.cluster.nthreads_per_group if self.cluster is not None else 1
        partitioned = proto.partition_type != S.kNone or any(
            l.HasField("partition_type") and l.partition_type != S.kNone for l in proto.layer) or \
            any(l.locationid != 0 for l in proto.layer)
        return NeuralNet(proto, gsize if partitioned else 1)

    def build(self, model) -> None:
        self.model = model
        if model.alg == S.enum_value("ModelProto.GradCalcAlg", "kContrastiveDivergence"):
            # declared in the reference (src/proto/model.proto:40-44, include/worker/base_layer.h:34,96-108) with no
            # layer or worker code behind it; training such a model with back-propagation would be silently wrong
            raise NotImplementedError(
                "ModelProto.alg = kContrastiveDivergence: neither the reference nor this framework has RBM layers or a "
                "CD training loop; use kBackPropagation")
        registry_param(model.updater.param_type)
        self.train_net = self.setup_neuralnet(model, S.kTrain)
        assert self.train_net is not None, "the model has no training data layer"
        has_test = model.test_steps > 0 and model.test_frequency > 0
        has_val = model.validation_steps > 0 and model.validation_frequency > 0
        if has_test:
            self.test_net = self.setup_neuralnet(model, S.kTest)
            if self.test_net is not None:
                self.test_net.share_weights(self.train_net)
        if has_val:
            self.validation_net = self.setup_neuralnet(model, S.kValidation)
            if self.validation_net is not None:
                self.validation_net.share_weights(self.train_net)
        self.pm = ParamManager(self.train_net, model.updater, self.cluster, self.backend, self.seed)
        for net in (self.test_net, self.validation_net):
            if net is not None:
                net.share_weights(self.train_net)   # re-alias after the arena re-pointed the blobs
        self.pm.init_params()
        self.executor = Executor(0, model, self.cluster, self.pm, self.train_net, self.test_net,
                                 self.validation_net, self.use_graph)
        self.perf = Performance(self.train_net)
        self._dump_graphs()

    def _dump_graphs(self) -> None:
        """Node-link JSON of every phase net into ``<workspace>/vis`` (the reference creates that folder,
        src/utils/cluster.cc:29-32, and logs ``NeuralNet::ToString()`` at INFO, src/worker/neuralnet.cc:109,195);
        ``python script/graph.py <workspace>/vis/train.json`` turns it into Graphviz DOT."""
        if self.cluster is None or self.cluster.groupid != 0 or self.cluster.group_procsid != 0:
            return
        try:
            folder = self.cluster.visualization_folder
            folder = folder() if callable(folder) else folder
            os.makedirs(folder, exist_ok=True)
            for name, net in (("train", self.train_net), ("test", self.test_net), ("validation", self.validation_net)):
                if net is not None:
                    with open(os.path.join(folder, f"{name}.json"), "w") as fh:
                        fh.write(net.to_string())
        except Exception as e:  # noqa: BLE001 - visualisation must never stop a job
            log.error("could not write the net graphs: %s", e)

    def resume(self) -> int:
        """Load the latest Param-proto checkpoint (if the model names one or the workspace holds one) and
        return the step to continue from — ``Worker::Resume`` is an empty TODO in the reference
        (src/worker/worker.cc:65-67)."""
        from . import checkpoint
        path = self.model.checkpoint if self.model.HasField("checkpoint") else None
        if path is None and self.cluster is not None:
            path = checkpoint.latest(self.cluster.checkpoint_folder)
        if path and os.path.exists(path):
            step = checkpoint.load(path, self.train_net, self.pm)
            log.error("Resumed from %s at step %d", path, step)
            return step
        return int(self.model.step)

    def start(self, model) -> int:
        log.error("Worker on %s is starting...", self.cluster.hostname if self.cluster else "localhost")
        self.build(model)
        ex, pm = self.executor, self.pm
        self.start_step = self.resume()
        ex.setup()
        # warm-up: time the compute to size RandomSync messages (worker.cc:41-48)
        warm = 0
        if pm.nservers > 0 and self.start_step == 0:
            warm = min(model.updater.warmup_steps, model.train_steps)
            _sync()
            t0 = time.monotonic()
            for step in range(warm):
                ex.run_one_batch(step, self.perf)
            _sync()
            if warm:
                pm.sync_config((time.monotonic() - t0) / warm)
            if self.cluster.groupid == 0:
                pm.send_params_to_servers()
            else:
                pm.get_params_from_servers()
        start = max(self.start_step, warm)
        nthreads = self.cluster.nthreads_per_procs if self.cluster is not None else 1
        if nthreads > 1 and not self.train_net.param_shares and pm.backend_name == "local":
            group = HogwildGroup(self, nthreads, pm.hogwild and model.updater.hogwild)
            group.run(max(0, model.train_steps - start), start)
            log.error("\t%s", Performance(self.train_net).to_string())
            pm.finish()
            return model.train_steps
        step = ex.run(self.perf, start, on_step=self._maybe_checkpoint)
        _sync()
        pm.finish()
        return step

    def _maybe_checkpoint(self, step: int) -> None:
        if self.executor.checkpoint_now(step + 1) and self.cluster is not None:
            from . import checkpoint
            if torch.cuda.is_available():
                torch.cuda.synchronize()      # the step's pushes have published every owner's master copy
            path = checkpoint.save(self.cluster.checkpoint_folder, self.train_net, self.pm, step + 1,
                                   self.model.name)
            log.error("Checkpoint written to %s", path)
