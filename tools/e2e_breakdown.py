"""Host-side breakdown of the end-to-end training step (where the gap between device-timed and e2e comes from)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from singa_b200 import device  # noqa: E402
from singa_b200.models import configs  # noqa: E402
from singa_b200.worker import Worker  # noqa: E402

device.set_context("cuda:0", "native")
m = configs.mnist_mlp()
m.display_frequency = 0
w = Worker(use_graph=True)
w.build(m)
ex, net = w.executor, w.train_net
loss = net.losslayers[0]
ex.setup()
for s in range(10):
    ex.ticks += 1
    ex.train_one_batch(s)
    loss.metric.cpu()
T = {"begin_step": 0.0, "join_prefetch": 0.0, "parse": 0.0, "start_prefetch": 0.0, "compute": 0.0, "metric": 0.0}
N = 300
t_all = time.perf_counter()
for s in range(10, 10 + N):
    ex.ticks += 1
    t0 = time.perf_counter(); w.pm.begin_step(s)
    t1 = time.perf_counter(); ex._join_prefetch()
    t2 = time.perf_counter()
    for p in net.parserlayers:
        p.compute_feature(True)
    t3 = time.perf_counter(); ex._start_prefetch(net, True)
    t4 = time.perf_counter(); ex._run_compute(s)
    t5 = time.perf_counter(); loss.metric.cpu()
    t6 = time.perf_counter()
    for k, v in zip(T, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t6 - t5)):
        T[k] += v
total = time.perf_counter() - t_all
ex._join_prefetch()
print(f"e2e {total / N * 1e3:.3f} ms/step")
for k, v in T.items():
    print(f"  {k:15s} {v / N * 1e6:8.1f} us")
