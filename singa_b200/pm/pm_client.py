"""Stack B worker side (reference: include/worker/pm_client.h:31-61, src/worker/pm_client.cc:20-291).

``SingaClient``: the worker group ``g = (id - nservers) / group_size`` talks to server group ``g``; one DEALER per
server of that group, an inproc ROUTER towards the client threads, requests routed by ``paramId % #servers``.
``PMClient``: ``put / get / update / collect`` (non-blocking send, polling collect) with a local-shard fast path.
``client_thread`` is the benchmark body that produced the reference's published numbers (benchmarks/*): thread 0 of
worker 0 PUTs the 12 MNIST-MLP tensors (11,972,510 floats), everyone GETs them, then ``rounds`` x (Update x12 +
Collect x12).
"""
from __future__ import annotations

import logging
import threading
import time
from typing import Dict, List, Optional

import numpy as np
import zmq

from ..param import kData, kGet, kPut, kStop, kUpdate
from .param_shard import ParamShard

log = logging.getLogger("singa.pm_client")

MNIST_MLP_SIZES = [784 * 2500, 2500, 2500 * 2000, 2000, 2000 * 1500, 1500, 1500 * 1000, 1000, 1000 * 500, 500,
                   500 * 10, 10]
assert sum(MNIST_MLP_SIZES) == 11972510


class PMClient:
    def __init__(self, thread_id: int, socket: zmq.Socket, shard: Optional[ParamShard] = None):
        self.tid = thread_id
        self.sock = socket
        self.shard = shard
        self.results: Dict[int, bytes] = {}

    def _send(self, mtype: int, pid: int, content=b"") -> None:
        # zero-copy for tensor payloads: ``content`` may be bytes, a memoryview or a contiguous numpy array (the
        # caller must not modify it until the matching Collect returned, as with the reference's zframe ownership)
        self.sock.send_multipart([str(mtype).encode(), str(pid).encode(), content], copy=False)

    def put(self, pid: int, payload: bytes) -> None:
        if self.shard is not None and self.shard.is_local(pid):
            self.shard.put(pid, payload)
        else:
            self._send(kPut, pid, payload)

    def get(self, pid: int) -> None:
        if self.shard is not None and self.shard.is_local(pid):
            self.results[pid] = self.shard.get(pid)
        else:
            self._send(kGet, pid)

    def update(self, pid: int, payload: bytes) -> None:
        if self.shard is not None and self.shard.is_local(pid):
            self.results[pid] = self.shard.update(pid, payload)
        else:
            self._send(kUpdate, pid, payload)

    def collect(self, pid: int, timeout_ms: int = 0) -> Optional[bytes]:
        """Return the reply for ``pid`` if it has arrived (poll 0 ms like the reference), else None."""
        if pid in self.results:
            return self.results.pop(pid)
        while self.sock.poll(timeout_ms):
            frames = self.sock.recv_multipart(copy=False)
            mtype, rid, content = int(frames[0].bytes), int(frames[1].bytes), frames[2].buffer
            if mtype == kData:
                self.results[rid] = content
            if rid == pid:
                return self.results.pop(pid)
            timeout_ms = 0
        return None

    def collect_blocking(self, pid: int, timeout_s: float = 60.0) -> bytes:
        t0 = time.monotonic()
        while True:
            r = self.collect(pid, 5)
            if r is not None:
                return r
            if time.monotonic() - t0 > timeout_s:
                raise TimeoutError(f"collect({pid}) timed out")


class SingaClient:
    def __init__(self, client_id: int, topology, hosts: List[str], context: Optional[zmq.Context] = None,
                 port_stride: int = 1):
        self.id = client_id                    # global id: servers first, then workers (test_pm.cc:82-88)
        self.ctx = context or zmq.Context.instance()
        n_servers = topology.nservers
        n_groups = max(1, len(topology.server_group))
        group_size = max(1, n_servers // n_groups)
        self.worker_threads = max(1, topology.worker_threads)
        self.local_id = client_id - n_servers
        n_worker_groups = max(1, topology.nworker_groups)
        # worker group -> server group (round-robin when there are fewer server groups than worker groups)
        self.group_id = (self.local_id * n_groups // n_worker_groups) % n_groups if n_worker_groups >= n_groups \
            else self.local_id % n_groups
        def group_of(lid):
            return (lid * n_groups // n_worker_groups) % n_groups if n_worker_groups >= n_groups else lid % n_groups
        # the first client of every server group seeds that group's shards (the reference seeds only from worker 0,
        # which leaves the other server groups empty until the first neighbour sync)
        self.is_group_seeder = self.local_id == 0 or group_of(self.local_id - 1) != self.group_id
        same_host = len(set(hosts[:n_servers])) == 1
        self.server_endpoints = []
        for k in range(group_size):
            sid = self.group_id * group_size + k
            port = topology.port + (sid * port_stride if same_host else 0)
            self.server_endpoints.append(f"tcp://{hosts[sid]}:{port}")
        self.backend_endpoint = f"inproc://singaclient{client_id}"
        self._stop = threading.Event()
        self._ready = threading.Event()

    def proxy(self) -> None:
        backend = self.ctx.socket(zmq.ROUTER)
        backend.setsockopt(zmq.LINGER, 0)
        backend.bind(self.backend_endpoint)
        servers = []
        for ep in self.server_endpoints:
            s = self.ctx.socket(zmq.DEALER)
            s.setsockopt(zmq.LINGER, 0)
            s.setsockopt(zmq.IDENTITY, f"client{self.id}".encode())
            s.connect(ep)
            servers.append(s)
        poller = zmq.Poller()
        for s in (backend, *servers):
            poller.register(s, zmq.POLLIN)
        self._ready.set()
        while not self._stop.is_set():
            events = dict(poller.poll(20))
            if events.get(backend):
                frames = backend.recv_multipart(copy=False)          # [tid][type][paramId][content]
                pid = int(frames[2].bytes)
                if int(frames[1].bytes) == kStop:
                    for s in servers:
                        s.send_multipart([frames[0], frames[1], frames[2]], copy=False)
                else:
                    servers[pid % len(servers)].send_multipart(frames, copy=False)   # frames forwarded, not copied
            for s in servers:
                if events.get(s):
                    # [tid][kData][paramId][content] -> thread tid
                    backend.send_multipart(s.recv_multipart(copy=False), copy=False)
        for s in (backend, *servers):
            s.close(0)

    def thread_socket(self, tid: int) -> zmq.Socket:
        s = self.ctx.socket(zmq.DEALER)
        s.setsockopt(zmq.LINGER, 0)
        s.setsockopt(zmq.IDENTITY, f"t{tid}".encode())
        s.connect(self.backend_endpoint)
        return s

    def stop(self, stop_servers: bool = False) -> None:
        if stop_servers:
            s = self.thread_socket(9999)
            s.send_multipart([str(kStop).encode(), b"0", b""])
            time.sleep(0.2)
            s.close(0)
        self._stop.set()


def client_thread(client: SingaClient, tid: int, sizes: List[int], rounds: int, is_putter: bool,
                  out: dict) -> None:
    sock = client.thread_socket(tid)
    pm = PMClient(tid, sock)
    rng = np.random.default_rng(tid)
    params = [rng.standard_normal(n, dtype=np.float32) * 0.01 for n in sizes]
    if is_putter:
        for i, p in enumerate(params):
            pm.put(i, p)
    for i in range(len(params)):
        pm.get(i)
    for i in range(len(params)):
        pm.collect_blocking(i)
    times = []
    for _ in range(rounds):
        t0 = time.monotonic()
        for i, p in enumerate(params):
            pm.update(i, p)
        for i in range(len(params)):
            pm.collect_blocking(i)
        times.append((time.monotonic() - t0) * 1e3)
        log.debug("Collected: %.3f", times[-1])
    out[tid] = times
    sock.close(0)
