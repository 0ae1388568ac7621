"""Worker side of Stack A over ZeroMQ (reference: src/utils/param_manager.cc:71-234).

Local updates every step; every ``sync_frequency`` steps each parameter is synchronised with the server that owns
it (``id % nservers``) by the Elastic or RandomSync rule.  ``sync_all`` sends without blocking; ``wait_layer``
(the ``WaitUpdate`` call site) collects the replies a layer needs right before its next forward.
"""
from __future__ import annotations

import logging
from typing import Dict, Set

from ..param import kGet, kPut, kStop, kSync
from ..router import Router

log = logging.getLogger("singa.pm")


class ZmqExchange:
    handles_update = False

    def __init__(self, pm):
        self.pm = pm
        c = pm.cluster
        self.nservers = c.nservers
        self.router = Router(c.router_port)
        for s in range(self.nservers):
            ok = self.router.connect(c.server_addr(s), c.server_router_port(s))
            if not ok:
                raise RuntimeError(f"worker: server {s} at {c.server_addr(s)} is unreachable")
        self.id2param = {p.id: p for p in pm.arena.params}
        self.pending: Set[int] = set()          # param ids with a kSync in flight
        self.layer_params: Dict[str, list] = {l.name: [p.id for p in l.get_params()] for l in pm._param_layers}

    def _server_of(self, pid: int) -> int:
        return pid % self.nservers

    def put_all(self) -> None:
        """kPut every parameter to its server shard (messages M2)."""
        for pid, p in self.id2param.items():
            self.router.send([str(kPut).encode(), str(pid).encode(), p.name.encode(),
                              ",".join(map(str, p.shape)).encode(), p.to_bytes()], self._server_of(pid))

    def get_all(self) -> None:
        """kGet every parameter and block until all replies arrived (messages M3)."""
        want = set(self.id2param)
        for pid in want:
            self.router.send([str(kGet).encode(), str(pid).encode(), self.id2param[pid].name.encode()],
                             self._server_of(pid))
        while want:
            frames = self.router.recv(timeout_ms=60000)
            if frames is None:
                raise TimeoutError(f"worker: no kGet reply for params {sorted(want)}")
            if int(frames[0]) != kGet:
                self._handle(frames)
                continue
            pid = int(frames[1])
            self.id2param[pid].from_bytes(frames[-1])
            want.discard(pid)
        self.pm.arena.sync_mirror()

    def sync_all(self, step: int) -> None:
        arg = self.pm.moving_rate if self.pm.param_type == "Elastic" else self.pm.sample_ratio
        for pid, p in self.id2param.items():
            if pid in self.pending:
                continue
            control, payload = p.gen_sync_msg_from_worker(arg)
            self.router.send([str(kSync).encode(), str(pid).encode(), control.encode(), payload],
                             self._server_of(pid))
            self.pending.add(pid)

    def _handle(self, frames) -> None:
        mtype = int(frames[0])
        if mtype == kSync:
            pid = int(frames[1])
            self.id2param[pid].parse_sync_msg_from_ps(frames[2].decode(), frames[3])
            self.pending.discard(pid)
            self.pm.paramid2version[pid] = self.pm.paramid2version.get(pid, 0) + 1
        else:
            log.error("worker: unexpected message type %d", mtype)

    def wait_layer(self, layer, step: int) -> None:
        ids = [pid for pid in self.layer_params.get(layer.name, []) if pid in self.pending]
        if not ids:
            return
        dirty = False
        while any(pid in self.pending for pid in ids):
            frames = self.router.recv(timeout_ms=60000)
            if frames is None:
                raise TimeoutError(f"worker: no kSync reply for params {ids}")
            self._handle(frames)
            dirty = True
        if dirty:
            self.pm.arena.sync_mirror()

    def close(self) -> None:
        # collect outstanding replies so the servers' sockets drain, then tell every server to stop (M6)
        while self.pending:
            frames = self.router.recv(timeout_ms=5000)
            if frames is None:
                break
            self._handle(frames)
        for s in range(self.nservers):
            self.router.send([str(kStop).encode()], s)
        # bounded linger: a job shorter than the warm-up ends right after put_all, with the stop message queued behind
        # the parameter payloads — closing with linger 0 dropped it and the servers never exited
        self.router.close(linger_ms=10000)
