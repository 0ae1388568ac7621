"""Stack B server (reference: include/server/pm_server.h:20-68, src/server/pm_server.cc:24-354).

``SingaServer``: frontend ROUTER (tcp) <-> backend DEALER (inproc) proxy with ``server_threads`` worker threads and
one DEALER per neighbour server group.  A worker thread decodes
``[identity][thread_id]([REQUEUE]...)[type][paramId][content]``, dispatches to ``PMServer.handle_*``, and re-queues
(through its own DEALER to the frontend) any request whose parameter is not there yet.  Replies are tagged ``REP``
or — when ``ParamShard.sync_now`` fires — ``SYNC``; the proxy duplicates SYNC replies to every neighbour as
``kSyncRequest``.  A ``kStop`` message (absent from the reference, which runs until ^C) ends the server.
"""
from __future__ import annotations

import logging
import threading
from typing import List, Optional

import zmq

from ..param import kData, kGet, kPut, kStop, kSyncRequest, kSyncResponse, kUpdate
from .param_shard import ParamShard

log = logging.getLogger("singa.pm_server")

RESPONSE_HEADER, SYNC_MSG, REQUEUE_ID = b"REP", b"SYNC", b"REQUEUE"


class PMServer:
    """Handlers; each returns ``(to_backend, to_requeue)`` frame lists (either may be None)."""

    def __init__(self, server_id: int, shard: ParamShard):
        self.id = server_id
        self.shard = shard

    @staticmethod
    def _hdr(t: int, pid: int) -> List[bytes]:
        return [str(t).encode(), str(pid).encode()]

    def handle_put(self, pid, ident, tid, content):
        self.shard.put(pid, content)
        return None, None

    def handle_get(self, pid, ident, tid, content):
        val = self.shard.get(pid)
        if val is not None:
            return [RESPONSE_HEADER, ident, tid, *self._hdr(kData, pid), val], None
        return None, [REQUEUE_ID, ident, tid, *self._hdr(kGet, pid)]

    def handle_update(self, pid, ident, tid, content):
        val = self.shard.update(pid, content)
        if val is not None:
            tag = SYNC_MSG if self.shard.sync_now(pid) else RESPONSE_HEADER
            return [tag, ident, tid, *self._hdr(kData, pid), val], None
        return None, [REQUEUE_ID, ident, tid, *self._hdr(kUpdate, pid), content]

    def handle_sync_request(self, pid, ident, tid, content):
        val = self.shard.sync_update(pid, content)
        if val is not None:
            # identity twice: the frontend ROUTER consumes one for routing, the requester's thread parser the other
            return [RESPONSE_HEADER, ident, ident, tid, *self._hdr(kSyncResponse, pid), val], None
        return None, [REQUEUE_ID, ident, tid, *self._hdr(kSyncRequest, pid), content]

    def handle_sync_response(self, pid, ident, tid, content):
        if self.shard.sync_update(pid, content) is not None:
            return None, None
        return None, [REQUEUE_ID, ident, tid, *self._hdr(kSyncResponse, pid), content]


class SingaServer:
    def __init__(self, server_id: int, topology, hosts: List[str], context: Optional[zmq.Context] = None,
                 sync_rule: str = "replace", port_stride: int = 1):
        self.id = server_id
        self.ctx = context or zmq.Context.instance()
        n_servers = topology.nservers
        n_groups = max(1, len(topology.server_group))
        self.group_size = max(1, n_servers // n_groups)
        self.server_threads = max(1, topology.server_threads)
        same_host = len(set(hosts[:n_servers])) == 1
        self._port = lambda sid: topology.port + (sid * port_stride if same_host else 0)
        self.frontend_endpoint = f"tcp://{hosts[server_id]}:{self._port(server_id)}"
        self.backend_endpoint = f"inproc://singanus{server_id}"
        local_id = server_id % self.group_size
        sync_interval = 1
        self.neighbors: List[str] = []
        for g in topology.server_group:
            if g.id == server_id // self.group_size:
                sync_interval = g.sync_interval
                for nb in g.neighbor:
                    sid = nb * self.group_size + local_id
                    self.neighbors.append(f"tcp://{hosts[sid]}:{self._port(sid)}")
                break
        self.shard = ParamShard(server_id, sync_interval, sync_rule)
        self._stop = threading.Event()
        self.nrequests = 0

    def _thread(self):
        backend = self.ctx.socket(zmq.DEALER)
        frontend = self.ctx.socket(zmq.DEALER)
        for s in (backend, frontend):
            s.setsockopt(zmq.LINGER, 0)
        backend.connect(self.backend_endpoint)
        frontend.connect(self.frontend_endpoint)
        pm = PMServer(self.id, self.shard)
        dispatch = {kPut: pm.handle_put, kGet: pm.handle_get, kUpdate: pm.handle_update,
                    kSyncRequest: pm.handle_sync_request, kSyncResponse: pm.handle_sync_response}
        poller = zmq.Poller()
        poller.register(backend, zmq.POLLIN)
        while not self._stop.is_set():
            if not dict(poller.poll(50)).get(backend):
                continue
            # zero-copy: tensor payloads stay in the zmq frames they arrived in (a Put / Update copies them once,
            # into the parameter blob; the echo of an Update re-sends the received frame)
            frames = backend.recv_multipart(copy=False)
            ident, tid = frames[0].bytes, frames[1].bytes
            rest = frames[2:]
            if tid == REQUEUE_ID:
                ident, tid, rest = rest[0].bytes, rest[1].bytes, rest[2:]
            mtype, pid = int(rest[0].bytes), int(rest[1].bytes)
            content = rest[2] if len(rest) > 2 else b""
            if mtype == kStop:
                self._stop.set()
                break
            self.nrequests += 1
            to_backend, to_requeue = dispatch[mtype](pid, ident, tid, content)
            if to_backend is not None:
                backend.send_multipart(to_backend, copy=False)
            if to_requeue is not None:
                frontend.send_multipart(to_requeue, copy=False)
        backend.close(0)
        frontend.close(0)

    def start(self) -> None:
        """Blocking proxy loop (``SingaServer::StartServer``)."""
        frontend = self.ctx.socket(zmq.ROUTER)
        backend = self.ctx.socket(zmq.DEALER)
        for s in (frontend, backend):
            s.setsockopt(zmq.LINGER, 0)
        frontend.bind(self.frontend_endpoint.replace(self.frontend_endpoint.split("//")[1].split(":")[0], "*"))
        backend.bind(self.backend_endpoint)
        nbs = []
        for ep in self.neighbors:
            s = self.ctx.socket(zmq.DEALER)
            s.setsockopt(zmq.LINGER, 0)
            s.setsockopt(zmq.IDENTITY, f"server{self.id}".encode())
            s.connect(ep)
            nbs.append(s)
        threads = [threading.Thread(target=self._thread, daemon=True) for _ in range(self.server_threads)]
        for t in threads:
            t.start()
        poller = zmq.Poller()
        for s in (frontend, backend, *nbs):
            poller.register(s, zmq.POLLIN)
        while not self._stop.is_set():
            events = dict(poller.poll(50))
            if events.get(frontend):
                backend.send_multipart(frontend.recv_multipart(copy=False), copy=False)
            if events.get(backend):
                frames = backend.recv_multipart(copy=False)
                first, msg = frames[0].bytes, frames[1:]
                if first == SYNC_MSG:
                    # [wid][tid][kData][pid][content] -> neighbours get [tid][kSyncRequest][pid][content]
                    for s in nbs:
                        s.send_multipart([msg[1], str(kSyncRequest).encode(), *msg[3:]], copy=False)
                frontend.send_multipart(msg, copy=False)
            for s in nbs:
                if events.get(s):
                    backend.send_multipart(s.recv_multipart(copy=False), copy=False)
        for t in threads:
            t.join(2)
        for s in (frontend, backend, *nbs):
            s.close(0)
