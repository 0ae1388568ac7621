"""Router: one ZeroMQ ROUTER socket per process — the control-plane endpoint.

Reference: include/utils/router.h:16-57, src/utils/router.cc:8-123 (CZMQ); here pyzmq, same frames.
``connect`` pings ``[endpoint]["PING"]`` once a second (up to ``retries`` times) until the server answers
``["PONG"]``; ``bind`` takes the identity ``tcp://addr:port``, waits for ``n`` distinct peers to ping and then
pongs them all — a start-up barrier (message M1 of the survey).  ``send`` prefixes the destination endpoint as
the routing frame, ``recv`` pops the sender identity into ``last_recv_node``.
"""
from __future__ import annotations

import logging
import time
from typing import List, Optional, Sequence

import zmq

log = logging.getLogger("singa.router")


class Router:
    def __init__(self, port: int, context: Optional[zmq.Context] = None, identity: Optional[bytes] = None):
        self.port = port
        self.ctx = context or zmq.Context.instance()
        self.sock = self.ctx.socket(zmq.ROUTER)
        self.sock.setsockopt(zmq.LINGER, 0)
        if identity:
            self.sock.setsockopt(zmq.IDENTITY, identity)
        self.nodes: List[bytes] = []            # endpoints of connected servers, index = server id
        self.last_recv_node: Optional[bytes] = None

    # -- worker side -----------------------------------------------------------------------------------
    def connect(self, addr: str, port: Optional[int] = None, retries: int = 10, interval_ms: int = 1000) -> bool:
        endpoint = f"tcp://{addr}:{port or self.port}".encode()
        self.sock.connect(endpoint.decode())
        poller = zmq.Poller()
        poller.register(self.sock, zmq.POLLIN)
        for _ in range(retries):
            log.debug("Ping %s", endpoint)
            self.sock.send_multipart([endpoint, b"PING"])
            if dict(poller.poll(interval_ms)).get(self.sock):
                frames = self.sock.recv_multipart()
                if len(frames) >= 2 and frames[0] == endpoint and frames[1] == b"PONG":
                    self.nodes.append(endpoint)
                    return True
                log.error("unexpected handshake reply %r", frames[:2])
        return False

    # -- server side -----------------------------------------------------------------------------------
    def bind(self, addr: str, expected_connections: int, timeout_s: float = 120.0) -> bool:
        endpoint = f"tcp://{addr}:{self.port}".encode()
        self.sock.setsockopt(zmq.IDENTITY, endpoint)
        self.sock.bind(f"tcp://*:{self.port}")
        identities = {}
        deadline = time.monotonic() + timeout_s
        poller = zmq.Poller()
        poller.register(self.sock, zmq.POLLIN)
        while len(identities) < expected_connections:
            left = deadline - time.monotonic()
            if left <= 0:
                return False
            if not dict(poller.poll(min(1000, left * 1000))).get(self.sock):
                continue
            frames = self.sock.recv_multipart()
            ident, control = frames[0], frames[1] if len(frames) > 1 else b""
            if control == b"PING":
                identities[ident] = True
            else:
                log.error("server received unexpected message %r during handshake", control)
                return False
        for ident in identities:
            self.sock.send_multipart([ident, b"PONG"])
        return True

    # -- traffic ------------------------------------------------------------------------------------------
    def send(self, frames: Sequence[bytes], serverid: int) -> None:
        self.sock.send_multipart([self.nodes[serverid], *frames])

    def reply(self, frames: Sequence[bytes], identity: Optional[bytes] = None) -> None:
        ident = identity if identity is not None else self.last_recv_node
        self.sock.send_multipart([ident, *frames])

    def recv(self, timeout_ms: Optional[int] = None) -> Optional[List[bytes]]:
        # Handshake leftovers are not traffic: a worker re-sends PING every retry interval until the barrier releases,
        # so duplicates can still be queued behind it (answered here: a late pinger may really be waiting), and the
        # answers to those duplicates reach the worker as extra PONGs (dropped).
        deadline = None if timeout_ms is None else time.monotonic() + timeout_ms / 1000.0
        while True:
            if deadline is not None:
                left = max(0.0, deadline - time.monotonic())
                if not self.sock.poll(int(left * 1000)):
                    return None
            frames = self.sock.recv_multipart()
            if len(frames) == 2 and frames[1] == b"PING":
                self.sock.send_multipart([frames[0], b"PONG"])
                continue
            if len(frames) == 2 and frames[1] == b"PONG":
                continue
            self.last_recv_node = frames[0]
            return frames[1:]

    def close(self, linger_ms: int = 0) -> None:
        """``linger_ms`` > 0 lets queued messages (e.g. the final kStop behind 48 MB of kPut payloads) reach the peer
        before the socket goes away; 0 drops whatever is still queued (the reference sets zsocket linger 0 to avoid
        hangs on exit, src/worker/pm_client.cc:64)."""
        self.sock.close(linger_ms)
