"""Parameter server, Stack A (reference: include/server/server.h, src/server/server.cc:18-214).

One poll loop on the ROUTER plus a pool of ``nthreads_per_server`` sync workers (the reference's zactors).
Messages (intended framing of the survey's §2.8 item 1 — ``[type][id][...]`` both ways):

  kPut  [1][id][name][shape csv][float32 payload]      create / overwrite the server copy
  kGet  [0][id][name]                                  reply [0][id][name][payload]; parked until the param exists
  kSync [2][id][control][payload]                      dispatched to a free sync worker -> Param.handle_sync_msg;
                                                       one param is handled by one worker at a time (lock map);
                                                       with no free worker the router is not polled (back-pressure)
  kStop [6]                                            the server exits after ``nworkers`` stops with all workers idle
"""
from __future__ import annotations

import logging
import time
from concurrent.futures import Future, ThreadPoolExecutor
from typing import Dict, List, Tuple

from .cluster import Cluster
from .param import Param, kGet, kPut, kStop, kSync, param_factory
from .router import Router

log = logging.getLogger("singa.server")


def handle_sync_from_worker(param: Param, control: str, payload: bytes):
    """Body of one sync worker (server.cc:18-43)."""
    return param.handle_sync_msg(control, payload)


class Server:
    def __init__(self, cluster: Cluster, server_id: int | None = None):
        self.cluster = cluster
        self.server_id = server_id if server_id is not None else cluster.global_procsid - cluster.nworkers
        self.params: Dict[int, Param] = {}
        self.bytes_received = 0
        self.nsync = 0

    def run(self, bind_timeout_s: float = 120.0) -> None:
        c = self.cluster
        binder = Router(c.server_router_port(self.server_id))
        ok = binder.bind(c.server_addr(self.server_id), c.nworkers, bind_timeout_s)
        if not ok:
            binder.close()
            raise RuntimeError("server: workers did not connect in time")
        nactors = c.nthreads_per_server
        pool = ThreadPoolExecutor(max_workers=nactors)
        locks: Dict[int, bool] = {}
        get_requests: List[Tuple[int, bytes, bytes]] = []
        sync_requests: List[Tuple[int, bytes, str, bytes]] = []
        running: Dict[Future, Tuple[int, bytes]] = {}
        nstop = 0
        t_start = time.monotonic()
        try:
            while True:
                # completed sync workers: forward their reply, unlock the param
                for fut in [f for f in running if f.done()]:
                    pid, ident = running.pop(fut)
                    control, payload = fut.result()
                    binder.reply([str(kSync).encode(), str(pid).encode(), control.encode(), payload], ident)
                    locks[pid] = False
                # drain parked gets whose param has arrived
                for req in list(get_requests):
                    pid, ident, name = req
                    if pid in self.params:
                        binder.reply([str(kGet).encode(), str(pid).encode(), name, self.params[pid].handle_get_msg()],
                                     ident)
                        get_requests.remove(req)
                # drain parked syncs whose param is free and a worker is available
                for req in list(sync_requests):
                    if len(running) >= nactors:
                        break
                    pid, ident, control, payload = req
                    if not locks.get(pid, False) and pid in self.params:
                        locks[pid] = True
                        running[pool.submit(handle_sync_from_worker, self.params[pid], control, payload)] = (pid, ident)
                        sync_requests.remove(req)
                if nstop >= c.nworkers and not running and not sync_requests:
                    break
                if len(running) >= nactors:
                    time.sleep(0.0002)      # back-pressure: no free worker -> do not read the router
                    continue
                frames = binder.recv(timeout_ms=1 if (running or sync_requests or get_requests) else 50)
                if frames is None:
                    continue
                ident = binder.last_recv_node
                mtype = int(frames[0])
                if mtype == kGet:
                    pid, name = int(frames[1]), frames[2] if len(frames) > 2 else b""
                    if pid in self.params:
                        binder.reply([frames[0], frames[1], name, self.params[pid].handle_get_msg()], ident)
                    else:
                        get_requests.append((pid, ident, name))
                elif mtype == kPut:
                    pid = int(frames[1])
                    shape = [int(x) for x in frames[3].decode().split(",") if x] if len(frames) > 4 else None
                    payload = frames[-1]
                    if pid not in self.params:
                        p = param_factory.create("Param")
                        p.id = pid
                        p.name = frames[2].decode() if len(frames) > 2 else ""
                        self.params[pid] = p
                        locks[pid] = False
                    self.params[pid].handle_put_msg(payload, shape)
                    self.bytes_received += len(payload)
                elif mtype == kSync:
                    pid, control, payload = int(frames[1]), frames[2].decode(), frames[3]
                    self.bytes_received += len(payload)
                    self.nsync += 1
                    if pid in self.params and not locks.get(pid, False) and len(running) < nactors:
                        locks[pid] = True
                        running[pool.submit(handle_sync_from_worker, self.params[pid], control, payload)] = (pid, ident)
                    else:
                        sync_requests.append((pid, ident, control, payload))
                elif mtype == kStop:
                    nstop += 1
                else:
                    log.error("server: unknown message type %d", mtype)
        finally:
            pool.shutdown(wait=True)
            binder.close()
            dt = max(time.monotonic() - t_start, 1e-9)
            log.info("server %d: %d syncs, %.1f MB received, %.1f MB/s", self.server_id, self.nsync,
                     self.bytes_received / 1e6, self.bytes_received / 1e6 / dt)
