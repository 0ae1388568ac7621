"""Cluster: the job's topology arithmetic (who is a worker, who is a server, groups, ports, folders).

Reference: include/utils/cluster.h:35-124, src/utils/cluster.cc:11-52.  Process ids ``[0, nworkers)`` are
workers and ``[nworkers, nworkers + nservers)`` servers; worker group = ``procsID // nprocs_per_group``.
On the single 8xB200 box one worker *group* is one GPU rank (or ``nprocs_per_group`` ranks) and the server
shards are co-located on the same GPUs (``server_shard_of_rank``); the host names of the hostfile are kept
for the ZeroMQ control plane and default to 127.0.0.1.
"""
from __future__ import annotations

import os
import socket
from typing import List, Optional

from .proto import schema as S


class Cluster:
    _instance: Optional["Cluster"] = None

    def __init__(self, cluster_proto, hostfile: str = "", procsid: int = 0):
        self.proto = S.ClusterProto()
        self.proto.CopyFrom(cluster_proto)
        self.global_procsid = int(procsid)
        self.hostname = socket.gethostname()
        self.addr: List[str] = []
        if hostfile and os.path.exists(hostfile) and (self.nworkers > 1 or self.nservers > 0):
            with open(hostfile) as f:
                self.addr = [ln.strip() for ln in f if ln.strip()]
        while len(self.addr) < self.nworkers + self.nservers:
            self.addr.append("127.0.0.1")
        self.setup_folders()

    # -- singleton (Cluster::Get) ------------------------------------------------------------------
    @classmethod
    def get(cls, cluster_proto=None, hostfile: str = "", procsid: int = 0) -> "Cluster":
        if cls._instance is None:
            if cluster_proto is None:
                raise RuntimeError("the first call to Cluster.get must provide the cluster config")
            cls._instance = Cluster(cluster_proto, hostfile, procsid)
        return cls._instance

    @classmethod
    def reset(cls) -> None:
        cls._instance = None

    def setup_folders(self) -> None:
        for sub in (self.proto.vis_subfolder, self.proto.log_subfolder, self.proto.checkpoint_subfolder):
            try:
                os.makedirs(os.path.join(self.workspace, sub), exist_ok=True)
            except OSError:
                pass   # read-only / non-existent workspace root: folders are created lazily by writers

    # -- roles ---------------------------------------------------------------------------------------
    @property
    def nworkers(self) -> int:
        return self.proto.nworkers

    @property
    def nservers(self) -> int:
        return self.proto.nservers

    def am_i_server(self) -> bool:
        return self.nworkers <= self.global_procsid < self.nworkers + self.nservers

    def am_i_worker(self) -> bool:
        return 0 <= self.global_procsid < self.nworkers

    @property
    def nprocs_per_group(self) -> int:
        return max(1, self.proto.nprocs_per_group)

    @property
    def nthreads_per_procs(self) -> int:
        return max(1, self.proto.nthreads_per_procs)

    @property
    def nthreads_per_server(self) -> int:
        return max(1, self.proto.nthreads_per_server)

    @property
    def groupid(self) -> int:
        return self.global_procsid // self.nprocs_per_group

    @property
    def ngroups(self) -> int:
        return max(1, self.nworkers // self.nprocs_per_group)

    @property
    def nthreads_per_group(self) -> int:
        return self.nthreads_per_procs * self.nprocs_per_group

    @property
    def group_procsid(self) -> int:
        return self.global_procsid % self.nprocs_per_group

    def group_threadid(self, local_threadid: int) -> int:
        return self.group_procsid * self.nthreads_per_procs + local_threadid

    def group_procsid_of(self, group_threadid: int) -> int:
        return group_threadid // self.nthreads_per_procs

    # -- addresses / ports ----------------------------------------------------------------------------
    def server_addr(self, server_procsid: Optional[int] = None) -> str:
        if server_procsid is None:
            assert self.am_i_server()
            return self.addr[self.global_procsid]
        assert 0 <= server_procsid < self.nservers
        return self.addr[self.nworkers + server_procsid]

    def group_thread_addr(self, group_threadid: int) -> str:
        assert 0 <= group_threadid < self.nthreads_per_group
        base = self.global_procsid - self.group_procsid
        return self.addr[base + self.group_procsid_of(group_threadid)]

    @property
    def pub_port(self) -> int:
        return self.proto.start_port

    @property
    def router_port(self) -> int:
        return self.proto.start_port + 1

    def pull_port(self, k: int) -> int:
        return self.proto.start_port + 2 + k

    def server_router_port(self, server_procsid: int) -> int:
        """Same-host deployments (every hostfile line equal, the reference's own smoke-test trick) need one
        port per server; distinct hosts keep the reference's single ``router_port``."""
        same_host = len(set(self.addr)) == 1
        return self.router_port + (100 + server_procsid if same_host else 0)

    # -- misc -------------------------------------------------------------------------------------------
    @property
    def synchronous(self) -> bool:
        return self.proto.synchronous

    @property
    def workspace(self) -> str:
        return self.proto.workspace

    @property
    def visualization_folder(self) -> str:
        return os.path.join(self.workspace, self.proto.vis_subfolder)

    @property
    def checkpoint_folder(self) -> str:
        return os.path.join(self.workspace, self.proto.checkpoint_subfolder)

    @property
    def log_folder(self) -> str:
        return os.path.join(self.workspace, self.proto.log_subfolder)

    @property
    def bandwidth(self) -> float:
        return self.proto.bandwidth

    # B200 mapping: shard s of the parameter arena lives on GPU rank s % world
    def server_shard_of_rank(self, rank: int, world: int) -> int:
        return rank % max(1, min(world, max(self.nservers, 1)))
