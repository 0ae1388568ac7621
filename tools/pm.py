#!/usr/bin/env python
"""``pm`` — the Stack-B parameter-manager benchmark (reference: src/test/test_pm.cc, launched by
examples/mnist/run.sh and mesos/pm_run.sh; numbers in benchmarks/*).

  python tools/pm.py --node_id I --topology_config T [--hostfile H]   one process, role by id (servers first)
  python tools/pm.py --local --nservers 2 --nclients 4 --threads 4    spawn everything on this host and report

Reports the mean milliseconds of one round (Update x12 + Collect x12 over the 12 MNIST-MLP tensors, 47.9 MB each
way) per client thread — the quantity the reference's charts label "time (ms)".
"""
import argparse
import json
import logging
import multiprocessing as mp
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def read_hosts(hostfile, n):
    hosts = []
    if hostfile and os.path.exists(hostfile):
        hosts = [l.strip() for l in open(hostfile) if l.strip()]
    while len(hosts) < n:
        hosts.append("127.0.0.1")
    return hosts


def run_node(node_id, topology, hosts, rounds, scale, result_q=None):
    from singa_b200.pm.pm_client import MNIST_MLP_SIZES, SingaClient, client_thread
    from singa_b200.pm.pm_server import SingaServer
    if node_id < topology.nservers:
        SingaServer(node_id, topology, hosts).start()
        return
    client = SingaClient(node_id, topology, hosts)
    pt = threading.Thread(target=client.proxy, daemon=True)
    pt.start()
    client._ready.wait(10)
    sizes = [max(1, int(s * scale)) for s in MNIST_MLP_SIZES]
    out = {}
    nthreads = max(1, topology.worker_threads)
    ths = [threading.Thread(target=client_thread,
                            args=(client, t, sizes, rounds, client.is_group_seeder and t == 0, out))
           for t in range(nthreads)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    means = {tid: sum(v) / max(len(v), 1) for tid, v in out.items()}
    if result_q is not None:
        result_q.put((node_id, means))
    else:
        print(json.dumps({"node": node_id, "ms_per_round": means}))
    client.stop()
    pt.join(2)


def local_bench(nservers, nclients, threads, server_threads, ngroups, sync_interval, rounds, scale, port):
    from singa_b200.proto import schema as S
    topo = S.Topology()
    topo.nservers, topo.nworker_groups = nservers, nclients
    topo.port, topo.server_threads, topo.worker_threads = port, server_threads, threads
    for g in range(ngroups):
        sg = topo.server_group.add()
        sg.id, sg.sync_interval = g, sync_interval
        if ngroups > 1:
            sg.neighbor.append((g + 1) % ngroups)
    hosts = ["127.0.0.1"] * (nservers + nclients)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    servers = [ctx.Process(target=run_node, args=(i, topo, hosts, rounds, scale)) for i in range(nservers)]
    for p in servers:
        p.start()
    time.sleep(0.5)
    clients = [ctx.Process(target=run_node, args=(nservers + i, topo, hosts, rounds, scale, q))
               for i in range(nclients)]
    for p in clients:
        p.start()
    res = [q.get(timeout=600) for _ in clients]
    for p in clients:
        p.join(10)
    for p in servers:
        p.terminate()
    allms = [ms for _, m in res for ms in m.values()]
    return {"nservers": nservers, "nclients": nclients, "threads": threads, "server_groups": ngroups,
            "sync_interval": sync_interval, "rounds": rounds, "payload_scale": scale,
            "mean_ms_per_round": sum(allms) / len(allms), "max_ms_per_round": max(allms)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--node_id", type=int, default=0)
    ap.add_argument("--topology_config", default="examples/mnist/topology.conf")
    ap.add_argument("--hostfile", default="examples/mnist/hostfile")
    ap.add_argument("--rounds", type=int, default=200)
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the 12 tensors (tests)")
    ap.add_argument("--local", action="store_true")
    ap.add_argument("--nservers", type=int, default=1)
    ap.add_argument("--nclients", type=int, default=1)
    ap.add_argument("--threads", type=int, default=1)
    ap.add_argument("--server_threads", type=int, default=1)
    ap.add_argument("--server_groups", type=int, default=1)
    ap.add_argument("--sync_interval", type=int, default=8)
    ap.add_argument("--port", type=int, default=60000)
    ap.add_argument("--v", type=int, default=0)
    args = ap.parse_args()
    logging.basicConfig(level=logging.DEBUG if args.v >= 3 else logging.ERROR)
    if args.local:
        print(json.dumps(local_bench(args.nservers, args.nclients, args.threads, args.server_threads,
                                     args.server_groups, args.sync_interval, args.rounds, args.scale, args.port)))
        return 0
    from singa_b200.proto import schema as S
    from singa_b200.utils.common import read_proto_from_text_file
    topo = read_proto_from_text_file(args.topology_config, S.Topology())
    hosts = read_hosts(args.hostfile, topo.nservers + topo.nworker_groups)
    run_node(args.node_id, topo, hosts, args.rounds, args.scale)
    return 0


if __name__ == "__main__":
    sys.exit(main())
