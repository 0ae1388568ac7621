"""``singa`` entry point: one process = one worker or one parameter server, chosen by ``--procsID``.

Reference: src/main.cc:13-62 (gflags ``--procsID --hostfile --cluster_conf --model_conf --v``) — ids
``[0, nworkers)`` are workers, ``[nworkers, nworkers+nservers)`` servers.  Extra flags select the B200 data
plane (``--push_pull fused|nccl|zmq``) and the compute backend.  Under ``torch.distributed.run`` the process id
defaults to ``$RANK`` and the GPU to ``$LOCAL_RANK``.
"""
from __future__ import annotations

import argparse
import logging
import os
import sys


def _flag(parser, name, **kw):
    # accept both gflags style (--procsID=1, -procsID 1) and argparse style
    parser.add_argument(f"--{name}", f"-{name}", **kw)


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="singa", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    _flag(ap, "procsID", type=int, default=int(os.environ.get("RANK", "0")), help="global process ID")
    _flag(ap, "hostfile", default="examples/mnist/hostfile")
    _flag(ap, "cluster_conf", default="examples/mnist/cluster.conf")
    _flag(ap, "model_conf", default="examples/mnist/mlp.conf")
    _flag(ap, "topology_config", default="examples/mnist/topology.conf")
    _flag(ap, "server_threads", type=int, default=8)
    _flag(ap, "client_threads", type=int, default=8)
    _flag(ap, "v", type=int, default=int(os.environ.get("GLOG_v", "0")), help="verbosity (3 = per-message logs)")
    _flag(ap, "push_pull", default="", help="zmq | nccl | fused (default: cluster.conf push_pull, zmq on CPU)")
    _flag(ap, "backend", default="", help="native | torch compute backend (default: native on a GPU)")
    _flag(ap, "device", default="", help="cpu | cuda | cuda:N")
    _flag(ap, "seed", type=int, default=0)
    _flag(ap, "watchdog_s", type=float, default=float(os.environ.get("SINGA_B200_WATCHDOG", "300")),
          help="terminate when no step completes / a peer rank is silent for this many seconds (0 = off)")
    _flag(ap, "debug_attach", action="store_true", help="spin until a debugger attaches (common.cc Debug())")
    return ap


def registry_classes(param_type: str) -> None:
    """``RegistryClasses`` (src/main.cc:24-31): users add their own Layer / Param classes here."""
    from .neuralnet import registry_layers, registry_param
    registry_layers()
    registry_param(param_type)


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    os.environ["SINGA_B200_WATCHDOG"] = str(args.watchdog_s)
    logging.basicConfig(level=logging.DEBUG if args.v >= 3 else (logging.INFO if args.v >= 1 else logging.ERROR),
                        format="%(levelname).1s%(asctime)s %(name)s] %(message)s", datefmt="%m%d %H:%M:%S")
    from .cluster import Cluster
    from .proto import schema as S
    from .utils.common import debug_attach_spin, read_proto_from_text_file
    if args.debug_attach:
        debug_attach_spin()
    pcluster = read_proto_from_text_file(args.cluster_conf, S.ClusterProto())
    cluster = Cluster.get(pcluster, args.hostfile, args.procsID)
    model = read_proto_from_text_file(args.model_conf, S.ModelProto())
    registry_classes(model.updater.param_type)

    if cluster.am_i_server():
        from .server import Server
        Server(cluster).run()
        return 0

    import torch
    from . import device
    dev = args.device
    if not dev:
        dev = f"cuda:{int(os.environ.get('LOCAL_RANK', '0'))}" if torch.cuda.is_available() else "cpu"
    if dev.startswith("cuda"):
        torch.cuda.set_device(torch.device(dev))
    device.set_context(dev, args.backend or None)
    push_pull = args.push_pull or (pcluster.push_pull if cluster.nservers > 0 else "")
    if cluster.nservers > 0 and dev.startswith("cuda") and not cluster.synchronous and push_pull in ("", "fused"):
        push_pull = "async" if not args.push_pull else push_pull    # asynchronous groups are the reference's default mode
    if cluster.nservers > 0 and not dev.startswith("cuda") and push_pull in ("fused", "nccl", "async", ""):
        push_pull = "zmq"
    if push_pull in ("fused", "nccl", "async") and cluster.nworkers > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", str(pcluster.start_port + 50))
            dist.init_process_group("nccl", rank=args.procsID, world_size=cluster.nworkers,
                                    device_id=torch.device(dev))
    from .worker import Worker
    backend = push_pull if cluster.nservers > 0 else "local"
    Worker(cluster, backend=backend, seed=args.seed).start(model)
    return 0


if __name__ == "__main__":
    sys.exit(main())
