"""Single-box launcher and topology discovery — replaces the Mesos scheduler and the ssh fan-out scripts
(mesos/SingaFramework.cc:40-180, examples/mnist/run.sh, script/node.sh).

``python -m singa_b200.launch --cluster_conf C --model_conf M`` reads the cluster config, discovers the GPUs of
the box (count, names, peer-access matrix), writes the hostfile (all 127.0.0.1), and starts one ``singa`` process
per worker (GPU i for worker i) and per ZeroMQ server, waits for them, and propagates the first failure.
``--discover`` only prints the topology as JSON.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time
from typing import List


def discover() -> dict:
    """GPU inventory + P2P matrix (NVLink/NVSwitch peers report access = true for every pair on an HGX box)."""
    info = {"gpus": [], "p2p": [], "nvml": False}
    try:
        import torch
        n = torch.cuda.device_count() if torch.cuda.is_available() else 0
        for i in range(n):
            pr = torch.cuda.get_device_properties(i)
            info["gpus"].append({"index": i, "name": pr.name, "sms": pr.multi_processor_count,
                                 "memory_gb": round(pr.total_memory / 2**30, 1),
                                 "cc": f"{pr.major}.{pr.minor}"})
        info["p2p"] = [[bool(i == j or torch.cuda.can_device_access_peer(i, j)) for j in range(n)]
                       for i in range(n)]
    except Exception as e:  # noqa: BLE001
        info["error"] = str(e)
    try:
        import pynvml
        pynvml.nvmlInit()
        info["nvml"] = True
        info["driver"] = pynvml.nvmlSystemGetDriverVersion()
        links = []
        for i in range(len(info["gpus"])):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            active = 0
            for l in range(18):
                try:
                    if pynvml.nvmlDeviceGetNvLinkState(h, l):
                        active += 1
                except Exception:  # noqa: BLE001
                    break
            links.append(active)
        info["nvlinks_active"] = links
    except Exception:  # noqa: BLE001
        pass
    return info


def launch(cluster_conf: str, model_conf: str, extra: List[str], workdir: str = "", timeout: float = 0) -> int:
    from .proto import schema as S
    from .utils.common import read_proto_from_text_file
    pc = read_proto_from_text_file(cluster_conf, S.ClusterProto())
    nprocs = pc.nworkers + pc.nservers
    workdir = workdir or pc.workspace or "."          # the generated hostfile lives with the job's other artefacts
    os.makedirs(workdir, exist_ok=True)
    hostfile = os.path.join(workdir, "hostfile")
    with open(hostfile, "w") as f:
        f.write("\n".join(["127.0.0.1"] * nprocs) + "\n")
    topo = discover()
    ngpu = len(topo["gpus"])
    procs = []
    env_base = dict(os.environ)
    env_base.setdefault("MASTER_ADDR", "127.0.0.1")
    env_base.setdefault("MASTER_PORT", str(pc.start_port + 50))
    # servers first so their ROUTER sockets are bound before the workers start pinging
    order = list(range(pc.nworkers, nprocs)) + list(range(pc.nworkers))
    for pid in order:
        env = dict(env_base)
        if pid < pc.nworkers:
            env["RANK"], env["WORLD_SIZE"] = str(pid), str(pc.nworkers)
            env["LOCAL_RANK"] = str(pid % ngpu) if ngpu else "0"
        cmd = [sys.executable, "-m", "singa_b200.cli", f"--procsID={pid}", f"--hostfile={hostfile}",
               f"--cluster_conf={cluster_conf}", f"--model_conf={model_conf}", *extra]
        procs.append((pid, subprocess.Popen(cmd, env=env)))
    rc = 0
    t0 = time.monotonic()
    alive = dict(procs)
    while alive:
        for pid, p in list(alive.items()):
            r = p.poll()
            if r is not None:
                del alive[pid]
                if r != 0 and rc == 0:
                    rc = r
                    print(f"[launch] process {pid} exited with {r}; stopping the job", file=sys.stderr)
                    for q in alive.values():
                        q.terminate()
        if timeout and time.monotonic() - t0 > timeout:
            for q in alive.values():
                q.kill()
            return 124
        time.sleep(0.05)
    return rc


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="singa-launch")
    ap.add_argument("--discover", action="store_true")
    ap.add_argument("--cluster_conf", default="examples/mnist/cluster.conf")
    ap.add_argument("--model_conf", default="examples/mnist/mlp.conf")
    ap.add_argument("--workdir", default="", help="where the generated hostfile goes (default: the cluster workspace)")
    ap.add_argument("--timeout", type=float, default=0)
    args, extra = ap.parse_known_args(argv)
    if args.discover:
        print(json.dumps(discover(), indent=1))
        return 0
    return launch(args.cluster_conf, args.model_conf, extra, args.workdir, args.timeout)


if __name__ == "__main__":
    sys.exit(main())
