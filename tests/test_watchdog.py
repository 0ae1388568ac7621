"""Failure detection (utils/watchdog.py): local stall detector and peer heartbeat over the rendezvous store."""
import os
import subprocess
import sys
import textwrap
import time

from singa_b200.utils.watchdog import Watchdog


def test_watchdog_fires_on_stall_and_not_on_progress():
    counter = {"v": 0}
    fired = []
    wd = Watchdog(lambda: counter["v"], timeout_s=0.4, on_stall=lambda s, idle: fired.append((s, idle)), poll_s=0.05).start()
    for _ in range(12):                 # progress every 100 ms: must stay quiet
        time.sleep(0.1)
        counter["v"] += 1
    assert not fired
    time.sleep(0.8)                     # stall
    wd.stop()
    assert fired and fired[0][0] == counter["v"] and fired[0][1] > 0.4


def test_watchdog_pause():
    fired = []
    wd = Watchdog(lambda: 0, timeout_s=0.2, on_stall=lambda s, idle: fired.append(s), poll_s=0.05).start()
    wd.pause(True)
    time.sleep(0.6)
    assert not fired
    wd.stop()


_RANK_SCRIPT = textwrap.dedent("""
    import os, sys, time
    import torch.distributed as dist
    from torch.distributed.distributed_c10d import _get_default_store
    from singa_b200.utils.watchdog import Heartbeat
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    step = {"v": 0}
    hb = Heartbeat(_get_default_store(), rank, world, lambda: step["v"], interval_s=0.2, dead_after_s=1.5).start()
    dist.barrier()
    for i in range(200):
        step["v"] = i
        time.sleep(0.05)
        if rank == 1 and i == 20:
            os._exit(3)              # rank 1 dies without saying goodbye
    hb.stop()
    sys.exit(0)
""")


def test_heartbeat_detects_dead_peer(tmp_path):
    script = tmp_path / "rank.py"
    script.write_text(_RANK_SCRIPT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(29650 + os.getpid() % 200), WORLD_SIZE="2",
               PYTHONPATH=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stderr=subprocess.PIPE)
             for r in range(2)]
    t0 = time.time()
    codes = [p.wait(timeout=60) for p in procs]
    err0 = procs[0].stderr.read().decode()
    assert codes[1] == 3
    assert codes[0] == 87, err0               # the survivor noticed and left with the heartbeat's exit code
    assert "rank 1 has been silent" in err0
    assert time.time() - t0 < 30
