"""Stack B: ParamShard semantics and the pm benchmark topology (server groups, neighbour sync, requeue)."""
import json
import os
import subprocess
import sys

import numpy as np

from singa_b200.pm.param_shard import ParamShard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_param_shard_put_get_update_sync_now():
    sh = ParamShard(0, sync_interval=3)
    assert sh.get(7) is None and sh.update(7, b"") is None and not sh.is_local(7)
    a = np.arange(10, dtype=np.float32)
    sh.put(7, a.tobytes())
    assert sh.is_local(7)
    assert np.array_equal(np.frombuffer(sh.get(7), dtype=np.float32), a)
    fired = []
    for k in range(7):
        echo = sh.update(7, (a + k).tobytes())
        assert np.array_equal(np.frombuffer(echo, dtype=np.float32), a + k)     # replace-and-echo
        fired.append(sh.sync_now(7))
    assert fired == [False, False, True, False, False, True, False]           # every 3rd update, counter reset
    sh2 = ParamShard(1, 1, sync_rule="average")
    sh2.put(1, np.ones(4, dtype=np.float32).tobytes())
    out = np.frombuffer(sh2.sync_update(1, (3 * np.ones(4, dtype=np.float32)).tobytes()), dtype=np.float32)
    assert np.allclose(out, 2.0)


def test_pm_benchmark_two_server_groups():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "pm.py"), "--local", "--nservers", "2",
                        "--nclients", "2", "--threads", "2", "--server_threads", "2", "--server_groups", "2",
                        "--sync_interval", "2", "--rounds", "4", "--scale", "0.002", "--port", "61555"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["server_groups"] == 2 and out["mean_ms_per_round"] > 0
    assert "Traceback" not in r.stderr
