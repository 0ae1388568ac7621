"""Stack A end to end on CPU with real OS processes over pyzmq (the reference's own multi-process-on-one-host
trick, examples/mnist/run.sh): 1 server + 2 worker groups train a tiny MLP with Elastic / RandomSync."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

MODEL = textwrap.dedent("""
    name: "tiny-mlp"
    train_steps: 24
    display_frequency: 8
    updater{ base_learning_rate: 0.05 type: kSGD param_type: "%s" sync_frequency: 4 moving_rate: 0.9 warmup_steps: 4 }
    neuralnet {
      layer { name: "data" type: "kShardData" data_param { source: "synthetic" batchsize: 32 synthetic_shape: 12 synthetic_shape: 12 } }
      layer { name: "mnist" type: "kMnistImage" srclayers: "data" mnist_param { norm_a: 127.5 norm_b: 1 } }
      layer { name: "label" type: "kLabel" srclayers: "data" }
      layer { name: "fc1" type: "kInnerProduct" srclayers: "mnist" inner_product_param { num_output: 32 }
              param { name: "weight" init_method: kUniform low: -0.05 high: 0.05 } param { name: "bias" init_method: kUniform low: -0.05 high: 0.05 } }
      layer { name: "tanh1" type: "kTanh" srclayers: "fc1" }
      layer { name: "fc2" type: "kInnerProduct" srclayers: "tanh1" inner_product_param { num_output: 10 }
              param { name: "weight" init_method: kUniform low: -0.05 high: 0.05 } param { name: "bias" init_method: kUniform low: -0.05 high: 0.05 } }
      layer { name: "loss" type: "kSoftmaxLoss" srclayers: "fc2" srclayers: "label" }
    }
""")


@pytest.mark.parametrize("param_type,port", [("Elastic", 17723), ("RandomSync", 17823)])
def test_two_workers_one_server(tmp_path, param_type, port):
    (tmp_path / "model.conf").write_text(MODEL % param_type)
    (tmp_path / "cluster.conf").write_text(
        f'nworkers: 2\nnservers: 1\nnthreads_per_server: 2\nstart_port: {port}\nworkspace: "{tmp_path}"\n')
    env = dict(os.environ, PYTHONPATH=ROOT, SINGA_B200_BACKEND="torch", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, "-m", "singa_b200.launch", f"--cluster_conf={tmp_path}/cluster.conf",
                        f"--model_conf={tmp_path}/model.conf", f"--workdir={tmp_path}", "--timeout=120",
                        "--device=cpu"], env=env, capture_output=True, text=True, timeout=180, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "Training at step 16" in r.stderr
    assert "Output from loss layer" in r.stderr


def test_job_shorter_than_the_warmup_still_stops_the_server(tmp_path):
    """train_steps < warmup_steps: the job ends right after the initial kPut of every parameter; the kStop queued behind
    those payloads must still reach the server (it was dropped when the worker closed its socket with linger 0, and
    the launcher then waited for the server forever)."""
    big = MODEL.replace("train_steps: 24", "train_steps: 3").replace("warmup_steps: 4", "warmup_steps: 60") \
               .replace("num_output: 32", "num_output: 4096") \
               .replace("synthetic_shape: 12 synthetic_shape: 12", "synthetic_shape: 64 synthetic_shape: 64")
    # 4096 x 4096 weights: 67 MB of kPut payload sit in the socket queue when the stop message is sent
    (tmp_path / "model.conf").write_text(big % "Elastic")
    (tmp_path / "cluster.conf").write_text(
        f'nworkers: 1\nnservers: 1\nnthreads_per_server: 2\nstart_port: 17923\nworkspace: "{tmp_path}"\n')
    env = dict(os.environ, PYTHONPATH=ROOT, SINGA_B200_BACKEND="torch", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, "-m", "singa_b200.launch", f"--cluster_conf={tmp_path}/cluster.conf",
                        f"--model_conf={tmp_path}/model.conf", "--timeout=90", "--device=cpu"],
                       env=env, capture_output=True, text=True, timeout=150, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    assert (tmp_path / "hostfile").exists()          # generated hostfile lives in the job workspace by default


def test_router_ping_pong_barrier():
    """dist_test/test_router.cc: the server's bind returns only after every worker pinged."""
    import threading
    from singa_b200.router import Router
    port = 17911
    res = {}

    def server():
        r = Router(port)
        res["bind"] = r.bind("127.0.0.1", 2, timeout_s=20)
        res["msg"] = r.recv(timeout_ms=5000)
        r.reply([b"ack"])
        r.close()

    t = threading.Thread(target=server)
    t.start()
    workers = [Router(port) for _ in range(2)]
    oks = [None, None]

    def conn(i):   # the barrier releases only when BOTH workers have pinged: connect concurrently
        oks[i] = workers[i].connect("127.0.0.1", retries=40, interval_ms=300)
    cts = [threading.Thread(target=conn, args=(i,)) for i in range(2)]
    for c in cts:
        c.start()
    for c in cts:
        c.join(30)
    workers[0].send([b"hello"], 0)
    ack = workers[0].recv(timeout_ms=5000)
    t.join(30)
    for w in workers:
        w.close()
    assert oks == [True, True] and res["bind"] is True
    assert res["msg"] == [b"hello"] and ack == [b"ack"]
