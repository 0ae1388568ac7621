"""End-to-end CPU scenarios through the `singa` entry point (bin/singa -> singa_b200.cli): periodic test passes,
checkpoints, resume from the workspace, and a data-partitioned net driven by nthreads_per_procs."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _conf(train_steps, extra="", partition=""):
    src = open(os.path.join(ROOT, "examples", "mnist", "conv.conf")).read()
    src = re.sub(r"train_steps: \d+", f"train_steps: {train_steps}", src)
    src = re.sub(r"display_frequency: \d+", "display_frequency: 4", src)
    src = re.sub(r"batchsize: \d+", "batchsize: 8", src)
    if partition:
        src = src.replace("neuralnet {", "neuralnet {\n  partition_type: %s" % partition, 1)
    return src + "\n" + extra


def _run(tmp_path, model, cluster):
    (tmp_path / "model.conf").write_text(model)
    (tmp_path / "cluster.conf").write_text(cluster)
    env = dict(os.environ, PYTHONPATH=ROOT, SINGA_B200_SYNTHETIC="1", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([os.path.join(ROOT, "bin", "singa"), "-procsID=0", f"-cluster_conf={tmp_path}/cluster.conf",
                        f"-model_conf={tmp_path}/model.conf", "--device=cpu", "--watchdog_s=120"],
                       env=env, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    return r.stderr


def test_test_passes_checkpoints_and_resume(tmp_path):
    ws = tmp_path / "ws"
    cluster = f'nworkers: 1\nnservers: 0\nworkspace: "{ws}"\n'
    extra = "test_frequency: 5\ntest_steps: 2\ncheckpoint_frequency: 6\n"
    log = _run(tmp_path, _conf(12, extra), cluster)
    assert "Test at step 5" in log and "Test at step 10" in log
    assert sorted(os.listdir(ws / "checkpoint")) == ["step12.bin", "step6.bin"]
    assert (ws / "vis" / "train.json").exists() and (ws / "vis" / "test.json").exists()
    log2 = _run(tmp_path, _conf(16, extra), cluster)                 # same workspace: continues from step 12
    assert "Resumed from" in log2 and "at step 12" in log2
    assert "Training at step 12" in log2 and "Training at step 0" not in log2


def test_data_partitioned_net_with_two_group_threads(tmp_path):
    cluster = f'nworkers: 1\nnservers: 0\nnthreads_per_procs: 2\nworkspace: "{tmp_path}/ws"\n'
    log = _run(tmp_path, _conf(6, partition="kDataPartition"), cluster)
    assert "loss-00" in log and "loss-01" in log                     # one loss replica per partition
    losses = [float(x) for x in re.findall(r"loss-00 layer 0 : ([0-9.]+)", log)]
    assert len(losses) >= 2 and all(0 < v < 5 for v in losses)
