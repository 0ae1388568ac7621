"""Several executor threads sharing one parameter set inside a process (hogwild and aggregated modes)."""
import pytest
import torch

from singa_b200 import device
from singa_b200.cluster import Cluster
from singa_b200.models import configs
from singa_b200.proto import schema as S
from singa_b200.worker import HogwildGroup, Worker


@pytest.mark.parametrize("hogwild", [True, False])
def test_threads_share_parameters_and_learn(tmp_path, hogwild):
    device.set_context("cpu", "torch")
    m = configs.mnist_mlp(batch=32, train_steps=30, hidden=(32,))
    m.display_frequency = 0
    m.updater.base_learning_rate = 0.05
    m.updater.hogwild = hogwild
    cp = S.ClusterProto()
    cp.nworkers, cp.nservers, cp.nthreads_per_procs = 1, 0, 3
    cp.workspace = str(tmp_path)
    Cluster.reset()
    w = Worker(Cluster.get(cp, "", 0))
    w.build(m)
    w0 = w.train_net.params[0].data.clone()
    g = HogwildGroup(w, 3, hogwild)
    assert len(g.replicas) == 3
    assert g.replicas[1].params[0].data.data_ptr() == w.train_net.params[0].data.data_ptr()
    assert g.replicas[1].params[0].grad.data_ptr() != w.train_net.params[0].grad.data_ptr()
    g.run(30)
    nparams = len(w.train_net.params)
    assert g.updates == (30 * 3 * nparams if hogwild else 30 * nparams)
    assert not torch.equal(w.train_net.params[0].data, w0)
    losses = [float(n.losslayers[0].metric[0]) / 30 for n in g.replicas]
    assert all(l < 2.3 for l in losses), losses          # below the ln(10) of an untrained net
    Cluster.reset()
    device.reset_context()
