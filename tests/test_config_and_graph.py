"""Job configs (field-by-field golden parse of the reference's shipped files), Graph, NeuralNet construction and
the partitioner goldens of src/test/test_neuralnet.cc:109-139."""
import json
import os

import pytest

from singa_b200.cluster import Cluster
from singa_b200.graph import Graph
from singa_b200.neuralnet import NeuralNet
from singa_b200.proto import schema as S
from singa_b200.utils.common import (int_vec_to_string, read_proto_from_binary_file, read_proto_from_text_file,
                                     write_proto_to_binary_file, write_proto_to_text_file)

REF = "/root/reference/examples/mnist"
have_ref = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not mounted")


@have_ref
def test_parse_reference_mlp_conf():
    m = read_proto_from_text_file(f"{REF}/mlp.conf", S.ModelProto())
    assert m.name == "deep-big-simple-mlp" and m.train_steps == 60000
    assert (m.test_steps, m.test_frequency, m.display_frequency, m.prefetch) == (10, 30, 30, True)
    u = m.updater
    assert S.enum_name("UpdaterProto.Type", u.type) == "kSGD"
    assert S.enum_name("UpdaterProto.ChangeProto", u.learning_rate_change_method) == "kStep"
    assert abs(u.base_learning_rate - 0.001) < 1e-9 and abs(u.gamma - 0.997) < 1e-7
    assert (u.learning_rate_change_frequency, u.sync_frequency, u.warmup_steps) == (60, 8, 60)
    assert u.param_type == "Elastic" and abs(u.moving_rate - 0.9) < 1e-7
    layers = m.neuralnet.layer
    assert len(layers) == 16
    assert [l.name for l in layers[:4]] == ["data", "data", "mnist", "label"]
    assert list(layers[0].exclude) == [S.kTest] and list(layers[1].exclude) == [S.kTrain]
    assert layers[0].data_param.batchsize == 1000 and layers[0].data_param.random_skip == 10000
    assert abs(layers[2].mnist_param.norm_a - 127.5) < 1e-6 and layers[2].mnist_param.norm_b == 1
    outs = [l.inner_product_param.num_output for l in layers if l.type == "kInnerProduct"]
    assert outs == [2500, 2000, 1500, 1000, 500, 10]
    p = layers[4].param[0]
    assert p.name == "weight" and S.enum_name("ParamProto.InitMethod", p.init_method) == "kUniform"
    assert abs(p.low + 0.05) < 1e-7 and abs(p.high - 0.05) < 1e-7
    # defaults of fields the file does not set
    assert p.learning_rate_multiplier == 1 and p.split_threshold == 5000000 and u.hogwild is True


@have_ref
def test_parse_reference_conv_cluster_topology():
    c = read_proto_from_text_file(f"{REF}/conv.conf", S.ModelProto())
    conv = [l for l in c.neuralnet.layer if l.type == "kConvolution"]
    assert [(l.convolution_param.num_filters, l.convolution_param.kernel, l.convolution_param.stride) for l in conv] \
        == [(20, 5, 1), (50, 5, 1)]
    pools = [l for l in c.neuralnet.layer if l.type == "kPooling"]
    assert all(l.pooling_param.kernel == 2 and l.pooling_param.stride == 2 for l in pools)
    assert c.debug is True and abs(c.updater.momentum - 0.9) < 1e-7 and abs(c.updater.weight_decay - 0.0005) < 1e-9
    cl = read_proto_from_text_file(f"{REF}/cluster.conf", S.ClusterProto())
    assert (cl.nworkers, cl.nservers, cl.nthreads_per_server) == (2, 1, 4)
    assert cl.start_port == 6723 and cl.nprocs_per_group == 1 and cl.synchronous is False and cl.bandwidth == 100
    t = read_proto_from_text_file(f"{REF}/topology.conf", S.Topology())
    assert (t.nservers, t.nworker_groups, t.port, t.server_threads, t.worker_threads) == (1, 1, 60000, 1, 1)
    assert t.server_group[0].id == 0 and t.server_group[0].sync_interval == 8


def test_proto_text_and_binary_roundtrip(tmp_path):
    from singa_b200.models import configs
    for name, fn in configs.CONFIGS.items():
        m = fn()
        write_proto_to_text_file(m, str(tmp_path / f"{name}.conf"))
        m2 = read_proto_from_text_file(str(tmp_path / f"{name}.conf"), S.ModelProto())
        assert m2 == m
        write_proto_to_binary_file(m, str(tmp_path / f"{name}.bin"))
        assert read_proto_from_binary_file(str(tmp_path / f"{name}.bin"), S.ModelProto()) == m
    assert int_vec_to_string([8, 20, 24, 24]) == "(8, 20, 24, 24, )"


def test_cluster_arithmetic(tmp_path):
    cp = S.ClusterProto()
    cp.nworkers, cp.nservers, cp.nprocs_per_group, cp.nthreads_per_procs = 8, 2, 2, 2
    cp.workspace = str(tmp_path)
    c = Cluster(cp, "", 5)
    assert c.am_i_worker() and not c.am_i_server()
    assert (c.groupid, c.ngroups, c.group_procsid, c.nthreads_per_group) == (2, 4, 1, 4)
    assert c.group_threadid(1) == 3 and c.group_procsid_of(3) == 1
    assert (c.pub_port, c.router_port, c.pull_port(2)) == (6723, 6724, 6727)
    s = Cluster(cp, "", 9)
    assert s.am_i_server() and s.server_addr() == "127.0.0.1"
    assert os.path.isdir(os.path.join(str(tmp_path), "vis"))


def test_graph_sort_and_json():
    g = Graph()
    for n in "abcde":
        g.add_node(n)
    for s, d in [("a", "b"), ("a", "c"), ("b", "d"), ("c", "d"), ("d", "e")]:
        g.add_edge(s, d)
    g.sort()
    assert g.check_acyclic() and g.nodes[0].name == "a" and g.nodes[-1].name == "e"
    d = json.loads(g.to_string({"a": "(1, )"}))
    assert d["directed"] == 1 and len(d["nodes"]) == 5 and len(d["links"]) == 5
    assert d["nodes"][0] == {"id": "a(1, )", "color": "red", "shape": "box"}
    g.insert_bridge_node(g.node("d"), g.node("e"))
    assert {"s-d-e", "d-d-e"} <= set(g.name2node)


def convnet():
    p = S.NetProto()
    l = p.layer.add(); l.name, l.type = "data", "kShardData"
    l.data_param.batchsize = 8; l.data_param.source = "synthetic"; l.data_param.synthetic_shape.extend([10, 10])
    l = p.layer.add(); l.name, l.type = "mnist", "kMnistImage"; l.srclayers.append("data")
    l = p.layer.add(); l.name, l.type = "label", "kLabel"; l.srclayers.append("data")
    l = p.layer.add(); l.name, l.type = "conv1", "kConvolution"; l.srclayers.append("mnist")
    l.param.add(); l.param.add(); l.convolution_param.num_filters = 8; l.convolution_param.kernel = 2
    l = p.layer.add(); l.name, l.type = "relu1", "kReLU"; l.srclayers.append("conv1")
    l = p.layer.add(); l.name, l.type = "pool1", "kPooling"; l.srclayers.append("relu1")
    l.pooling_param.kernel = 4; l.pooling_param.stride = 2
    l = p.layer.add(); l.name, l.type = "fc1", "kInnerProduct"; l.srclayers.append("pool1")
    l.param.add(); l.param.add(); l.inner_product_param.num_output = 10
    l = p.layer.add(); l.name, l.type = "loss", "kSoftmaxLoss"; l.srclayers.extend(["fc1", "label"])
    return p


def test_no_partition_golden():
    net = NeuralNet(convnet())
    assert len(net.layers) == 8
    assert net.layers[0].name == "data" and net.layers[-1].name == "loss"
    shapes = {l.name: l.shape() for l in net.layers}
    assert shapes["conv1"] == [8, 8, 9, 9] and shapes["pool1"] == [8, 8, 4, 4] and shapes["fc1"] == [8, 10]
    assert [p.id for p in net.params] == [0, 1, 2, 3]
    assert net.name2layer("conv1").weight.shape == [8, 4] and net.name2layer("fc1").weight.shape == [128, 10]
    d = json.loads(net.to_string())
    assert len(d["nodes"]) == 8


def test_data_partition_golden_28_layers():
    p = convnet()
    p.partition_type = S.kDataPartition
    net = NeuralNet(p, 3)
    assert len(net.layers) == 28 and net.layers[0].name == "data"
    names = {l.name for l in net.layers}
    assert {"slice-mnist", "slice-label", "conv1-00", "conv1-02", "loss-01"} <= names
    assert sum(1 for l in net.layers if l.is_bridgesrclayer) == 4
    assert net.name2layer("conv1-02").shape() == [4, 8, 9, 9]      # 8 = 2 + 2 + 4 (remainder on the last)
    assert net.name2layer("conv1-00").shape() == [2, 8, 9, 9]


def test_layer_and_hybrid_partition_construct():
    p = convnet()
    p.partition_type = S.kLayerPartition
    net = NeuralNet(p, 2)
    assert net.name2layer("conv1-00").shape() == [8, 4, 9, 9]
    assert any(l.type_id == "kConcate" for l in net.layers)
    p = convnet()
    p.layer[-2].partition_type = S.kDataPartition
    p.layer[-1].partition_type = S.kDataPartition
    p.partition_type = S.kLayerPartition
    net = NeuralNet(p, 2)
    assert any(l.type_id == "kSlice" for l in net.layers) and any(l.type_id == "kConcate" for l in net.layers)


def test_share_weights_and_phase_nets():
    from singa_b200.models import configs
    from singa_b200.worker import Worker
    m = configs.mnist_mlp(batch=8, hidden=(16,))
    # add a test-phase data layer like examples/mnist/mlp.conf
    m.neuralnet.layer[0].exclude.append(S.kTest)
    tl = m.neuralnet.layer.add()
    tl.CopyFrom(m.neuralnet.layer[0])
    del tl.exclude[:]
    tl.exclude.append(S.kTrain)
    m.test_steps, m.test_frequency = 2, 5
    w = Worker()
    w.build(m)
    assert w.test_net is not None and len(w.test_net.layers) == len(w.train_net.layers)
    for a, b in zip(w.test_net.params, w.train_net.params):
        assert a.data.data_ptr() == b.data.data_ptr()
    vals = w.executor.test(w.test_net, 2, disperf=False)
    assert len(vals[0]) == 2


def test_contrastive_divergence_is_rejected_loudly():
    """ModelProto.alg = kContrastiveDivergence is declared by the reference without any implementation; training such a
    model by back-propagation would be silently wrong, so building it fails."""
    import pytest
    from singa_b200.models import configs
    from singa_b200.worker import Worker
    m = configs.mnist_mlp(batch=8, hidden=(16,))
    m.alg = S.enum_value("ModelProto.GradCalcAlg", "kContrastiveDivergence")
    with pytest.raises(NotImplementedError, match="kContrastiveDivergence"):
        Worker().build(m)
