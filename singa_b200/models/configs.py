"""Job configs authored in the reference's ``ModelProto`` schema for the five BASELINE configurations.

``mnist_mlp`` restates examples/mnist/mlp.conf (784-2500-2000-1500-1000-500-10, tanh, batch 1000, SGD step
decay, Elastic sync) and ``lenet`` restates examples/mnist/conv.conf with a synthetic data source; the
CIFAR-10 convnet, AlexNet and wide-MLP configs do not exist in the reference tree and are
written here.
"""
from __future__ import annotations

from ..proto import schema as S


def _data(net, batch, shape, classes, name="data", exclude=None):
    l = net.layer.add()
    l.name, l.type = name, "kShardData"
    l.data_param.source = "synthetic"
    l.data_param.batchsize = batch
    l.data_param.synthetic_shape.extend(shape)
    l.data_param.synthetic_classes = classes
    if exclude is not None:
        l.exclude.append(exclude)
    return l


def _param(layer, name, method="kUniform", **kw):
    p = layer.param.add()
    p.name = name
    p.init_method = S.enum_value("ParamProto.InitMethod", method)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _ip(net, name, src, nout, low=-0.05, high=0.05):
    l = net.layer.add()
    l.name, l.type = name, "kInnerProduct"
    l.srclayers.append(src)
    l.inner_product_param.num_output = nout
    _param(l, "weight", low=low, high=high)
    _param(l, "bias", low=low, high=high)
    return l


def _act(net, name, src, typ):
    l = net.layer.add()
    l.name, l.type = name, typ
    l.srclayers.append(src)
    return l


def _loss(net, src, label="label", topk=1):
    l = net.layer.add()
    l.name, l.type = "loss", "kSoftmaxLoss"
    l.srclayers.extend([src, label])
    l.softmaxloss_param.topk = topk
    return l


def _conv(net, name, src, nf, k, stride=1, pad=0, std=0.01, bias=0.0, lr_b=2.0, wd_b=0.0):
    l = net.layer.add()
    l.name, l.type = name, "kConvolution"
    l.srclayers.append(src)
    cp = l.convolution_param
    cp.num_filters, cp.kernel, cp.stride, cp.pad = nf, k, stride, pad
    _param(l, "weight", "kGaussain", mean=0.0, std=std)
    _param(l, "bias", "kConstant", value=bias, learning_rate_multiplier=lr_b, weight_decay_multiplier=wd_b)
    return l


def _pool(net, name, src, k, stride, method="MAX"):
    l = net.layer.add()
    l.name, l.type = name, "kPooling"
    l.srclayers.append(src)
    l.pooling_param.kernel, l.pooling_param.stride = k, stride
    l.pooling_param.pool = S.enum_value("PoolingProto.PoolMethod", method)
    return l


def _lrn(net, name, src, size=3, alpha=5e-5, beta=0.75):
    l = net.layer.add()
    l.name, l.type = name, "kLRN"
    l.srclayers.append(src)
    l.lrn_param.local_size, l.lrn_param.alpha, l.lrn_param.beta = size, alpha, beta
    return l


def _sgd(model, lr, momentum=0.0, wd=0.0):
    u = model.updater
    u.type = S.enum_value("UpdaterProto.Type", "kSGD")
    u.base_learning_rate = lr
    u.momentum = momentum
    u.weight_decay = wd
    u.param_type = "Elastic"
    return u


def mnist_mlp(batch=1000, train_steps=60000, hidden=(2500, 2000, 1500, 1000, 500)):
    m = S.ModelProto()
    m.name = "deep-big-simple-mlp"
    m.train_steps = train_steps
    m.display_frequency = 30
    m.prefetch = True
    u = _sgd(m, 0.001)
    u.learning_rate_change_method = S.enum_value("UpdaterProto.ChangeProto", "kStep")
    u.learning_rate_change_frequency = 60
    u.gamma = 0.997
    u.sync_frequency = 8
    u.moving_rate = 0.9
    u.warmup_steps = 60
    net = m.neuralnet
    _data(net, batch, [28, 28], 10)
    l = net.layer.add(); l.name, l.type = "mnist", "kMnistImage"; l.srclayers.append("data")
    l.mnist_param.norm_a, l.mnist_param.norm_b = 127.5, 1.0
    l = net.layer.add(); l.name, l.type = "label", "kLabel"; l.srclayers.append("data")
    src = "mnist"
    for i, h in enumerate(hidden, 1):
        _ip(net, f"fc{i}", src, h)
        _act(net, f"tanh{i}", f"fc{i}", "kTanh")
        src = f"tanh{i}"
    _ip(net, f"fc{len(hidden) + 1}", src, 10)
    _loss(net, f"fc{len(hidden) + 1}")
    return m


def wide_mlp(batch=4096, hidden=16384, nlayers=4, classes=1000, train_steps=1000):
    """Server-sharding stress: `nlayers` hidden x hidden inner products (push/pull bandwidth bound)."""
    m = S.ModelProto()
    m.name = "wide-mlp-16k"
    m.train_steps = train_steps
    m.display_frequency = 50
    _sgd(m, 0.001, momentum=0.9)
    net = m.neuralnet
    _data(net, batch, [128, 128], classes)
    l = net.layer.add(); l.name, l.type = "mnist", "kMnistImage"; l.srclayers.append("data")
    l.mnist_param.norm_a, l.mnist_param.norm_b = 127.5, 1.0
    l = net.layer.add(); l.name, l.type = "label", "kLabel"; l.srclayers.append("data")
    s = 0.5 / (hidden ** 0.5)
    src = "mnist"
    for i in range(1, nlayers + 1):
        _ip(net, f"fc{i}", src, hidden, low=-s, high=s)
        _act(net, f"relu{i}", f"fc{i}", "kReLU")
        src = f"relu{i}"
    _ip(net, "fc_out", src, classes, low=-s, high=s)
    _loss(net, "fc_out")
    return m


def lenet(batch=64, train_steps=10000):
    """examples/mnist/conv.conf: conv 20x5x5 - pool 2/2 - conv 50x5x5 - pool 2/2 - ip 500 - relu - ip 10."""
    m = S.ModelProto()
    m.name = "lenet"
    m.train_steps = train_steps
    m.display_frequency = 100
    u = _sgd(m, 0.01, momentum=0.9, wd=0.0005)
    u.learning_rate_change_method = S.enum_value("UpdaterProto.ChangeProto", "kInverse")
    u.gamma, u.pow = 0.0001, 0.75
    net = m.neuralnet
    _data(net, batch, [28, 28], 10)
    l = net.layer.add(); l.name, l.type = "mnist", "kMnistImage"; l.srclayers.append("data")
    l.mnist_param.norm_a = 255.0
    l = net.layer.add(); l.name, l.type = "label", "kLabel"; l.srclayers.append("data")
    for name, src, nf in (("conv1", "mnist", 20), ("conv2", "pool1", 50)):
        c = _conv(net, name, src, nf, 5)
        c.param[0].init_method = S.enum_value("ParamProto.InitMethod", "kUniformSqrtFanIn")
        c.param[0].low, c.param[0].high = -1.0, 1.0
        _pool(net, "pool1" if name == "conv1" else "pool2", name, 2, 2)
    ip1 = _ip(net, "ip1", "pool2", 500)
    ip1.param[0].init_method = S.enum_value("ParamProto.InitMethod", "kUniformSqrtFanIn")
    ip1.param[0].low, ip1.param[0].high = -1.0, 1.0
    _act(net, "relu1", "ip1", "kReLU")
    ip2 = _ip(net, "ip2", "relu1", 10)
    ip2.param[0].init_method = S.enum_value("ParamProto.InitMethod", "kUniformSqrtFanIn")
    ip2.param[0].low, ip2.param[0].high = -1.0, 1.0
    _loss(net, "ip2")
    return m


def cifar10_convnet(batch=128, train_steps=70000):
    """(conv 5x5 - pool 3/2 - LRN) x3 + inner product + softmax — the "CIFAR-10 quick" family."""
    m = S.ModelProto()
    m.name = "cifar10-convnet"
    m.train_steps = train_steps
    m.display_frequency = 100
    _sgd(m, 0.001, momentum=0.9, wd=0.004)
    net = m.neuralnet
    _data(net, batch, [3, 32, 32], 10)
    l = net.layer.add(); l.name, l.type = "rgb", "kRGBImage"; l.srclayers.append("data")
    l.rgbimage_param.scale = 1.0 / 128
    l = net.layer.add(); l.name, l.type = "label", "kLabel"; l.srclayers.append("data")
    _conv(net, "conv1", "rgb", 32, 5, 1, 2, std=1e-4)
    _pool(net, "pool1", "conv1", 3, 2, "MAX")
    _act(net, "relu1", "pool1", "kReLU")
    _lrn(net, "norm1", "relu1")
    _conv(net, "conv2", "norm1", 32, 5, 1, 2, std=0.01)
    _act(net, "relu2", "conv2", "kReLU")
    _pool(net, "pool2", "relu2", 3, 2, "AVE")
    _lrn(net, "norm2", "pool2")
    _conv(net, "conv3", "norm2", 64, 5, 1, 2, std=0.01)
    _act(net, "relu3", "conv3", "kReLU")
    _pool(net, "pool3", "relu3", 3, 2, "AVE")
    _lrn(net, "norm3", "pool3")
    ip = _ip(net, "ip1", "norm3", 10)
    ip.param[0].init_method = S.enum_value("ParamProto.InitMethod", "kGaussain")
    ip.param[0].std = 0.01
    ip.param[1].init_method = S.enum_value("ParamProto.InitMethod", "kConstant")
    ip.param[1].value = 0.0
    _loss(net, "ip1")
    return m


def alexnet(batch=256, train_steps=450000, classes=1000):
    """AlexNet, ImageNet shape 3x256x256 cropped to 227, ungrouped convolutions (the current
    ConvolutionProto has no group field; layer sizes as in the legacy src/test/data/model.conf:71-118)."""
    m = S.ModelProto()
    m.name = "alexnet"
    m.train_steps = train_steps
    m.display_frequency = 20
    u = _sgd(m, 0.01, momentum=0.9, wd=0.0005)
    u.learning_rate_change_method = S.enum_value("UpdaterProto.ChangeProto", "kStep")
    u.learning_rate_change_frequency = 100000
    u.gamma = 0.1
    net = m.neuralnet
    _data(net, batch, [3, 256, 256], classes)
    l = net.layer.add(); l.name, l.type = "rgb", "kRGBImage"; l.srclayers.append("data")
    l.rgbimage_param.cropsize, l.rgbimage_param.mirror, l.rgbimage_param.scale = 227, True, 1.0 / 128
    l = net.layer.add(); l.name, l.type = "label", "kLabel"; l.srclayers.append("data")
    _conv(net, "conv1", "rgb", 96, 11, 4, 0, std=0.01)
    _act(net, "relu1", "conv1", "kReLU")
    _pool(net, "pool1", "relu1", 3, 2)
    _lrn(net, "norm1", "pool1", 5, 1e-4, 0.75)
    _conv(net, "conv2", "norm1", 256, 5, 1, 2, std=0.01, bias=1.0)
    _act(net, "relu2", "conv2", "kReLU")
    _pool(net, "pool2", "relu2", 3, 2)
    _lrn(net, "norm2", "pool2", 5, 1e-4, 0.75)
    _conv(net, "conv3", "norm2", 384, 3, 1, 1, std=0.01)
    _act(net, "relu3", "conv3", "kReLU")
    _conv(net, "conv4", "relu3", 384, 3, 1, 1, std=0.01, bias=1.0)
    _act(net, "relu4", "conv4", "kReLU")
    _conv(net, "conv5", "relu4", 256, 3, 1, 1, std=0.01, bias=1.0)
    _act(net, "relu5", "conv5", "kReLU")
    _pool(net, "pool5", "relu5", 3, 2)
    src = "pool5"
    for i, n in ((6, 4096), (7, 4096)):
        ip = _ip(net, f"fc{i}", src, n)
        ip.param[0].init_method = S.enum_value("ParamProto.InitMethod", "kGaussain")
        ip.param[0].std = 0.005
        ip.param[1].init_method = S.enum_value("ParamProto.InitMethod", "kConstant")
        ip.param[1].value = 1.0
        _act(net, f"relu{i}", f"fc{i}", "kReLU")
        d = net.layer.add(); d.name, d.type = f"drop{i}", "kDropout"; d.srclayers.append(f"relu{i}")
        d.dropout_param.dropout_ratio = 0.5
        src = f"drop{i}"
    ip = _ip(net, "fc8", src, classes)
    ip.param[0].init_method = S.enum_value("ParamProto.InitMethod", "kGaussain")
    ip.param[0].std = 0.01
    ip.param[1].init_method = S.enum_value("ParamProto.InitMethod", "kConstant")
    ip.param[1].value = 0.0
    _loss(net, "fc8")
    return m


CONFIGS = {
    "mnist_mlp": mnist_mlp,
    "wide_mlp": wide_mlp,
    "lenet": lenet,
    "cifar10": cifar10_convnet,
    "alexnet": alexnet,
}
