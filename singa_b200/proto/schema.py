"""Job-configuration schemas (``package singa``) materialised at run time — there is no ``protoc`` on the box.

The message/field names, numbers, labels and defaults are the reference's job API
(src/proto/model.proto, cluster.proto, topology.proto) so its text-format ``.conf`` files parse unchanged.
Additions (all on unused field numbers, so reference files stay valid):

* ``ParamProto.blob`` (6, the slot of the reference's commented-out value field, model.proto:69-70),
  ``ParamProto.aux_blob`` (15) and message ``ParamCheckpoint`` — the Param-proto checkpoint layout.
* ``ConvolutionProto.group`` is *not* added; AlexNet is authored ungrouped.
* ``ClusterProto.server_update`` / ``staleness`` / ``nserver_groups`` knobs for the B200 parameter server.

The schema is written as a compact table and turned into a ``FileDescriptorProto``.
"""
from __future__ import annotations

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_F = descriptor_pb2.FieldDescriptorProto
_TYPES = {
    "int32": _F.TYPE_INT32, "uint32": _F.TYPE_UINT32, "int64": _F.TYPE_INT64, "float": _F.TYPE_FLOAT,
    "double": _F.TYPE_DOUBLE, "bool": _F.TYPE_BOOL, "string": _F.TYPE_STRING, "bytes": _F.TYPE_BYTES,
}
_LABELS = {"opt": _F.LABEL_OPTIONAL, "req": _F.LABEL_REQUIRED, "rep": _F.LABEL_REPEATED}

# name -> list of values (number assigned explicitly)
ENUMS = {
    "Phase": [("kTrain", 0), ("kValidation", 1), ("kTest", 2)],
    "PartitionType": [("kDataPartition", 0), ("kLayerPartition", 1), ("kNone", 2)],
    "ConnectionType": [("kOneToOne", 0), ("kOneToAll", 1)],
}

# message -> (nested enums, fields).  field = (label, type, name, number[, default[, packed]])
MESSAGES = {
    "ModelProto": ({"GradCalcAlg": [("kBackPropagation", 1), ("kContrastiveDivergence", 2)]}, [
        ("opt", "string", "name", 1),
        ("opt", "string", "train_folder", 2, "train"),
        ("opt", "string", "test_folder", 3, "test"),
        ("opt", "string", "validation_folder", 4, "validation"),
        ("opt", "int32", "display_after_steps", 6, 0),
        ("opt", "int32", "display_frequency", 7, 0),
        ("opt", "int32", "validation_after_steps", 10, 0),
        ("opt", "int32", "validation_frequency", 11, 0),
        ("opt", "int32", "test_after_steps", 13, 0),
        ("opt", "int32", "test_frequency", 14, 0),
        ("opt", "bool", "prefetch", 15, True),
        ("opt", "int32", "train_steps", 20),
        ("opt", "int32", "validation_steps", 21),
        ("opt", "int32", "test_steps", 22),
        ("opt", "int32", "step", 29, 0),
        ("opt", "UpdaterProto", "updater", 31),
        ("opt", "ModelProto.GradCalcAlg", "alg", 32, "kBackPropagation"),
        ("opt", "NetProto", "neuralnet", 40),
        ("opt", "bool", "debug", 41, False),
        # additions: checkpointing schedule (the reference only declares `step`, model.proto:34-35)
        ("opt", "int32", "checkpoint_after_steps", 50, 0),
        ("opt", "int32", "checkpoint_frequency", 51, 0),
        ("opt", "string", "checkpoint", 52),
    ]),
    "NetProto": ({}, [
        ("rep", "LayerProto", "layer", 1),
        ("opt", "PartitionType", "partition_type", 3, "kNone"),
    ]),
    "ParamProto": ({"InitMethod": [("kConstant", 0), ("kGaussain", 1), ("kUniform", 2), ("kPretrained", 3),
                                   ("kGaussainSqrtFanIn", 4), ("kUniformSqrtFanIn", 5),
                                   ("kUniformSqrtFanInOut", 6)]}, [
        ("opt", "string", "name", 1),
        ("opt", "int32", "id", 2),
        ("rep", "int32", "shape", 3),
        ("opt", "int32", "split_threshold", 4, 5000000),
        ("opt", "int32", "partition_dim", 5, -1),
        ("rep", "BlobProto", "blob", 6),
        ("opt", "ParamProto.InitMethod", "init_method", 7, "kConstant"),
        ("opt", "float", "value", 8, 1.0),
        ("opt", "float", "low", 9, -1.0),
        ("opt", "float", "high", 10, 1.0),
        ("opt", "float", "mean", 11, 0.0),
        ("opt", "float", "std", 12, 1.0),
        ("opt", "float", "learning_rate_multiplier", 13, 1.0),
        ("opt", "float", "weight_decay_multiplier", 14, 1.0),
        ("rep", "BlobProto", "aux_blob", 15),
        ("opt", "int32", "version", 16, 0),
    ]),
    "LayerProto": ({}, [
        ("opt", "string", "name", 1),
        ("opt", "string", "type", 2),
        ("rep", "string", "srclayers", 3),
        ("opt", "int32", "locationid", 4, 0),
        ("opt", "int32", "partitionid", 5, 0),
        ("opt", "PartitionType", "partition_type", 6),
        ("rep", "string", "share_ary", 11),
        ("rep", "ParamProto", "param", 12),
        ("rep", "string", "share_param", 13),
        ("rep", "Phase", "exclude", 20),
        ("opt", "ConvolutionProto", "convolution_param", 21),
        ("opt", "ConcateProto", "concate_param", 31),
        ("opt", "DataProto", "data_param", 22),
        ("opt", "DropoutProto", "dropout_param", 23),
        ("opt", "InnerProductProto", "inner_product_param", 24),
        ("opt", "LRNProto", "lrn_param", 25),
        ("opt", "MnistProto", "mnist_param", 26),
        ("opt", "PoolingProto", "pooling_param", 27),
        ("opt", "SliceProto", "slice_param", 32),
        ("opt", "SplitProto", "split_param", 33),
        ("opt", "ReLUProto", "relu_param", 28),
        ("opt", "RGBImage", "rgbimage_param", 34),
        ("opt", "SoftmaxLossProto", "softmaxloss_param", 29),
        ("opt", "TanhProto", "tanh_param", 30),
    ]),
    "RGBImage": ({}, [
        ("opt", "float", "scale", 1, 1.0),
        ("opt", "int32", "cropsize", 2, 0),
        ("opt", "bool", "mirror", 3, False),
    ]),
    "SplitProto": ({}, [("opt", "int32", "num_splits", 1)]),
    "TanhProto": ({}, [
        ("opt", "float", "outer_scale", 1, 1.0),
        ("opt", "float", "inner_scale", 2, 1.0),
    ]),
    "SoftmaxLossProto": ({}, [
        ("opt", "int32", "topk", 1, 1),
        ("opt", "float", "scale", 2, 1.0),
    ]),
    "ConvolutionProto": ({}, [
        ("opt", "uint32", "num_filters", 1),
        ("opt", "bool", "bias_term", 2, True),
        ("opt", "uint32", "pad", 3, 0),
        ("opt", "uint32", "stride", 4, 1),
        ("req", "uint32", "kernel", 5),
    ]),
    "ConcateProto": ({}, [
        ("opt", "int32", "concate_dimension", 1),
        ("opt", "int32", "concate_num", 2),
    ]),
    "DataProto": ({}, [
        ("opt", "string", "source", 1),
        ("opt", "string", "path", 2),
        ("opt", "uint32", "batchsize", 4),
        ("opt", "uint32", "random_skip", 5, 0),
        # additions: shape/classes of the synthetic source ("source: synthetic")
        ("rep", "int32", "synthetic_shape", 10),
        ("opt", "int32", "synthetic_classes", 11, 10),
    ]),
    "MnistProto": ({}, [
        ("opt", "int32", "kernel", 1, 0),
        ("opt", "float", "sigma", 2, 0.0),
        ("opt", "float", "alpha", 3, 0.0),
        ("opt", "float", "beta", 4, 0.0),
        ("opt", "float", "gamma", 5, 0.0),
        ("opt", "int32", "resize", 6, 0),
        ("opt", "int32", "elastic_freq", 7, 0),
        ("opt", "float", "norm_a", 8, 1.0),
        ("opt", "float", "norm_b", 9, 0.0),
    ]),
    "DropoutProto": ({}, [("opt", "float", "dropout_ratio", 1, 0.5)]),
    "InnerProductProto": ({}, [
        ("opt", "uint32", "num_output", 1),
        ("opt", "bool", "bias_term", 2, True),
    ]),
    "LRNProto": ({"NormRegion": [("ACROSS_CHANNELS", 0), ("WITHIN_CHANNEL", 1)]}, [
        ("opt", "uint32", "local_size", 1, 5),
        ("opt", "float", "alpha", 2, 1.0),
        ("opt", "float", "beta", 3, 0.75),
        ("opt", "LRNProto.NormRegion", "norm_region", 4, "ACROSS_CHANNELS"),
        ("opt", "float", "knorm", 5, 1.0),
    ]),
    "PoolingProto": ({"PoolMethod": [("MAX", 0), ("AVE", 1)]}, [
        ("opt", "PoolingProto.PoolMethod", "pool", 1, "MAX"),
        ("req", "uint32", "kernel", 2),
        ("opt", "uint32", "pad", 4, 0),
        ("opt", "uint32", "stride", 3, 1),
    ]),
    "SliceProto": ({}, [
        ("opt", "int32", "slice_dimension", 1),
        ("opt", "int32", "slice_num", 2),
    ]),
    "ReLUProto": ({}, [("opt", "float", "negative_slope", 1, 0.0)]),
    "Record": ({"Type": [("kSingleLabelImage", 0)]}, [
        ("opt", "Record.Type", "type", 1, "kSingleLabelImage"),
        ("opt", "SingleLabelImageRecord", "image", 2),
    ]),
    "Datum": ({}, [
        ("opt", "int32", "channels", 1),
        ("opt", "int32", "height", 2),
        ("opt", "int32", "width", 3),
        ("opt", "bytes", "data", 4),
        ("opt", "int32", "label", 5),
        ("rep", "float", "float_data", 6),
        ("opt", "bool", "encoded", 7, False),
    ]),
    "SingleLabelImageRecord": ({}, [
        ("rep", "int32", "shape", 1),
        ("opt", "int32", "label", 2),
        ("opt", "bytes", "pixel", 3),
        ("rep", "float", "data", 4),
    ]),
    "UpdaterProto": ({"Type": [("kAdaGrad", 1), ("kAdaDelta", 2), ("kNesterov", 3), ("kSGD", 4), ("kRMSProp", 5)],
                      "ChangeProto": [("kFixed", 0), ("kInverse_t", 1), ("kInverse", 2), ("kExponential", 3),
                                      ("kLinear", 4), ("kStep", 5)]}, [
        ("opt", "UpdaterProto.Type", "type", 1, "kAdaGrad"),
        ("opt", "bool", "hogwild", 2, True),
        ("opt", "float", "momentum", 4, 0.0),
        ("opt", "float", "weight_decay", 5, 0.0),
        ("opt", "float", "gamma", 6, 1.0),
        ("opt", "float", "pow", 7, 0.0),
        ("opt", "float", "delta", 8, 0.0000001),
        ("opt", "float", "rho", 9, 0.9),
        ("opt", "float", "base_learning_rate", 12),
        ("opt", "float", "final_learning_rate", 13),
        ("opt", "int32", "learning_rate_change_frequency", 14),
        ("opt", "UpdaterProto.ChangeProto", "learning_rate_change_method", 16, "kFixed"),
        ("opt", "int32", "sync_frequency", 17, 1),
        ("opt", "int32", "warmup_steps", 25, 10),
        ("opt", "float", "moving_rate", 26, 0.0),
        ("opt", "string", "param_type", 27, "Elastic"),
    ]),
    "BlobProto": ({}, [
        ("opt", "int32", "num", 1, 0),
        ("opt", "int32", "channels", 2, 0),
        ("opt", "int32", "height", 3, 0),
        ("opt", "int32", "width", 4, 0),
        ("rep", "float", "data", 5, None, True),
        ("rep", "float", "diff", 6, None, True),
        ("opt", "int64", "offset", 7, 0),      # addition: first element of this segment within the flattened parameter
    ]),
    # ---- checkpoint container (addition) -----------------------------------------------------
    "ParamCheckpoint": ({}, [
        ("opt", "int32", "step", 1, 0),
        ("rep", "ParamProto", "param", 2),
        ("opt", "int32", "shard_id", 3, 0),
        ("opt", "int32", "nshards", 4, 1),
        ("opt", "string", "model_name", 5),
    ]),
    # ---- cluster.proto -----------------------------------------------------------------------
    "ClusterProto": ({}, [
        ("opt", "int32", "nworkers", 1),
        ("opt", "int32", "nservers", 2),
        ("opt", "int32", "start_port", 3, 6723),
        ("opt", "int32", "nprocs_per_group", 5, 1),
        ("opt", "int32", "nthreads_per_procs", 6, 1),
        ("opt", "int32", "nthreads_per_server", 7, 1),
        ("req", "string", "workspace", 10),
        ("opt", "string", "vis_subfolder", 11, "vis"),
        ("opt", "string", "log_subfolder", 12, "log"),
        ("opt", "bool", "synchronous", 15, False),
        ("opt", "int32", "largest_message", 20, 1048576),
        ("opt", "float", "bandwidth", 21, 100.0),
        # additions for the single-box B200 parameter server
        ("opt", "int32", "staleness", 30, 0),           # max steps a worker group may run ahead of a shard
        ("opt", "string", "push_pull", 31, "fused"),    # "fused" (peer-memory kernels) | "nccl" (baseline)
        ("opt", "string", "checkpoint_subfolder", 32, "checkpoint"),
        # server groups on one box (Topology.nserver_groups / ServerGroup.sync_interval of the reference's Stack B):
        # the ranks are cut into this many groups, each holding its own sharded copy of the parameters; a shard owner
        # averages with the same shard of the other groups every `server_sync_interval` worker exchanges
        ("opt", "int32", "nserver_groups", 33, 1),
        ("opt", "int32", "server_sync_interval", 34, 1),
    ]),
    # ---- topology.proto ----------------------------------------------------------------------
    "Topology": ({}, [
        ("req", "int32", "nservers", 2),
        ("req", "int32", "nworker_groups", 3),
        ("rep", "int32", "nserver_groups", 4),
        ("rep", "ServerGroup", "server_group", 5),
        ("req", "int32", "port", 6),
        ("req", "int32", "server_threads", 7),
        ("req", "int32", "worker_threads", 8),
    ]),
    "ServerGroup": ({}, [
        ("req", "int32", "id", 9),
        ("req", "int32", "sync_interval", 10),
        ("rep", "int32", "neighbor", 11),
    ]),
}


def _default_str(v) -> str:
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, float):
        return repr(v)
    return str(v)


def _build_file() -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "singa_b200/job.proto"
    fd.package = "singa"
    fd.syntax = "proto2"
    for ename, values in ENUMS.items():
        e = fd.enum_type.add()
        e.name = ename
        for vn, vi in values:
            ev = e.value.add()
            ev.name, ev.number = vn, vi
    enum_names = set(ENUMS)
    for mname, (nested, _fields) in MESSAGES.items():
        for en in nested:
            enum_names.add(f"{mname}.{en}")
    for mname, (nested, fields) in MESSAGES.items():
        m = fd.message_type.add()
        m.name = mname
        for en, values in nested.items():
            e = m.enum_type.add()
            e.name = en
            for vn, vi in values:
                ev = e.value.add()
                ev.name, ev.number = vn, vi
        for spec in fields:
            label, typ, name, number = spec[:4]
            default = spec[4] if len(spec) > 4 else None
            packed = spec[5] if len(spec) > 5 else False
            f = m.field.add()
            f.name, f.number, f.label = name, number, _LABELS[label]
            if typ in _TYPES:
                f.type = _TYPES[typ]
            elif typ in enum_names:
                f.type = _F.TYPE_ENUM
                f.type_name = f".singa.{typ}"
            else:
                assert typ in MESSAGES, typ
                f.type = _F.TYPE_MESSAGE
                f.type_name = f".singa.{typ}"
            if default is not None:
                f.default_value = _default_str(default)
            if packed:
                f.options.packed = True
    return fd


_pool = descriptor_pool.DescriptorPool()
_file = _pool.Add(_build_file())


def message_class(name: str):
    return message_factory.GetMessageClass(_pool.FindMessageTypeByName(f"singa.{name}"))


def enum_value(enum_path: str, name: str) -> int:
    return _pool.FindEnumTypeByName(f"singa.{enum_path}").values_by_name[name].number


def enum_name(enum_path: str, number: int) -> str:
    return _pool.FindEnumTypeByName(f"singa.{enum_path}").values_by_number[number].name


ModelProto = message_class("ModelProto")
NetProto = message_class("NetProto")
LayerProto = message_class("LayerProto")
ParamProto = message_class("ParamProto")
UpdaterProto = message_class("UpdaterProto")
BlobProto = message_class("BlobProto")
Record = message_class("Record")
Datum = message_class("Datum")
SingleLabelImageRecord = message_class("SingleLabelImageRecord")
ClusterProto = message_class("ClusterProto")
Topology = message_class("Topology")
ServerGroup = message_class("ServerGroup")
ParamCheckpoint = message_class("ParamCheckpoint")

# enum constants used throughout the framework
kTrain, kValidation, kTest = 0, 1, 2
kDataPartition, kLayerPartition, kNone = 0, 1, 2
kOneToOne, kOneToAll = 0, 1
