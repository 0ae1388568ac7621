"""The layer zoo (18 registered types, src/worker/neuralnet.cc:13-33)."""
from .base import (BridgeDstLayer, BridgeSrcLayer, ConcateLayer, Layer, SliceLayer, SplitLayer,
                   layer_factory)
from .data import (DataLayer, LabelLayer, LMDBDataLayer, MnistImageLayer, ParserLayer, RGBImageLayer,
                   ShardDataLayer)
from .neuron import (ConvolutionLayer, DropoutLayer, InnerProductLayer, LossLayer, LRNLayer, PoolingLayer,
                     ReLULayer, SoftmaxLossLayer, TanhLayer)

ALL_LAYERS = [
    BridgeDstLayer, BridgeSrcLayer, ConvolutionLayer, ConcateLayer, DropoutLayer, InnerProductLayer,
    LabelLayer, LMDBDataLayer, LRNLayer, MnistImageLayer, PoolingLayer, ReLULayer, RGBImageLayer,
    ShardDataLayer, SliceLayer, SoftmaxLossLayer, SplitLayer, TanhLayer,
]
