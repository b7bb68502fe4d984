"""Layer factory: type id -> layer object (reference ``src/layer/layer_impl-inl.hpp:36-76``).

``softplus`` and ``maxout`` parse to an id but have no implementation in the reference
(``layer_impl-inl.hpp:44-75``); they are rejected here with the same message.  The ``caffe``
plugin layer (reference ``src/plugin/caffe_adapter-inl.hpp``) is out of scope.
"""
from __future__ import annotations

from .. import netconfig as nc
from . import native, zoo
from .base import Layer
from .pairtest import PairTestLayer

_NATIVE = {
    nc.kFullConnect: native.FullConnectLayer,
    nc.kConv: native.ConvolutionLayer,
    nc.kRectifiedLinear: native.ReluLayer,
    nc.kSigmoid: native.SigmoidLayer,
    nc.kTanh: native.TanhLayer,
    nc.kXelu: native.XeluLayer,
    nc.kMaxPooling: native.MaxPoolingLayer,
    nc.kSumPooling: native.SumPoolingLayer,
    nc.kAvgPooling: native.AvgPoolingLayer,
    nc.kReluMaxPooling: native.ReluMaxPoolingLayer,
    nc.kLRN: native.LRNLayer,
    nc.kFlatten: native.FlattenLayer,
    nc.kDropout: native.DropoutLayer,
    nc.kSoftmax: native.SoftmaxLayer,
    nc.kL2Loss: native.L2LossLayer,
    nc.kMultiLogistic: native.MultiLogisticLayer,
    nc.kBias: native.BiasLayer,
    nc.kConcat: native.ConcatLayer,
    nc.kChConcat: native.ChConcatLayer,
    nc.kSplit: native.SplitLayer,
    nc.kPRelu: native.PReluLayer,
    nc.kBatchNorm: native.BatchNormLayer,
    nc.kInsanity: native.InsanityLayer,
    nc.kFixConnect: native.FixConnectLayer,
    nc.kInsanityPooling: native.InsanityPoolingLayer,
}

_REF_ONLY = {}


def create_layer(type_id: int) -> Layer:
    if type_id >= nc.kPairTestGap:
        return PairTestLayer(create_layer(type_id // nc.kPairTestGap), create_layer(type_id % nc.kPairTestGap))
    if type_id in _NATIVE:
        return _NATIVE[type_id]()
    if type_id in _REF_ONLY:
        return _REF_ONLY[type_id]()
    if type_id == nc.kCaffe:
        raise ValueError("caffe plugin layer is not supported by cxxnet_b200")
    raise ValueError("unknown layer type id : \"%d\"" % type_id)
