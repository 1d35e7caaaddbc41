"""Layer-builder DSL: host-side mirror of src/main/scala/libs/Layers.scala (RDDLayer :17-41,
ConvolutionLayer :43-58, PoolingLayer :66-87, InnerProductLayer :89-102, ReLULayer :104-116,
SoftmaxWithLoss :118-134, NetParam :136-145), widened with the optional arguments the four
benchmark nets need (pad/stride/group/fillers/param specs, LRN, Dropout, Accuracy)."""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

from .caffe_proto import ENUMS, Msg

TRAIN, TEST = ENUMS["Phase"]["TRAIN"], ENUMS["Phase"]["TEST"]


def _base(type_: str, name: str, bottom: Sequence[str], top: Sequence[str]) -> Msg:
    l = Msg("LayerParameter")
    l.name = name
    l.type = type_
    if bottom:
        l.bottom = list(bottom)
    if top:
        l.top = list(top)
    return l


def _include(l: Msg, phase: Optional[int]) -> Msg:
    if phase is not None:
        l.include = [Msg("NetStateRule", phase=phase)]
    return l


def _params(l: Msg, specs: Optional[Sequence[Tuple[Optional[float], Optional[float]]]]) -> None:
    if specs:
        out = []
        for lr, decay in specs:
            p = Msg("ParamSpec")
            if lr is not None:
                p.lr_mult = lr
            if decay is not None:
                p.decay_mult = decay
            out.append(p)
        l.param = out


def _filler(type_: str, std: Optional[float] = None, value: Optional[float] = None) -> Msg:
    f = Msg("FillerParameter", type=type_)
    if value is not None:
        f.value = value
    if std is not None:
        f.std = std
    return f


def RDDLayer(name: str, shape: Sequence[int], include: Optional[int]) -> Msg:
    l = _base("JavaData", name, [], [name])
    l.java_data_param = Msg("JavaDataParameter", shape=Msg("BlobShape", dim=list(shape)))
    return _include(l, include)


def ConvolutionLayer(name: str, bottom: List[str], kernel: int, num_output: int, *, pad: Optional[int] = None,
                     stride: Optional[int] = None, group: Optional[int] = None, weight_filler: Optional[Msg] = None,
                     bias_filler: Optional[Msg] = None, params=None) -> Msg:
    l = _base("Convolution", name, bottom, [name])
    _params(l, params)
    c = Msg("ConvolutionParameter", num_output=num_output)
    if pad is not None:
        c.pad = [pad]
    c.kernel_size = [kernel]
    if group is not None:
        c.group = group
    if stride is not None:
        c.stride = [stride]
    if weight_filler is not None:
        c.weight_filler = weight_filler
    if bias_filler is not None:
        c.bias_filler = bias_filler
    l.convolution_param = c
    return l


def PoolingLayer(name: str, bottom: List[str], pooling: str, kernel: int, stride: int, top: Optional[str] = None) -> Msg:
    l = _base("Pooling", name, bottom, [top or name])
    l.pooling_param = Msg("PoolingParameter", pool=ENUMS["PoolMethod"][pooling], kernel_size=kernel, stride=stride)
    return l


def InnerProductLayer(name: str, bottom: List[str], num_output: int, *, weight_filler: Optional[Msg] = None,
                      bias_filler: Optional[Msg] = None, params=None) -> Msg:
    l = _base("InnerProduct", name, bottom, [name])
    _params(l, params)
    p = Msg("InnerProductParameter", num_output=num_output)
    if weight_filler is not None:
        p.weight_filler = weight_filler
    if bias_filler is not None:
        p.bias_filler = bias_filler
    l.inner_product_param = p
    return l


def ReLULayer(name: str, bottom: List[str]) -> Msg:
    return _base("ReLU", name, bottom, list(bottom))           # in place, as in all four nets


def DropoutLayer(name: str, bottom: List[str], ratio: float) -> Msg:
    l = _base("Dropout", name, bottom, list(bottom))
    l.dropout_param = Msg("DropoutParameter", dropout_ratio=ratio)
    return l


def LRNLayer(name: str, bottom: List[str], local_size: int, alpha: float, beta: float,
             norm_region: Optional[str] = None) -> Msg:
    l = _base("LRN", name, bottom, [name])
    p = Msg("LRNParameter", local_size=local_size, alpha=alpha, beta=beta)
    if norm_region is not None:
        p.norm_region = ENUMS["NormRegion"][norm_region]
    l.lrn_param = p
    return l


def AccuracyLayer(name: str, bottom: List[str], include: Optional[int] = TEST) -> Msg:
    return _include(_base("Accuracy", name, bottom, [name]), include)


def SoftmaxWithLoss(name: str, bottom: List[str]) -> Msg:
    return _base("SoftmaxWithLoss", name, bottom, [name])


def NetParam(name: str, *layers: Msg) -> Msg:
    return Msg("NetParameter", name=name, layer=list(layers))


gaussian = lambda std: _filler("gaussian", std=std)
constant = lambda value=None: _filler("constant", value=value)
xavier = lambda: _filler("xavier")
