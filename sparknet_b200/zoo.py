"""The four benchmark nets of BASELINE.json, built with the layer DSL so that bench.py and the
GPU tests need no file from the reference tree.  Each builder reproduces, byte for byte after
serialization, what SparkNet sends to load_solver_from_protobuf for the reference's prototxts
(caffe/examples/mnist/lenet_{train_test,solver}.prototxt, caffe/examples/cifar10/
cifar10_{quick,full}_{train_test,solver}.prototxt, caffe/models/bvlc_reference_caffenet/
{train_val,solver}.prototxt) after ProtoLoader.replaceDataLayers +
loadSolverPrototxtWithNet; tests/test_proto.py pins that against tests/golden/*.binpb."""
from __future__ import annotations

from typing import Callable, Dict, List, Tuple

from .caffe_proto import ENUMS, Msg
from .layers import (TEST, TRAIN, AccuracyLayer, ConvolutionLayer, DropoutLayer, InnerProductLayer, LRNLayer,
                     NetParam, PoolingLayer, RDDLayer, ReLULayer, SoftmaxWithLoss, constant, gaussian, xavier)

W12 = [(1.0, None), (2.0, None)]            # param { lr_mult: 1 } param { lr_mult: 2 }
W12D = [(1.0, 1.0), (2.0, 0.0)]             # + decay_mult 1 / 0


def _data(train_bs: int, test_bs: int, c: int, h: int, w: int) -> List[Msg]:
    # order produced by ProtoLoader.replaceDataLayers (ProtoLoader.scala:50-57): TEST pair first
    return [RDDLayer("data", [test_bs, c, h, w], TEST), RDDLayer("label", [test_bs, 1], TEST),
            RDDLayer("data", [train_bs, c, h, w], TRAIN), RDDLayer("label", [train_bs, 1], TRAIN)]


def lenet_net(train_bs: int = 64, test_bs: int = 100) -> Msg:
    conv = lambda n, b, k, o: ConvolutionLayer(n, [b], k, o, stride=1, weight_filler=xavier(),
                                               bias_filler=constant(), params=W12)
    ip = lambda n, b, o: InnerProductLayer(n, [b], o, weight_filler=xavier(), bias_filler=constant(), params=W12)
    return NetParam("LeNet", *_data(train_bs, test_bs, 1, 28, 28),
                    conv("conv1", "data", 5, 20), PoolingLayer("pool1", ["conv1"], "MAX", 2, 2),
                    conv("conv2", "pool1", 5, 50), PoolingLayer("pool2", ["conv2"], "MAX", 2, 2),
                    ip("ip1", "pool2", 500), ReLULayer("relu1", ["ip1"]), ip("ip2", "ip1", 10),
                    AccuracyLayer("accuracy", ["ip2", "label"]), SoftmaxWithLoss("loss", ["ip2", "label"]))


def _cifar_conv(n: str, b: str, o: int, std: float, params=W12) -> Msg:
    return ConvolutionLayer(n, [b], 5, o, pad=2, stride=1, weight_filler=gaussian(std), bias_filler=constant(),
                            params=params)


def cifar10_quick_net(train_bs: int = 100, test_bs: int = 100) -> Msg:
    ip = lambda n, b, o: InnerProductLayer(n, [b], o, weight_filler=gaussian(0.1), bias_filler=constant(), params=W12)
    return NetParam("CIFAR10_quick", *_data(train_bs, test_bs, 3, 32, 32),
                    _cifar_conv("conv1", "data", 32, 0.0001), PoolingLayer("pool1", ["conv1"], "MAX", 3, 2),
                    ReLULayer("relu1", ["pool1"]),
                    _cifar_conv("conv2", "pool1", 32, 0.01), ReLULayer("relu2", ["conv2"]),
                    PoolingLayer("pool2", ["conv2"], "AVE", 3, 2),
                    _cifar_conv("conv3", "pool2", 64, 0.01), ReLULayer("relu3", ["conv3"]),
                    PoolingLayer("pool3", ["conv3"], "AVE", 3, 2),
                    ip("ip1", "pool3", 64), ip("ip2", "ip1", 10),
                    AccuracyLayer("accuracy", ["ip2", "label"]), SoftmaxWithLoss("loss", ["ip2", "label"]))


def cifar10_full_net(train_bs: int = 100, test_bs: int = 100) -> Msg:
    lrn = lambda n, b: LRNLayer(n, [b], 3, 5e-05, 0.75, "WITHIN_CHANNEL")
    return NetParam("CIFAR10_full", *_data(train_bs, test_bs, 3, 32, 32),
                    _cifar_conv("conv1", "data", 32, 0.0001), PoolingLayer("pool1", ["conv1"], "MAX", 3, 2),
                    ReLULayer("relu1", ["pool1"]), lrn("norm1", "pool1"),
                    _cifar_conv("conv2", "norm1", 32, 0.01), ReLULayer("relu2", ["conv2"]),
                    PoolingLayer("pool2", ["conv2"], "AVE", 3, 2), lrn("norm2", "pool2"),
                    _cifar_conv("conv3", "norm2", 64, 0.01, params=None), ReLULayer("relu3", ["conv3"]),
                    PoolingLayer("pool3", ["conv3"], "AVE", 3, 2),
                    InnerProductLayer("ip1", ["pool3"], 10, weight_filler=gaussian(0.01), bias_filler=constant(),
                                      params=[(1.0, 250.0), (2.0, 0.0)]),
                    AccuracyLayer("accuracy", ["ip1", "label"]), SoftmaxWithLoss("loss", ["ip1", "label"]))


def caffenet_net(train_bs: int = 256, test_bs: int = 50) -> Msg:
    def conv(n, b, k, o, bias, pad=None, stride=None, group=None):
        return ConvolutionLayer(n, [b], k, o, pad=pad, stride=stride, group=group, weight_filler=gaussian(0.01),
                                bias_filler=constant(bias), params=W12D)

    def fc(n, b, o, std, bias):
        return InnerProductLayer(n, [b], o, weight_filler=gaussian(std), bias_filler=constant(bias), params=W12D)

    lrn = lambda n, b: LRNLayer(n, [b], 5, 0.0001, 0.75)
    return NetParam("CaffeNet", *_data(train_bs, test_bs, 3, 227, 227),
                    conv("conv1", "data", 11, 96, 0.0, stride=4), ReLULayer("relu1", ["conv1"]),
                    PoolingLayer("pool1", ["conv1"], "MAX", 3, 2), lrn("norm1", "pool1"),
                    conv("conv2", "norm1", 5, 256, 1.0, pad=2, group=2), ReLULayer("relu2", ["conv2"]),
                    PoolingLayer("pool2", ["conv2"], "MAX", 3, 2), lrn("norm2", "pool2"),
                    conv("conv3", "norm2", 3, 384, 0.0, pad=1), ReLULayer("relu3", ["conv3"]),
                    conv("conv4", "conv3", 3, 384, 1.0, pad=1, group=2), ReLULayer("relu4", ["conv4"]),
                    conv("conv5", "conv4", 3, 256, 1.0, pad=1, group=2), ReLULayer("relu5", ["conv5"]),
                    PoolingLayer("pool5", ["conv5"], "MAX", 3, 2),
                    fc("fc6", "pool5", 4096, 0.005, 1.0), ReLULayer("relu6", ["fc6"]), DropoutLayer("drop6", ["fc6"], 0.5),
                    fc("fc7", "fc6", 4096, 0.005, 1.0), ReLULayer("relu7", ["fc7"]), DropoutLayer("drop7", ["fc7"], 0.5),
                    fc("fc8", "fc7", 1000, 0.01, 0.0),
                    AccuracyLayer("accuracy", ["fc8", "label"]), SoftmaxWithLoss("loss", ["fc8", "label"]))


def _solver(**kw) -> Msg:
    return Msg("SolverParameter", **kw)


GPU = ENUMS["SolverMode"]["GPU"]

# solver settings of the reference's shipped solver prototxts, minus snapshot/net (ProtoLoader.scala:34-39)
SOLVERS: Dict[str, Callable[[], Msg]] = {
    "lenet": lambda: _solver(test_iter=[100], test_interval=500, base_lr=0.01, display=100, max_iter=10000,
                             lr_policy="inv", gamma=0.0001, power=0.75, momentum=0.9, weight_decay=0.0005,
                             solver_mode=GPU),
    "cifar10_quick": lambda: _solver(test_iter=[100], test_interval=500, base_lr=0.001, display=100, max_iter=4000,
                                     lr_policy="fixed", momentum=0.9, weight_decay=0.004, solver_mode=GPU,
                                     snapshot_format=ENUMS["SnapshotFormat"]["HDF5"]),
    "cifar10_full": lambda: _solver(test_iter=[100], test_interval=1000, base_lr=0.001, display=200, max_iter=60000,
                                    lr_policy="fixed", momentum=0.9, weight_decay=0.004, solver_mode=GPU,
                                    snapshot_format=ENUMS["SnapshotFormat"]["HDF5"]),
    "caffenet": lambda: _solver(test_iter=[1000], test_interval=1000, base_lr=0.01, display=20, max_iter=450000,
                                lr_policy="step", gamma=0.1, momentum=0.9, weight_decay=0.0005, stepsize=100000,
                                solver_mode=GPU),
}

NETS: Dict[str, Callable[..., Msg]] = {
    "lenet": lenet_net, "cifar10_quick": cifar10_quick_net, "cifar10_full": cifar10_full_net,
    "caffenet": caffenet_net,
}

# (channels, height, width, classes, default train batch, default test batch)
GEOMETRY: Dict[str, Tuple[int, int, int, int, int, int]] = {
    "lenet": (1, 28, 28, 10, 64, 100), "cifar10_quick": (3, 32, 32, 10, 100, 100),
    "cifar10_full": (3, 32, 32, 10, 100, 100), "caffenet": (3, 227, 227, 1000, 256, 50),
}


def solver_with_net(name: str, train_bs: int | None = None, test_bs: int | None = None, **overrides) -> Msg:
    """SolverParameter with embedded NetParameter, i.e. what CaffeNet.apply serializes
    (Net.scala:253-262).  `overrides` set extra solver fields (e.g. random_seed=1701, display=0)."""
    c, h, w, _, trb, teb = GEOMETRY[name]
    sol = SOLVERS[name]()
    sol.net_param = NETS[name](train_bs or trb, test_bs or teb)
    for k, v in overrides.items():
        setattr(sol, k, v)
    return sol
