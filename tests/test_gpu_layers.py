"""GPU parity, layer by layer, through the C ABI: product (libccaffe.so, CUDA) vs the CPU oracle on
identical inputs and weights.  FFMA validation mode must agree within 1e-5 relative (north_star);
integer outputs (pooling argmax masks) and averaging must agree exactly.  Run on the B200 box:
    python -m pytest tests -m gpu
"""
import ctypes

import numpy as np
import pytest

from sparknet_b200.caffe_proto import Msg
from sparknet_b200.layers import (TEST, TRAIN, DropoutLayer, InnerProductLayer, LRNLayer, PoolingLayer, ReLULayer,
                                  SoftmaxWithLoss, gaussian)
from tests.harness import LayerNet, fptr, rel_err

pytestmark = pytest.mark.gpu
F = np.float32
TOL_FFMA = 1e-5
TOL_TF32 = 1e-3


def tol(product_net):
    return TOL_TF32 if product_net.lib.sparknet_set_precision(product_net.state, PRECISION) == 1 else TOL_FFMA


PRECISION = 0   # these tests arbitrate in FFMA mode; test_gpu_tc.py repeats the contractions in TF32 mode


def conv_layer(name, bottom, num_output, **kw):
    l = Msg("LayerParameter", name=name, type="Convolution", bottom=[bottom], top=[name])
    l.convolution_param = Msg("ConvolutionParameter", num_output=num_output, **kw)
    return l


def run_pair(oracle, product, inputs, layers, top, wlayer=None, phase=TRAIN, seed=3, precision=PRECISION,
             masks=None):
    """Build the same tiny net on both libraries, copy the oracle's weights, run fwd + bwd with a
    fixed random top gradient.  Returns dict of (oracle, product) arrays."""
    o = LayerNet(oracle, inputs, layers, phase=phase)
    p = LayerNet(product, inputs, layers, phase=phase)
    p.lib.sparknet_set_precision(p.state, precision)
    if wlayer:
        for wl in ([wlayer] if isinstance(wlayer, str) else wlayer):
            p.set_weights(wl, o.weights(wl))
    for idx, m in (masks or {}).items():
        mp = m.ctypes.data_as(ctypes.POINTER(ctypes.c_uint))
        assert oracle.sparknet_inject_dropout_mask(o.state, idx, mp, m.size) == 0
        assert product.sparknet_inject_dropout_mask(p.state, idx, mp, m.size) == 0
    yo, yp = o.forward(), p.forward()
    out = {"fwd": (yo, yp)}
    if phase == TRAIN:
        G = np.random.default_rng(seed).standard_normal(yo[top].shape).astype(F)
        for n in (o, p):
            n.zero_param_diffs()
            n.set_diff(top, G)
        out["bwd"] = (o.backward(), p.backward())
        if wlayer:
            out["wdiff"] = {wl: (o.weights(wl, True), p.weights(wl, True))
                            for wl in ([wlayer] if isinstance(wlayer, str) else wlayer)}
    out["nets"] = (o, p)
    return out


def assert_close(a, b, t, what=""):
    assert a.shape == b.shape, what
    e = rel_err(b, a)
    assert e <= t, f"{what}: rel err {e:.3e} > {t:g}"


CONV_CASES = [
    # (input shape, num_output, kwargs)  — reference test shapes first, then the four nets' layer classes
    ((2, 3, 6, 4), 4, dict(kernel_size=[3], stride=[2])),
    ((2, 3, 6, 4), 3, dict(kernel_size=[3], stride=[2], group=3)),
    ((2, 3, 6, 4), 4, dict(kernel_size=[1])),
    ((3, 1, 28, 28), 20, dict(kernel_size=[5], stride=[1])),                 # LeNet conv1
    ((3, 20, 12, 12), 50, dict(kernel_size=[5], stride=[1])),                # LeNet conv2
    ((4, 3, 32, 32), 32, dict(kernel_size=[5], pad=[2], stride=[1])),        # cifar conv1
    ((3, 32, 16, 16), 32, dict(kernel_size=[5], pad=[2], stride=[1])),       # cifar conv2
    ((3, 32, 8, 8), 64, dict(kernel_size=[5], pad=[2], stride=[1])),         # cifar conv3
    ((2, 3, 227, 227), 96, dict(kernel_size=[11], stride=[4])),              # CaffeNet conv1
    ((2, 96, 27, 27), 256, dict(kernel_size=[5], pad=[2], group=2)),         # CaffeNet conv2
    ((2, 256, 13, 13), 384, dict(kernel_size=[3], pad=[1])),                 # CaffeNet conv3
    ((2, 384, 13, 13), 384, dict(kernel_size=[3], pad=[1], group=2)),        # CaffeNet conv4
    ((2, 384, 13, 13), 256, dict(kernel_size=[3], pad=[1], group=2)),        # CaffeNet conv5
    ((2, 5, 9, 7), 6, dict(kernel_h=3, kernel_w=2, stride_h=2, stride_w=1, pad_h=1, pad_w=0)),   # rectangular
]


@pytest.mark.parametrize("shape,num_output,kw", CONV_CASES)
def test_convolution_fwd_bwd(oracle, product, shape, num_output, kw):
    x = np.random.default_rng(1701).standard_normal(shape).astype(F)
    l = conv_layer("conv", "data", num_output, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5), **kw)
    r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv")
    assert_close(r["fwd"][0]["conv"], r["fwd"][1]["conv"], TOL_FFMA, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL_FFMA, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["conv"])):
        assert_close(a, b, TOL_FFMA, f"param diff {j}")


def test_convolution_relu_fusion_matches_unfused_oracle(oracle, product):
    x = np.random.default_rng(5).standard_normal((2, 8, 9, 9)).astype(F)
    l = conv_layer("conv", "data", 16, kernel_size=[3], pad=[1], weight_filler=gaussian(0.1), bias_filler=gaussian(0.1))
    r = run_pair(oracle, product, {"data": x}, [l, ReLULayer("relu", ["conv"])], "conv", "conv")
    assert np.all(r["fwd"][1]["conv"] >= 0)
    assert_close(r["fwd"][0]["conv"], r["fwd"][1]["conv"], TOL_FFMA, "relu(conv)")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL_FFMA, "bottom diff")
    for a, b in zip(*r["wdiff"]["conv"]):
        assert_close(a, b, TOL_FFMA, "param diff")


def test_backward_accumulates_param_diffs_like_the_reference(oracle, product):
    # raw backward() never clears (libccaffe/ccaffe.cpp:224-228): two calls double the weight gradient
    x = np.random.default_rng(2).standard_normal((2, 3, 6, 4)).astype(F)
    l = conv_layer("conv", "data", 4, kernel_size=[3], weight_filler=gaussian(0.1), bias_filler=gaussian(0.1))
    r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv")
    o, p = r["nets"]
    once = p.weights("conv", True)
    o.backward(); p.backward()
    for a, b, c in zip(o.weights("conv", True), p.weights("conv", True), once):
        assert_close(a, b, TOL_FFMA, "accumulated")
        assert_close(2 * c, b, 1e-6, "doubled")


@pytest.mark.parametrize("shape,n_out", [((7, 50), 13), ((64, 50, 4, 4), 500), ((8, 256, 6, 6), 512),
                                         ((100, 64), 10), ((5, 4096), 1000)])
def test_inner_product_fwd_bwd(oracle, product, shape, n_out):
    x = np.random.default_rng(11).standard_normal(shape).astype(F)
    l = InnerProductLayer("ip", ["data"], n_out, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5))
    r = run_pair(oracle, product, {"data": x}, [l], "ip", "ip")
    assert_close(r["fwd"][0]["ip"], r["fwd"][1]["ip"], TOL_FFMA, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL_FFMA, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["ip"])):
        assert_close(a, b, TOL_FFMA, f"param diff {j}")


POOL_CASES = [
    ((2, 2, 3, 5), dict(kernel_size=2)),                       # reference literal shape
    ((2, 2, 6, 6), dict(kernel_h=3, kernel_w=2)),
    ((1, 1, 3, 3), dict(kernel_size=3, stride=2, pad=2)),      # padded
    ((2, 3, 6, 5), dict(kernel_size=3, stride=2)),             # ceil rule 6x5 -> 3x2
    ((3, 20, 24, 24), dict(kernel_size=2, stride=2)),          # LeNet pool1
    ((3, 32, 32, 32), dict(kernel_size=3, stride=2)),          # cifar pool1 (32 -> 16, ceil)
    ((2, 96, 55, 55), dict(kernel_size=3, stride=2)),          # CaffeNet pool1 (55 -> 27)
    ((2, 256, 13, 13), dict(kernel_size=3, stride=2)),         # CaffeNet pool5 (13 -> 6)
    # the tile kernels' other regimes: planes too large for one shared-memory tile are cut into row bands ...
    ((1, 3, 120, 100), dict(kernel_size=3, stride=2, pad=1)),  # 12000 floats per plane: bands, padded
    ((2, 2, 90, 80), dict(kernel_size=5, stride=3)),           # window wider than the stride, ceil rule on both axes
    ((1, 2, 70, 90), dict(kernel_h=70, kernel_w=90)),          # global pooling: one window = the whole plane
    ((2, 3, 33, 35), dict(kernel_h=4, kernel_w=3, stride=1, pad=1)),   # stride 1: every input under up to 12 windows
    ((3, 5, 7, 6), dict(kernel_size=2, stride=2, pad=1)),      # ... and many small planes per tile, odd sizes
    ((1, 1, 200, 160), dict(kernel_size=2, stride=2, pad=1)),  # 2x2 / stride 2 WITH padding: not the fixed-shape kernel
]


@pytest.mark.parametrize("method", [0, 1])
@pytest.mark.parametrize("shape,pp", POOL_CASES)
def test_pooling_fwd_bwd_and_mask(oracle, product, shape, pp, method):
    rng = np.random.default_rng(21)
    x = rng.standard_normal(shape).astype(F)
    x[..., ::3] = np.round(x[..., ::3])                        # plant ties: first-max-wins must agree
    l = Msg("LayerParameter", name="pool", type="Pooling", bottom=["data"], top=["pool"])
    l.pooling_param = Msg("PoolingParameter", pool=method, **pp)
    r = run_pair(oracle, product, {"data": x}, [l], "pool")
    yo, yp = r["fwd"][0]["pool"], r["fwd"][1]["pool"]
    assert yo.shape == yp.shape
    if method == 0:
        assert np.array_equal(yo, yp)
        o, p = r["nets"]
        mo = np.ctypeslib.as_array(oracle.oracle_pool_mask(o.state, 1), shape=(yo.size,)).copy()
        mp = np.empty(yo.size, dtype=np.int32)
        assert product.sparknet_get_pool_mask(p.state, 1, mp.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), mp.size) == 0
        assert np.array_equal(mo, mp)
        assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], 1e-6, "bottom diff")
    else:
        assert_close(yo, yp, TOL_FFMA, "top")
        assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL_FFMA, "bottom diff")


@pytest.mark.parametrize("shape", [(3, 32, 32, 32), (2, 96, 55, 55), (2, 256, 13, 13), (2, 8, 27, 27), (1, 3, 6, 5), (1, 5, 64, 63), (3, 7, 9, 16)])
def test_max_pool_3x3s2_backward_strip_and_block_forms_agree(oracle, product, monkeypatch, shape):
    """The 3x3 / stride-2 max-pool backward has three kernels (kernels.cu: a 2x2 block per thread; a column strip per thread
    that carries the pooled row above in registers; a lane group per plane with shuffled, contiguous stores — large layers
    take the last).  All must give the
    oracle's bottom diff (pooling_layer.cpp:253-270: every input sums the top diffs of the windows whose argmax it is,
    ties planted) and be BIT-identical to each other."""
    rng = np.random.default_rng(33)
    x = rng.standard_normal(shape).astype(F)
    x[..., ::3] = np.round(x[..., ::3])
    l = Msg("LayerParameter", name="pool", type="Pooling", bottom=["data"], top=["pool"])
    l.pooling_param = Msg("PoolingParameter", pool=0, kernel_size=3, stride=2)
    out = {}
    for mode in ("0", "2", "3"):                               # block, strip, row form
        monkeypatch.setenv("SPARKNET_POOL_STRIP", mode)
        r = run_pair(oracle, product, {"data": x}, [l], "pool")
        assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], 1e-6, "bottom diff, mode " + mode)
        out[mode] = r["bwd"][1]["data"]
    assert np.array_equal(out["0"], out["2"]) and np.array_equal(out["0"], out["3"])


def test_pool_literals_on_device(product):
    # caffe/src/caffe/test/test_pooling_layer.cpp:45-115 directly against the CUDA path
    plane = np.array([[1, 2, 5, 2, 3], [9, 4, 1, 4, 8], [1, 2, 5, 2, 3]], dtype=F)
    x = np.broadcast_to(plane, (2, 2, 3, 5)).copy()
    l = Msg("LayerParameter", name="pool", type="Pooling", bottom=["data"], top=["pool"])
    l.pooling_param = Msg("PoolingParameter", pool=0, kernel_size=2)
    net = LayerNet(product, {"data": x}, [l])
    top = net.forward()["pool"]
    assert np.array_equal(top, np.broadcast_to(np.array([[9, 5, 5, 8], [9, 5, 5, 8]], dtype=F), top.shape))
    m = np.empty(top.size, dtype=np.int32)
    assert product.sparknet_get_pool_mask(net.state, 1, m.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), m.size) == 0
    assert np.array_equal(m.reshape(top.shape), np.broadcast_to(np.array([[5, 2, 2, 9], [5, 12, 12, 9]]), top.shape))


@pytest.mark.parametrize("shape", [(2, 3, 4, 5), (3, 96, 55, 55), (1, 7)])
def test_relu_fwd_bwd(oracle, product, shape):
    x = np.random.default_rng(31).standard_normal(shape).astype(F)
    l = Msg("LayerParameter", name="relu", type="ReLU", bottom=["data"], top=["relu"])
    r = run_pair(oracle, product, {"data": x}, [l], "relu")
    assert np.array_equal(r["fwd"][0]["relu"], r["fwd"][1]["relu"])
    assert np.array_equal(r["bwd"][0]["data"], r["bwd"][1]["data"])


@pytest.mark.parametrize("shape,size,alpha,beta,within", [
    ((2, 7, 3, 3), 5, 1.0, 0.75, False),                      # reference test shape (test_lrn_layer.cpp:24-37)
    ((2, 7, 3, 3), 3, 1.0, 0.75, True),
    ((2, 96, 27, 27), 5, 1e-4, 0.75, False),                  # CaffeNet norm1
    ((2, 256, 13, 13), 5, 1e-4, 0.75, False),                 # CaffeNet norm2
    ((3, 32, 16, 16), 3, 5e-5, 0.75, True),                   # cifar10_full norm1
    ((3, 32, 8, 8), 3, 5e-5, 0.75, True),                     # cifar10_full norm2
    ((2, 20, 5, 6), 7, 1e-2, 0.75, False),                    # the chunk kernels' other window sizes (odd, up to 11) ...
    ((2, 37, 4, 3), 9, 1e-2, 0.5, False),                     # ... with a channel count that is no multiple of the chunk
    ((1, 11, 6, 6), 11, 1e-2, 1.25, False),                   # window as wide as the tensor is deep
])
def test_lrn_fwd_bwd(oracle, product, shape, size, alpha, beta, within):
    x = (np.random.default_rng(1701).standard_normal(shape) * (30 if alpha < 1e-3 else 1)).astype(F)
    l = LRNLayer("norm", ["data"], size, alpha, beta, "WITHIN_CHANNEL" if within else None)
    r = run_pair(oracle, product, {"data": x}, [l], "norm")
    assert_close(r["fwd"][0]["norm"], r["fwd"][1]["norm"], TOL_FFMA, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL_FFMA, "bottom diff")


def test_dropout_injected_mask_and_statistics(oracle, product):
    x = np.random.default_rng(41).standard_normal((8, 4096)).astype(F)
    mask = (np.random.default_rng(42).random(x.size) > 0.5).astype(np.uint32)
    l = DropoutLayer("drop", ["data"], 0.5)
    r = run_pair(oracle, product, {"data": x}, [l], "data", masks={1: mask})
    assert np.array_equal(r["fwd"][0]["data"], r["fwd"][1]["data"])
    assert np.array_equal(r["bwd"][0]["data"], r["bwd"][1]["data"])
    # own Philox stream: keep probability and scale (test_neuron_layer.cpp:467-512)
    p = LayerNet(product, {"data": np.ones((64, 4096), dtype=F)}, [DropoutLayer("drop", ["data"], 0.5)])
    y1 = p.forward()["data"].copy()
    y2 = p.forward()["data"].copy()
    kept = y1 != 0
    assert np.allclose(y1[kept], 2.0) and abs(kept.mean() - 0.5) < 0.01
    assert not np.array_equal(y1, y2)                           # a fresh mask every pass
    t = LayerNet(product, {"data": x}, [DropoutLayer("drop", ["data"], 0.5)], phase=TEST, force_backward=False)
    assert np.array_equal(t.forward()["data"], x)               # TEST phase: identity


@pytest.mark.parametrize("rows,classes", [(10, 5), (100, 10), (64, 1000), (256, 1000)])
def test_softmax_loss_fwd_bwd(oracle, product, rows, classes):
    rng = np.random.default_rng(1701)
    x = (rng.standard_normal((rows, classes)) * 10).astype(F)
    lab = rng.integers(0, classes, size=(rows, 1)).astype(F)
    l = SoftmaxWithLoss("loss", ["data", "label"])
    o = LayerNet(oracle, {"data": x, "label": lab}, [l])
    p = LayerNet(product, {"data": x, "label": lab}, [l])
    lo, lp_ = float(o.forward()["loss"]), float(p.forward()["loss"])
    assert abs(lo - lp_) <= 1e-5 * max(1.0, abs(lo))
    assert abs(product.sparknet_last_loss(p.state) - lo) <= 1e-5 * max(1.0, abs(lo))
    do, dp = o.backward()["data"], p.backward()["data"]
    assert_close(do, dp, TOL_FFMA, "bottom diff")


def test_accuracy_and_split_in_test_phase(oracle, product):
    rng = np.random.default_rng(7)
    x = rng.standard_normal((50, 10)).astype(F)
    x[:, 3] = x[:, 7]                                          # ties: (value, index) ordering must agree
    lab = rng.integers(0, 10, size=(50, 1)).astype(F)
    layers = [InnerProductLayer("ip", ["data"], 10, weight_filler=gaussian(1.0)),
              Msg("LayerParameter", name="accuracy", type="Accuracy", bottom=["ip", "label"], top=["accuracy"]),
              SoftmaxWithLoss("loss", ["ip", "label"])]
    o = LayerNet(oracle, {"data": x, "label": lab}, layers, phase=TEST, force_backward=False)
    p = LayerNet(product, {"data": x, "label": lab}, layers, phase=TEST, force_backward=False)
    p.set_weights("ip", o.weights("ip"))
    yo, yp = o.forward(), p.forward()
    assert sorted(yo) == sorted(yp)                            # incl. the auto-inserted split blobs
    assert float(yo["accuracy"]) == float(yp["accuracy"])
    assert abs(float(yo["loss"]) - float(yp["loss"])) < 1e-5 * max(1, abs(float(yo["loss"])))
    assert o.net.layerNames == p.net.layerNames
