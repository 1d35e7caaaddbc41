"""TF32 tensor-core mode (tcgen05/TMEM/TMA kernels of csrc/gemm_tc.cu) against the CPU oracle.
north_star tolerance: per-layer forward and backward outputs within 1e-3 relative (normwise:
max|a-b| / max|b|).  The same cases run in FFMA mode at 1e-5 in test_gpu_layers.py."""
import numpy as np
import pytest

from sparknet_b200 import zoo
from sparknet_b200.layers import InnerProductLayer, LRNLayer, PoolingLayer, ReLULayer, gaussian
from sparknet_b200.net import CaffeNet, MinibatchSource
from tests.harness import LayerNet, assert_mostly_close, fptr, rel_err, synthetic_batches
from tests.test_gpu_layers import CONV_CASES, assert_close, conv_layer, run_pair

pytestmark = pytest.mark.gpu
F = np.float32
TOL = 1e-3


def test_tc_kernels_are_built(product):
    assert product.sparknet_tc_available() == 1


@pytest.mark.parametrize("shape,num_output,kw", CONV_CASES)
def test_convolution_tf32(oracle, product, shape, num_output, kw):
    x = np.random.default_rng(1701).standard_normal(shape).astype(F)
    l = conv_layer("conv", "data", num_output, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5), **kw)
    r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv", precision=1)
    assert_close(r["fwd"][0]["conv"], r["fwd"][1]["conv"], TOL, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["conv"])):
        assert_close(a, b, TOL, f"param diff {j}")


# stride-1 layers whose channel counts are multiples of 16 take the flat-grid TMA path (csrc/gemm_tc.cu)
FLAT_CASES = [
    ((5, 16, 9, 11), 32, dict(kernel_size=[3])),                                   # no padding: Ho = H - 2
    ((3, 32, 12, 7), 16, dict(kernel_h=3, kernel_w=5, pad_h=1, pad_w=1)),          # rectangular, pad < (k-1)/2 along w
    ((2, 48, 10, 10), 32, dict(kernel_size=[5], pad=[4])),                         # "full" padding: Ho = H + 4
    ((40, 64, 13, 13), 96, dict(kernel_size=[3], pad=[1], group=2)),               # many tiles + split reduction
    ((2, 16, 6, 6), 16, dict(kernel_size=[1])),                                    # 1x1
    ((3, 96, 8, 8), 48, dict(kernel_size=[2], pad=[1], group=3)),                  # even filter, three groups
]


@pytest.mark.parametrize("flat", ["1", "0"])
@pytest.mark.parametrize("shape,num_output,kw", FLAT_CASES)
def test_convolution_flat_grid_and_gather_paths_tf32(oracle, product, monkeypatch, shape, num_output, kw, flat):
    monkeypatch.setenv("SPARKNET_CONV_FLAT", flat)
    x = np.random.default_rng(77).standard_normal(shape).astype(F)
    l = conv_layer("conv", "data", num_output, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5), **kw)
    r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv", precision=1)
    assert_close(r["fwd"][0]["conv"], r["fwd"][1]["conv"], TOL, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["conv"])):
        assert_close(a, b, TOL, f"param diff {j}")


# stride-1 first layers with fewer than 16 input channels CAN run on the window kernels over a channel-padded grid (runtime.cpp
# ConvolutionLayer::use_s2d kind 2, SPARKNET_CONV_CPAD=1); the default keeps them on the gather kernels, which are faster there
CPAD_CASES = [
    ((4, 3, 32, 32), 32, dict(kernel_size=[5], pad=[2])),                           # cifar conv1
    ((7, 1, 12, 9), 16, dict(kernel_size=[3])),                                    # one channel, no padding
    ((3, 8, 10, 10), 48, dict(kernel_h=3, kernel_w=5, pad_h=1, pad_w=2)),          # rectangular
    ((50, 3, 32, 32), 32, dict(kernel_size=[5], pad=[2])),                          # many tiles: persistent CTAs, resident filter
]


@pytest.mark.parametrize("cpad", ["1", "0"])
@pytest.mark.parametrize("shape,num_output,kw", CPAD_CASES)
def test_convolution_channel_padded_first_layer_tf32(oracle, product, monkeypatch, shape, num_output, kw, cpad):
    monkeypatch.setenv("SPARKNET_CONV_CPAD", cpad)
    oracle.oracle_seed(4243)
    x = np.random.default_rng(78).standard_normal(shape).astype(F)
    l = conv_layer("conv", "data", num_output, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5), **kw)
    r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv", precision=1)
    assert_close(r["fwd"][0]["conv"], r["fwd"][1]["conv"], TOL, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["conv"])):
        assert_close(a, b, TOL, f"param diff {j}")


# strided, unpadded layers whose folded channel count C*s*s is a multiple of 16 run as stride-1 window convolutions on
# the space-to-depth grid (csrc/conv_flat.cu); SPARKNET_CONV_S2D=0 keeps them on the gather kernels
S2D_CASES = [
    ((3, 3, 35, 39), 16, dict(kernel_size=[11], stride=[4])),                      # conv1-like, H and W not multiples of 4
    ((2, 4, 17, 15), 32, dict(kernel_size=[5], stride=[2])),                       # stride 2, 4 -> 16 folded channels
    ((5, 16, 12, 12), 48, dict(kernel_size=[2], stride=[2])),                      # filter == stride (1x1 after folding)
    ((2, 1, 29, 29), 16, dict(kernel_size=[7], stride=[4])),                       # one input channel -> 16
]


@pytest.mark.parametrize("s2d", ["1", "0"])
@pytest.mark.parametrize("shape,num_output,kw", S2D_CASES)
def test_convolution_space_to_depth_tf32(oracle, product, monkeypatch, shape, num_output, kw, s2d):
    monkeypatch.setenv("SPARKNET_CONV_S2D", s2d)
    x = np.random.default_rng(78).standard_normal(shape).astype(F)
    l = conv_layer("conv", "data", num_output, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5), **kw)
    r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv", precision=1)
    assert_close(r["fwd"][0]["conv"], r["fwd"][1]["conv"], TOL, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["conv"])):
        assert_close(a, b, TOL, f"param diff {j}")


def test_convolution_relu_fusion_tf32(oracle, product):
    x = np.random.default_rng(5).standard_normal((2, 8, 9, 9)).astype(F)
    l = conv_layer("conv", "data", 16, kernel_size=[3], pad=[1], weight_filler=gaussian(0.1), bias_filler=gaussian(0.1))
    r = run_pair(oracle, product, {"data": x}, [l, ReLULayer("relu", ["conv"])], "conv", "conv", precision=1)
    assert np.all(r["fwd"][1]["conv"] >= 0)
    assert_close(r["fwd"][0]["conv"], r["fwd"][1]["conv"], TOL, "relu(conv)")


def test_convolution_stride2_dgrad_generic_path_tf32(oracle, product):
    x = np.random.default_rng(8).standard_normal((2, 6, 11, 9)).astype(F)
    l = conv_layer("conv", "data", 10, kernel_size=[3], stride=[2], pad=[1], weight_filler=gaussian(0.1), bias_filler=gaussian(0.1))
    r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv", precision=1)
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL, "bottom diff")


@pytest.mark.parametrize("shape,n_out", [((7, 50), 13), ((64, 50, 4, 4), 500), ((8, 256, 6, 6), 512),
                                         ((100, 64), 10), ((5, 4096), 1000), ((256, 1024), 384), ((33, 500), 10)])
def test_inner_product_tf32(oracle, product, shape, n_out):
    x = np.random.default_rng(11).standard_normal(shape).astype(F)
    l = InnerProductLayer("ip", ["data"], n_out, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5))
    r = run_pair(oracle, product, {"data": x}, [l], "ip", "ip", precision=1)
    assert_close(r["fwd"][0]["ip"], r["fwd"][1]["ip"], TOL, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["ip"])):
        assert_close(a, b, TOL, f"param diff {j}")


# shapes that take the all-TMA matrix path (csrc/conv_flat.cu tcw_matmul*): single CTA, CTA pair, two accumulators per
# CTA, partial last column tile, reduction not a multiple of 32; SPARKNET_IP_TMA=0 keeps them on the gather kernels
IP_TMA_CASES = [((96, 200), 64), ((129, 512), 320), ((300, 96), 272), ((256, 36, 4, 4), 1000), ((40, 100), 16)]


@pytest.mark.parametrize("tma", ["1", "0"])
@pytest.mark.parametrize("shape,n_out", IP_TMA_CASES)
def test_inner_product_tma_and_gather_paths_tf32(oracle, product, monkeypatch, shape, n_out, tma):
    monkeypatch.setenv("SPARKNET_IP_TMA", tma)
    x = np.random.default_rng(12).standard_normal(shape).astype(F)
    l = InnerProductLayer("ip", ["data"], n_out, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5))
    r = run_pair(oracle, product, {"data": x}, [l], "ip", "ip", precision=1)
    assert_close(r["fwd"][0]["ip"], r["fwd"][1]["ip"], TOL, "top")
    assert_close(r["bwd"][0]["data"], r["bwd"][1]["data"], TOL, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["ip"])):
        assert_close(a, b, TOL, f"param diff {j}")


@pytest.mark.parametrize("shape,kernel,stride", [((3, 16, 27, 27), 3, 2), ((2, 16, 32, 30), 3, 2), ((2, 16, 12, 12), 2, 2)])
@pytest.mark.parametrize("strip", ["0", "2", "3"])
def test_relu_backward_folded_into_max_pool_backward(oracle, product, monkeypatch, shape, kernel, stride, strip):
    """conv -> in-place ReLU -> MAX pool -> LRN, the CaffeNet pattern.  With SPARKNET_FUSE_RELU_POOL=1 (default) the pool's
    backward applies the ReLU gate and the ReLU layer's own backward is skipped: the result must be BIT-identical to the
    unfused product (same forward, same gates), and match the oracle."""
    x = np.random.default_rng(9).standard_normal(shape).astype(F)
    layers = [conv_layer("conv", "data", 32, kernel_size=[3], pad=[1], weight_filler=gaussian(0.1), bias_filler=gaussian(0.1)),
              ReLULayer("relu", ["conv"]), PoolingLayer("pool", ["conv"], "MAX", kernel, stride),
              LRNLayer("norm", ["pool"], 5, 1e-4, 0.75)]
    # fp32 FFMA mode: the forward passes agree to ~1e-6, so oracle and product open the same gates
    monkeypatch.setenv("SPARKNET_POOL_STRIP", strip)     # both forms of the 3x3 / stride-2 backward kernel apply the gate
    monkeypatch.setenv("SPARKNET_FUSE_RELU_POOL", "1")
    r = run_pair(oracle, product, {"data": x}, layers, "norm", "conv", precision=0)
    # the unfused product on the same weights and the same top gradient (the switch is read when a net is built)
    monkeypatch.setenv("SPARKNET_FUSE_RELU_POOL", "0")
    o, fused = r["nets"]
    u = LayerNet(product, {"data": x}, layers)
    product.sparknet_set_precision(u.state, 0)
    u.set_weights("conv", o.weights("conv"))
    u.forward()
    u.zero_param_diffs()
    u.set_diff("norm", np.random.default_rng(3).standard_normal(r["fwd"][0]["norm"].shape).astype(F))
    ub = u.backward()
    for k in ("conv", "data"):
        assert np.array_equal(r["bwd"][1][k], ub[k]), k
    for a, b in zip(r["wdiff"]["conv"][1], u.weights("conv", True)):
        assert np.array_equal(a, b)
    assert_close(r["fwd"][0]["norm"], r["fwd"][1]["norm"], 1e-5, "top")
    assert_mostly_close(r["bwd"][1]["conv"], r["bwd"][0]["conv"], 1e-4, "diff behind the ReLU")
    assert_mostly_close(r["bwd"][1]["data"], r["bwd"][0]["data"], 1e-4, "bottom diff")
    for j, (a, b) in enumerate(zip(*r["wdiff"]["conv"])):
        assert_mostly_close(b, a, 1e-4, f"param diff {j}")


@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("shape,group", [((3, 32, 13, 13), 1), ((2, 32, 13, 13), 2), ((2, 16, 9, 11), 1)])
def test_relu_backward_folded_into_convolution_backward_data(oracle, product, monkeypatch, shape, group, precision):
    """conv -> in-place ReLU -> conv, the CaffeNet conv3/relu3/conv4 pattern.  With SPARKNET_FUSE_RELU_CONV=1 (default) the
    second convolution's backward-data epilogue applies dx = dy * (y > 0) (relu_layer.cu:35-41) while it writes the
    gradient and the ReLU layer's own backward is skipped: BIT-identical to the unfused product in both precision modes
    (window kernels in TF32, gather / FFMA kernels with a separate gating pass otherwise), and within tolerance of the oracle."""
    x = np.random.default_rng(19).standard_normal(shape).astype(F)
    layers = [conv_layer("conva", "data", 64, kernel_size=[3], pad=[1], weight_filler=gaussian(0.1), bias_filler=gaussian(0.1)),
              ReLULayer("relu", ["conva"]),
              conv_layer("convb", "conva", 32, kernel_size=[3], pad=[1], group=group, weight_filler=gaussian(0.1), bias_filler=gaussian(0.1))]
    monkeypatch.setenv("SPARKNET_FUSE_RELU_CONV", "1")
    oracle.oracle_seed(4242)                            # the oracle draws the filters: same draw on every run
    r = run_pair(oracle, product, {"data": x}, layers, "convb", ["conva", "convb"], precision=precision)
    monkeypatch.setenv("SPARKNET_FUSE_RELU_CONV", "0")
    o, fused = r["nets"]
    u = LayerNet(product, {"data": x}, layers)
    product.sparknet_set_precision(u.state, precision)
    for wl in ("conva", "convb"):
        u.set_weights(wl, o.weights(wl))
    u.forward()
    u.zero_param_diffs()
    u.set_diff("convb", np.random.default_rng(3).standard_normal(r["fwd"][0]["convb"].shape).astype(F))
    ub = u.backward()
    for k in ("conva", "data"):
        assert np.array_equal(r["bwd"][1][k], ub[k]), k
    for wl in ("conva", "convb"):
        for a, b in zip(r["wdiff"][wl][1], u.weights(wl, True)):
            assert np.array_equal(a, b), wl
    # against the oracle.  fp32 FFMA mode: same gates, tight.  TF32: a pre-activation within rounding noise of zero may
    # open a different gate than the oracle's fp32 forward did, and ONE flipped gate moves a bias-gradient sum of ~500
    # random-sign terms by several percent, so only the relative L2 error is bounded there (bit-identity with the
    # unfused product above is the strict check of the fusion itself).
    pairs = [(r["bwd"][1]["conva"], r["bwd"][0]["conva"], "diff behind the ReLU"), (r["bwd"][1]["data"], r["bwd"][0]["data"], "bottom diff")]
    pairs += [(b, a, f"param diff {j}") for j, (a, b) in enumerate(zip(*r["wdiff"]["conva"]))]
    for got, want, what in pairs:
        if precision == 0:
            assert_mostly_close(got, want, 1e-4, what)
        else:
            l2 = float(np.linalg.norm(got.astype(np.float64) - want) / max(np.linalg.norm(want.astype(np.float64)), 1e-30))
            assert l2 <= 5e-2, (what, l2)


def test_weight_shadow_follows_host_writes_and_updates(oracle, product):
    """The TMA-fed operands read a tf32-rounded copy of the weights; it must track setWeights and SGD."""
    sol = zoo.solver_with_net("lenet", train_bs=16, test_bs=16, random_seed=1701, display=0)
    o, p = CaffeNet.from_solver(sol, oracle), CaffeNet.from_solver(sol, product)
    product.sparknet_set_precision(p.state, 1)
    imgs, labs = synthetic_batches("lenet", 1, 16)
    o.setTrainData(MinibatchSource(imgs, labs)); p.setTrainData(MinibatchSource(imgs, labs))
    p.forward()                                    # shadow built from the product's own fillers
    p.setWeights(o.getWeights())                   # host write: shadow must be rebuilt
    o.forward(); p.forward()
    do, dp = o.getData(), p.getData()
    for k in ("conv1", "conv2", "ip1", "ip2"):
        assert rel_err(dp[k], do[k]) <= TOL, k
    p.train(3)                                     # device-side updates: the shadow must follow
    p.forward()
    tf32 = p.getData()
    product.sparknet_set_precision(p.state, 0)     # same trained weights, fp32 FFMA kernels (no shadow involved)
    p.forward()
    fp32 = p.getData()
    for k in ("conv1", "conv2", "ip1", "ip2"):
        assert rel_err(tf32[k], fp32[k]) <= 3 * TOL, k
    w_before = p.getWeights().allWeights["ip1"][0]
    assert not np.array_equal(w_before, o.getWeights().allWeights["ip1"][0])   # training did change the weights
    o.close(); p.close()


@pytest.mark.parametrize("name,batch", [("lenet", 16), ("cifar10_quick", 10), ("cifar10_full", 10)])
def test_nets_forward_tf32(oracle, product, name, batch):
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0)
    o, p = CaffeNet.from_solver(sol, oracle), CaffeNet.from_solver(sol, product)
    product.sparknet_set_precision(p.state, 1)
    imgs, labs = synthetic_batches(name, 1, batch)
    o.setTrainData(MinibatchSource(imgs, labs)); p.setTrainData(MinibatchSource(imgs, labs))
    p.setWeights(o.getWeights())
    o.forward(); p.forward()
    do, dp = o.getData(), p.getData()
    for k in do:
        assert rel_err(dp[k], do[k]) <= 3 * TOL, (k, rel_err(dp[k], do[k]))     # error compounds over the depth
    o.close(); p.close()


@pytest.mark.parametrize("name,batch", [("caffenet", 3), ("cifar10_full", 10)])
def test_producer_side_packing_is_bit_identical_to_the_pack_kernels(oracle, product, monkeypatch, name, batch):
    """The packed (channels-last, tf32-rounded) input of a window convolution is written by the layer that produces the
    tensor where that is possible — a convolution's or an across-channel LRN's forward writes the next convolution's
    flat grid from its epilogue (SPARKNET_FUSE_PACK) — instead of a separate pack kernel.  Same values, same rounding:
    every blob, every blob gradient and every parameter gradient must be BIT-identical with the fusion on and off, and
    two solver steps must leave identical weights."""
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0)
    imgs, labs = synthetic_batches(name, 2, batch)
    out = []
    for fuse in ("0", "1"):
        monkeypatch.setenv("SPARKNET_FUSE_PACK", fuse)
        p = CaffeNet.from_solver(sol, product)
        product.sparknet_set_precision(p.state, 1)
        p.setTrainData(MinibatchSource(imgs, labs))
        if fuse == "0":
            w0 = p.getWeights()
        p.setWeights(w0)
        l0 = product.sparknet_kernel_launches(p.state)
        p.forward(); p.backward()
        launches = product.sparknet_kernel_launches(p.state) - l0
        data, diff, wd = p.getData(), p.getData(diff=True), p.getWeightDiffs()
        p.train(2)
        n = int(product.sparknet_num_params(p.state))
        flat = np.empty(n, dtype=F)
        assert product.sparknet_get_weights_flat(p.state, fptr(flat), n) == 0
        out.append((data, diff, wd, flat, launches))
        p.close()
    (d0, g0, w0d, f0, n0), (d1, g1, w1d, f1, n1) = out
    assert n1 == n0 - 4 if name == "caffenet" else n1 == n0      # CaffeNet: the input packs of conv2 .. conv5 are gone
    for k in d0:
        assert np.array_equal(d0[k], d1[k]), k
        assert np.array_equal(g0[k], g1[k]), k
    for ln in w0d:
        for a, b in zip(w0d[ln], w1d[ln]):
            assert np.array_equal(a, b), ln
    assert np.array_equal(f0, f1)
