import os, sys
import numpy as np
sys.path.insert(0, ".")
from tests.harness import load_oracle
from tests.test_gpu_layers import conv_layer, run_pair
from sparknet_b200.layers import gaussian, ReLULayer, PoolingLayer, LRNLayer
from sparknet_b200.library import load_library
oracle = load_oracle(); product = load_library()
F = np.float32
x = np.random.default_rng(9).standard_normal((3, 16, 27, 27)).astype(F)
layers = [conv_layer("conv", "data", 32, kernel_size=[3], pad=[1], weight_filler=gaussian(0.1), bias_filler=gaussian(0.1)),
          ReLULayer("relu", ["conv"]), PoolingLayer("pool", ["conv"], "MAX", 3, 2), LRNLayer("norm", ["pool"], 5, 1e-4, 0.75)]
runs = {}
for fuse in ("1", "0", "1"):
    os.environ["SPARKNET_FUSE_RELU_POOL"] = fuse
    r = run_pair(oracle, product, {"data": x}, layers, "norm", "conv", precision=1)
    runs.setdefault(fuse, []).append(r)
a, b, c = runs["1"][0], runs["0"][0], runs["1"][1]
for k in ("conv", "pool", "data"):
    print(k, "fused vs unfused", np.abs(a["bwd"][1][k] - b["bwd"][1][k]).max(), "fused vs fused", np.abs(a["bwd"][1][k] - c["bwd"][1][k]).max(),
          "n differing", int((a["bwd"][1][k] != b["bwd"][1][k]).sum()), "of", a["bwd"][1][k].size)
print("fwd conv equal", np.array_equal(a["fwd"][1]["conv"], b["fwd"][1]["conv"]))
d = a["bwd"][1]["conv"] != b["bwd"][1]["conv"]
idx = np.argwhere(d)[:5]
for i in idx:
    i = tuple(i)
    print(i, a["bwd"][1]["conv"][i], b["bwd"][1]["conv"][i], "y=", a["fwd"][1]["conv"][i])
