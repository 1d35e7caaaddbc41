"""Debug aid for the window convolution kernels (csrc/conv_flat.cu): one convolution layer, TF32 product vs the
CPU oracle, for each descriptor base-offset mode (SPARKNET_FLAT_BO) and forced accumulator count.
usage: python tools/debug_window.py"""
import os
import subprocess
import sys

CHILD = r'''
import os, sys
import numpy as np
sys.path.insert(0, ".")
from tests.harness import load_oracle, rel_err
from tests.test_gpu_layers import conv_layer, run_pair
from sparknet_b200.layers import gaussian
from sparknet_b200.library import load_library
oracle = load_oracle(); product = load_library()
product.sparknet_debug_set(1, int(os.environ.get("MT", "0")))
product.sparknet_debug_set(3, int(os.environ.get("CG", "0")))
shape, nout, kw = eval(os.environ["CASE"])
x = np.random.default_rng(1).standard_normal(shape).astype(np.float32)
l = conv_layer("conv", "data", nout, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5), **kw)
r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv", precision=1)
print("  fwd %.2e dx %.2e" % (rel_err(r["fwd"][1]["conv"], r["fwd"][0]["conv"]), rel_err(r["bwd"][1]["data"], r["bwd"][0]["data"])),
      "dw", ["%.2e" % rel_err(b, a) for a, b in zip(*r["wdiff"]["conv"])], flush=True)
'''
CASES = [
    "((3, 32, 16, 16), 32, dict(kernel_size=[5], pad=[2]))",
    "((4, 96, 27, 27), 256, dict(kernel_size=[5], pad=[2], group=2))",
    "((6, 256, 13, 13), 384, dict(kernel_size=[3], pad=[1]))",
    "((2, 16, 6, 6), 16, dict(kernel_size=[1]))",
]
for bo in (os.environ.get("CGS") or "1 2").split():
    for mt in (os.environ.get("MTS") or "0 1").split():
        for case in CASES:
            print("cg", bo, "mt", mt, case, flush=True)
            env = dict(os.environ, SPARKNET_CONV_FLAT="7", CG=bo, MT=mt, CASE=case)
            p = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=300)
            print(p.stdout[-600:], p.stderr[-400:] if p.returncode else "", flush=True)
