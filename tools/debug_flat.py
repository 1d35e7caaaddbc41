"""Debug aid: one convolution layer, TF32 product vs oracle, flat-grid paths enabled one at a time.
usage: python tools/debug_flat.py [mask ...]   (mask bits: 1 fwd, 2 wgrad, 4 dgrad)"""
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
shape, nout, kw = eval(os.environ["CASE"])
x = np.random.default_rng(1).standard_normal(shape).astype(np.float32)
l = conv_layer("conv", "data", nout, weight_filler=gaussian(0.05), bias_filler=gaussian(0.5), **kw)
r = run_pair(oracle, product, {"data": x}, [l], "conv", "conv", precision=1)
print("  fwd", rel_err(r["fwd"][1]["conv"], r["fwd"][0]["conv"]), "dx", rel_err(r["bwd"][1]["data"], r["bwd"][0]["data"]),
      "dw", [rel_err(b, a) for a, b in zip(*r["wdiff"]["conv"])], flush=True)
'''
CASES = [
    "((3, 32, 16, 16), 32, dict(kernel_size=[5], pad=[2]))",
    "((2, 96, 27, 27), 256, dict(kernel_size=[5], pad=[2], group=2))",
    "((2, 16, 6, 6), 16, dict(kernel_size=[1]))",
]
masks = (os.environ.get("MASKS") or "0 1 2 4").split()
if len(sys.argv) > 1:
    CASES = sys.argv[1:]
for case in CASES:
    for m in masks:
        print(case, "mask", m, flush=True)
        env = dict(os.environ, SPARKNET_CONV_FLAT=m, CASE=case)
        p = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=300)
        print(p.stdout[-600:], p.stderr[-300:] if p.returncode else "", flush=True)
