"""One CaffeNet-shaped convolution layer (forward + backward) on the CUDA path, for ncu captures.
usage: python tools/prof_conv.py [conv1|conv2|conv3|conv4|conv5|fc6|fc7] [iters]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sparknet_b200.caffe_proto import Msg
from sparknet_b200.layers import InnerProductLayer, gaussian
from sparknet_b200.library import load_library
from tests.harness import LayerNet

SHAPES = {
    "conv1": ((256, 3, 227, 227), 96, dict(kernel_size=[11], stride=[4])),
    "conv2": ((256, 96, 27, 27), 256, dict(kernel_size=[5], pad=[2], group=2)),
    "conv3": ((256, 256, 13, 13), 384, dict(kernel_size=[3], pad=[1])),
    "conv4": ((256, 384, 13, 13), 384, dict(kernel_size=[3], pad=[1], group=2)),
    "conv5": ((256, 384, 13, 13), 256, dict(kernel_size=[3], pad=[1], group=2)),
}
which = sys.argv[1] if len(sys.argv) > 1 else "conv2"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
lib = load_library()
lib.set_mode_gpu()
rng = np.random.default_rng(0)
if which.startswith("conv"):
    shape, nout, kw = SHAPES[which]
    l = Msg("LayerParameter", name="l", type="Convolution", bottom=["data"], top=["l"])
    l.convolution_param = Msg("ConvolutionParameter", num_output=nout, weight_filler=gaussian(0.01), **kw)
else:
    shape, nout = {"fc6": ((256, 9216), 4096), "fc7": ((256, 4096), 4096)}[which]
    l = InnerProductLayer("l", ["data"], nout, weight_filler=gaussian(0.01))
x = rng.standard_normal(shape).astype(np.float32)
net = LayerNet(lib, {"data": x}, [l])
lib.sparknet_set_precision(net.state, 1)
for _ in range(iters):
    net.net.forward()
    net.net.backward()
lib.sparknet_sync(net.state)
print("done", which)
