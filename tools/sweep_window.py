"""Sweep the window kernels' tile shape (column width n_mma, accumulators MT) per CaffeNet layer and print the kernel
duration in clocks (max over CTAs of the persistent kernel's lifetime, from the clock stamps).
usage: python tools/sweep_window.py [conv1 conv2 ...]"""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sparknet_b200.caffe_proto import Msg
from sparknet_b200.layers import gaussian
from sparknet_b200.library import load_library
from tests.harness import LayerNet

SHAPES = {
    "conv1": ((256, 3, 227, 227), 96, dict(kernel_size=[11], stride=[4])),
    "conv2": ((256, 96, 27, 27), 256, dict(kernel_size=[5], pad=[2], group=2)),
    "conv3": ((256, 256, 13, 13), 384, dict(kernel_size=[3], pad=[1])),
    "conv4": ((256, 384, 13, 13), 384, dict(kernel_size=[3], pad=[1], group=2)),
    "conv5": ((256, 384, 13, 13), 256, dict(kernel_size=[3], pad=[1], group=2)),
}
CAND = {  # (fwd n_mma list, dgrad n_mma list)
    "conv1": ([96, 48], []), "conv2": ([128, 64], [48]), "conv3": ([192, 128], [256, 128]),
    "conv4": ([192, 96], [192, 96]), "conv5": ([128, 64], [192, 96]),
}
lib = load_library()
lib.set_mode_gpu()
lib.sparknet_debug_read_stamps.restype = ctypes.c_int
lib.sparknet_debug_read_stamps.argtypes = [ctypes.POINTER(ctypes.c_longlong), ctypes.c_int]
rng = np.random.default_rng(0)


def clocks():
    buf = np.zeros((4096, 8), dtype=np.int64)
    n = lib.sparknet_debug_read_stamps(buf.ctypes.data_as(ctypes.POINTER(ctypes.c_longlong)), 4096)
    t = buf[:n]
    return int((t[:, 7] - t[:, 0]).max()), n


for which in (sys.argv[1:] or list(SHAPES)):
    shape, nout, kw = SHAPES[which]
    l = Msg("LayerParameter", name="l", type="Convolution", bottom=["data"], top=["l"])
    l.convolution_param = Msg("ConvolutionParameter", num_output=nout, weight_filler=gaussian(0.01), **kw)
    x = rng.standard_normal(shape).astype(np.float32)
    net = LayerNet(lib, {"data": x}, [l])
    lib.sparknet_set_precision(net.state, 1)
    net.net.forward(); net.net.backward()
    for direction, ns in zip(("fwd", "dgrad"), CAND[which]):
        for n_mma in ns:
            for mt in (1, 2, 4):
                if mt * n_mma > 512:
                    continue
                lib.sparknet_debug_set(4, n_mma); lib.sparknet_debug_set(1, mt)
                try:
                    net.net.forward(); net.net.backward()        # warm
                    lib.sparknet_debug_set(2, 1)
                    net.net.forward()
                    c = clocks()
                    if direction == "dgrad":
                        net.net.backward()
                        c = clocks()
                    lib.sparknet_debug_set(2, 0)
                    print(f"{which} {direction} n_mma {n_mma:3d} MT {mt}: {c[0]:8d} clk ({c[1]} CTAs)", flush=True)
                except Exception as e:
                    print(which, direction, n_mma, mt, "failed", e, flush=True)
    lib.sparknet_debug_set(4, 0); lib.sparknet_debug_set(1, 0)
