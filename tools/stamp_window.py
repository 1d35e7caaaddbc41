"""Per-CTA phase timing of the window convolution kernels (csrc/conv_flat.cu), from clock64 stamps:
enter, set-up done, first window landed, first filter chunk landed, all MMAs issued, accumulator complete (epilogue
starts), epilogue done (warp 0), all warps done.
usage: python tools/stamp_window.py [conv2|conv3|conv4|conv5 ...]"""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sparknet_b200.caffe_proto import Msg
from sparknet_b200.layers import gaussian
from sparknet_b200.library import load_library
from tests.harness import LayerNet

SHAPES = {
    "conv2": ((256, 96, 27, 27), 256, dict(kernel_size=[5], pad=[2], group=2)),
    "conv3": ((256, 256, 13, 13), 384, dict(kernel_size=[3], pad=[1])),
    "conv4": ((256, 384, 13, 13), 384, dict(kernel_size=[3], pad=[1], group=2)),
    "conv5": ((256, 384, 13, 13), 256, dict(kernel_size=[3], pad=[1], group=2)),
}
lib = load_library()
lib.set_mode_gpu()
lib.sparknet_debug_read_stamps.restype = ctypes.c_int
lib.sparknet_debug_read_stamps.argtypes = [ctypes.POINTER(ctypes.c_longlong), ctypes.c_int]
rng = np.random.default_rng(0)
lib.sparknet_debug_set(3, int(os.environ.get('CG', '0')))
lib.sparknet_debug_set(1, int(os.environ.get('MT', '0')))
NAMES = ["setup", "window0", "filter0", "mma issue", "mma drain", "epilogue(w0)", "tail"]


def report(tag):
    buf = np.zeros((4096, 8), dtype=np.int64)
    n = lib.sparknet_debug_read_stamps(buf.ctypes.data_as(ctypes.POINTER(ctypes.c_longlong)), 4096)
    t = buf[:n]
    lead = t[::2] if os.environ.get('CG', '0') != '1' else t      # CTAs that issued MMAs (rank 0 of a pair, or all)
    tot = (t[:, 7] - t[:, 0]).astype(np.float64)
    print(f"{tag}: {n} CTAs, total {tot.mean():.0f} clk (min {tot.min():.0f} max {tot.max():.0f}); issuing CTAs {len(lead)}: "
          f"setup {np.mean(lead[:, 1] - lead[:, 0]):.0f}  mainloop {np.mean(lead[:, 4] - lead[:, 1]):.0f} "
          f"(waiting for window {lead[:, 2].mean():.0f}, for filter chunks {lead[:, 3].mean():.0f})  "
          f"epilogue {np.mean(t[:, 6] - t[:, 5]):.0f}  tail {np.mean(t[:, 7] - t[:, 6]):.0f}", flush=True)


for which in (sys.argv[1:] or ["conv3"]):
    shape, nout, kw = SHAPES[which]
    l = Msg("LayerParameter", name="l", type="Convolution", bottom=["data"], top=["l"])
    l.convolution_param = Msg("ConvolutionParameter", num_output=nout, weight_filler=gaussian(0.01), **kw)
    x = rng.standard_normal(shape).astype(np.float32)
    net = LayerNet(lib, {"data": x}, [l])
    lib.sparknet_set_precision(net.state, 1)
    for _ in range(2):
        net.net.forward(); net.net.backward()
    lib.sparknet_debug_set(2, 1)
    net.net.forward()
    report(which + " fwd")
    net.net.backward()
    report(which + " dgrad")
    lib.sparknet_debug_set(2, 0)
