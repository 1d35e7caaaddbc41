"""Shared test harness: drives the CPU oracle and the product through the *same* C ABI.

The oracle (oracle/liboracle_ccaffe.so) is test infrastructure; the product is
sparknet_b200/csrc/libccaffe.so.  `LayerNet` wraps one or a few layers in a NetParameter with
JavaData inputs so single layers can be exercised through load_net_from_protobuf / forward /
backward exactly as the reference's FeaturizerApp uses them (libccaffe/ccaffe.cpp:136-140,218-228).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Dict, List, Optional, Sequence

import numpy as np

from sparknet_b200.caffe_proto import Msg, encode
from sparknet_b200.layers import RDDLayer, TRAIN, TEST
from sparknet_b200.library import java_callback_t, load_library
from sparknet_b200.net import CaffeNet, MinibatchSource

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle_ccaffe.so")


def load_oracle():
    src = os.path.join(ROOT, "oracle", "oracle_ccaffe.cpp")
    hdr = os.path.join(ROOT, "sparknet_b200", "csrc", "caffe_proto.hpp")
    if (not os.path.exists(ORACLE_SO)
            or os.path.getmtime(ORACLE_SO) < max(os.path.getmtime(src), os.path.getmtime(hdr))):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    return load_library(ORACLE_SO, strict=False)


def find_openblas():
    """The OpenBLAS inside the scipy / numpy wheel and the name of its cblas_sgemm (32-bit-int interface only is usable
    by oracle_use_blas).  The reference links OpenBLAS; the image has no other BLAS."""
    import glob
    import sys
    for pat in ("scipy.libs/libscipy_openblas*.so", "numpy.libs/libscipy_openblas*.so"):
        for sp in sys.path:
            hits = glob.glob(os.path.join(sp, pat))
            if hits:
                sym = "scipy_cblas_sgemm64_" if "openblas64_" in os.path.basename(hits[0]) else "scipy_cblas_sgemm"
                return hits[0], sym
    return None, None


def fptr(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


class LayerNet:
    """A tiny net: JavaData inputs -> `layers`, loaded with load_net_from_protobuf.

    inputs: dict name -> float32 array (full blob incl. batch dim), fed through the callback ABI.
    force_backward makes every layer (and every non-label bottom) take part in backward()."""

    def __init__(self, lib, inputs: Dict[str, np.ndarray], layers: Sequence[Msg], phase: int = TRAIN,
                 force_backward: bool = True):
        self.lib = lib
        self.inputs = {k: np.ascontiguousarray(v, dtype=np.float32) for k, v in inputs.items()}
        net = Msg("NetParameter", name="t")
        if force_backward:
            net.force_backward = True
        net.state = Msg("NetState", phase=phase)
        net.layer = [RDDLayer(k, list(v.shape), None) for k, v in self.inputs.items()] + list(layers)
        self.param = net
        self.state = lib.create_state()
        raw = encode(net)
        lib.load_net_from_protobuf(self.state, raw, len(raw))
        self._cbs = []
        for idx, (k, v) in enumerate(self.inputs.items()):
            def make(arr):
                def invoke(data, batch, ndims, shape):
                    assert batch == arr.shape[0]
                    ctypes.memmove(data, arr.ctypes.data, arr.nbytes)
                return java_callback_t(invoke)
            cb = make(v)
            self._cbs.append(cb)
            assert lib.set_train_data_callback(self.state, idx, cb) == 0
        self.net = CaffeNet(self.state, lib)

    def set_input(self, name: str, value: np.ndarray) -> None:
        self.inputs[name][...] = value

    def forward(self) -> Dict[str, np.ndarray]:
        self.net.forward()
        return self.net.getData()

    def blob(self, name: str):
        for i in range(self.lib.num_data_blobs(self.state)):
            if self.lib.get_data_blob_name(self.state, i).decode() == name:
                return self.lib.get_data_blob(self.state, i)
        raise KeyError(name)

    def set_diff(self, name: str, value: np.ndarray) -> None:
        v = self.net._view(self.blob(name), diff=True)
        v[...] = value

    def zero_param_diffs(self) -> None:
        for _, _, blob in self.net.weight_blobs():
            self.net._view(blob, diff=True)[...] = 0

    def set_weights(self, layer: str, arrays: Sequence[np.ndarray]) -> None:
        i = self.net.layerNames.index(layer)
        for j, a in enumerate(arrays):
            v = self.net._view(self.lib.get_weight_blob(self.state, i, j))
            v[...] = np.asarray(a, dtype=np.float32).reshape(v.shape)

    def weights(self, layer: str, diff: bool = False) -> List[np.ndarray]:
        i = self.net.layerNames.index(layer)
        return [self.net._view(self.lib.get_weight_blob(self.state, i, j), diff).copy()
                for j in range(self.net.layerNumBlobs[i])]

    def backward(self) -> Dict[str, np.ndarray]:
        self.net.backward()
        return self.net.getData(diff=True)

    def close(self):
        self.net.close()


def synthetic_batches(name: str, n_batches: int, batch: Optional[int] = None, seed: int = 1701):
    """Synthetic minibatches of the config's shape: uniform [0,255)-128 for
    cifar/CaffeNet, uniform [0,1) for LeNet; labels uniform ints stored as float."""
    from sparknet_b200.zoo import GEOMETRY
    c, h, w, classes, trb, _ = GEOMETRY[name]
    b = batch or trb
    rng = np.random.default_rng(seed)
    imgs, labs = [], []
    for _ in range(n_batches):
        x = rng.random((b, c, h, w), dtype=np.float32)
        if name != "lenet":
            x = x * np.float32(255.0) - np.float32(128.0)
        imgs.append(np.ascontiguousarray(x, dtype=np.float32))
        labs.append(rng.integers(0, classes, size=(b,)).astype(np.float32))
    return imgs, labs


def rel_err(a: np.ndarray, b: np.ndarray) -> float:
    """Normwise relative error max|a-b| / max(max|b|, tiny) used for the 1e-5 / 1e-3 gates."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    denom = max(float(np.max(np.abs(b))) if b.size else 0.0, 1e-30)
    return float(np.max(np.abs(a - b))) / denom if a.size else 0.0


def assert_mostly_close(got: np.ndarray, want: np.ndarray, tol: float, what="", max_bad_frac: float = 5e-4) -> None:
    """Net-level comparison that tolerates max-pool argmax flips: when two inputs of a pooling window
    differ by less than the summation-order noise of the convolution before it, oracle and device may
    route the gradient to different (equally valid) elements.  Every element must agree within
    tol * max|want| except for at most `max_bad_frac` of them, and the relative L2 error must stay
    below 100 * tol.  (Layer-level tests use exact masks and the strict metric.)"""
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    assert got.shape == want.shape, what
    if want.size == 0:
        return
    scale = max(float(np.max(np.abs(want))), 1e-30)
    bad = np.abs(got - want) > tol * scale
    frac = float(np.mean(bad))
    l2 = float(np.linalg.norm(got - want) / max(np.linalg.norm(want), 1e-30))
    assert frac <= max_bad_frac and l2 <= 100 * tol, f"{what}: {frac:.2e} of elements off by > {tol:g}, rel L2 {l2:.2e}"
