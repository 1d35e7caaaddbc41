"""Host-side mirror of src/main/scala/libs/Net.scala: `WeightCollection` (:14-47) and `CaffeNet`
(:67-262) on top of the C ABI, with the same method names and semantics.  The per-element JNA
loops of getWeights/setWeights (:132-172) become one bulk copy through the same `get_data`
pointer; everything else calls exactly the entry points the Scala code calls, in the same order
(`set_mode_gpu` before train/test/forward/backward, :103,108,123,128)."""
from __future__ import annotations

import ctypes
from typing import Callable, Dict, Iterator, List, Optional, Sequence, Tuple

import numpy as np

from .caffe_proto import Msg, encode
from .library import java_callback_t, load_library


class WeightCollection:
    """Net.scala:14-47.  allWeights: layer name -> list of float32 arrays (weights, bias)."""

    def __init__(self, all_weights: Dict[str, List[np.ndarray]], layer_names: List[str]):
        self.allWeights = all_weights
        self.layerNames = layer_names
        self.numLayers = len(layer_names)

    def scalarDivide(self, v: float) -> None:                      # Net.scala:17-23 -> JavaNDArray.java:108-117
        d = np.float32(v)
        for name in self.layerNames:
            for w in self.allWeights[name]:
                np.divide(w, d, out=w)                             # true fp32 division, in place

    @staticmethod
    def add(wc1: "WeightCollection", wc2: "WeightCollection") -> "WeightCollection":   # Net.scala:27-46
        assert wc1.layerNames == wc2.layerNames
        new: Dict[str, List[np.ndarray]] = {}
        for name in wc1.layerNames:
            assert len(wc1.allWeights[name]) == len(wc2.allWeights[name])
            new[name] = []
            for a, b in zip(wc1.allWeights[name], wc2.allWeights[name]):
                assert a.shape == b.shape
                new[name].append(a + b)                            # NDArray.plus: copy then fp32 add
        return WeightCollection(new, wc1.layerNames)


class MinibatchSource:
    """What Net.scala's callbacks pull from (MinibatchSampler.scala:3-59): strictly alternating
    image / label minibatches.  `images[i]` is float32 (batch, C, H, W), `labels[i]` float32 (batch,)."""

    def __init__(self, images: Sequence[np.ndarray], labels: Sequence[np.ndarray]):
        self.images, self.labels = images, labels
        self.i_img = self.i_lab = 0

    def next_image_minibatch(self) -> np.ndarray:
        a = self.images[self.i_img % len(self.images)]
        self.i_img += 1
        return a

    def next_label_minibatch(self) -> np.ndarray:
        a = self.labels[self.i_lab % len(self.labels)]
        self.i_lab += 1
        return a


class CaffeNet:
    """Net.scala:67-262."""

    def __init__(self, state: int, lib):
        self.state, self.lib = state, lib
        self.numLayers = lib.num_layers(state)
        self.layerNames = [lib.layer_name(state, i).decode() for i in range(self.numLayers)]
        self.layerNumBlobs = [lib.num_layer_weights(state, i) for i in range(self.numLayers)]
        self._callbacks: Dict[str, object] = {}        # keep callbacks alive (Net.scala:72-76)
        self.dtypeSize = lib.get_dtype_size()
        self.intSize = lib.get_int_size()
        self.numTestBatches: Optional[int] = None
        self.copy_threads = 1          # > 1: image callbacks copy with sparknet_host_copy on that many host threads

    # -- construction (object CaffeNet.apply, Net.scala:253-262) -----------------------------
    @classmethod
    def from_solver(cls, solver: Msg | bytes, lib=None) -> "CaffeNet":
        lib = lib or load_library()
        state = lib.create_state()
        raw = solver if isinstance(solver, (bytes, bytearray)) else encode(solver)
        lib.load_solver_from_protobuf(state, bytes(raw), len(raw))
        return cls(state, lib)

    def close(self) -> None:
        if self.state:
            self.lib.destroy_state(self.state)
            self.state = None

    # -- data ---------------------------------------------------------------------------------
    def _image_cb(self, src: MinibatchSource):                      # makeImageCallback, Net.scala:194-221
        def invoke(data, batch_size, num_dims, shape):
            batch = src.next_image_minibatch()
            assert batch.shape[0] == batch_size
            size = 1
            for i in range(num_dims):
                size *= shape[i]
            assert batch.dtype == np.float32 and batch.size == batch_size * size
            if self.copy_threads > 1 and batch.nbytes >= (1 << 22):
                # large minibatch (CaffeNet: 158 MB): the product's multi-threaded host copy; the oracle has none
                self.lib.sparknet_host_copy(data, batch.ctypes.data, batch.nbytes, self.copy_threads)
            else:
                ctypes.memmove(data, batch.ctypes.data, batch.nbytes)
        return java_callback_t(invoke)

    def _byte_image_cb(self, src: MinibatchSource):
        """Callback of the byte-pixel path (sparknet_set_train_data_callback_bytes): the minibatches are int8 / uint8
        NCHW arrays, as the reference's ByteImage holds them, and are copied as they are — the device widens them."""
        def invoke(data, batch_size, num_dims, shape):
            batch = src.next_image_minibatch()
            assert batch.shape[0] == batch_size and batch.dtype.itemsize == 1
            if self.copy_threads > 1 and batch.nbytes >= (1 << 22):
                self.lib.sparknet_host_copy(data, batch.ctypes.data, batch.nbytes, self.copy_threads)
            else:
                ctypes.memmove(data, batch.ctypes.data, batch.nbytes)
        return java_callback_t(invoke)

    def _widening_image_cb(self, src: MinibatchSource, signed: bool):
        """makeImageCallback as the reference runs it (Net.scala:194-221): the source holds bytes, the callback widens
        them to float into the library's buffer (ByteImage.cropInto), here on `copy_threads` native threads."""
        def invoke(data, batch_size, num_dims, shape):
            batch = src.next_image_minibatch()
            assert batch.shape[0] == batch_size and batch.dtype.itemsize == 1
            self.lib.sparknet_host_widen(data, batch.ctypes.data, batch.size, 1 if signed else 0, max(1, self.copy_threads))
        return java_callback_t(invoke)

    def _label_cb(self, src: MinibatchSource):                      # makeLabelCallback, Net.scala:223-233
        def invoke(data, batch_size, num_dims, shape):
            lab = src.next_label_minibatch()
            assert lab.size == batch_size and lab.dtype == np.float32
            ctypes.memmove(data, lab.ctypes.data, lab.nbytes)
        return java_callback_t(invoke)

    def setTrainData(self, src: MinibatchSource) -> None:           # Net.scala:83-88
        self._callbacks["img_train"], self._callbacks["lab_train"] = self._image_cb(src), self._label_cb(src)
        assert self.lib.set_train_data_callback(self.state, 0, self._callbacks["img_train"]) == 0
        assert self.lib.set_train_data_callback(self.state, 1, self._callbacks["lab_train"]) == 0

    def setTrainDataBytes(self, src: MinibatchSource, signed: bool = True, mean: Optional[np.ndarray] = None,
                          scale: float = 1.0) -> None:
        """setTrainData for byte pixels: `src` yields int8 (signed, the JVM's byte) or uint8 NCHW minibatches; the
        library widens them on the device and subtracts `mean` (C*H*W or C float32 values) there."""
        self._callbacks["img_train"], self._callbacks["lab_train"] = self._byte_image_cb(src), self._label_cb(src)
        m = None if mean is None else np.ascontiguousarray(mean, dtype=np.float32).ravel()
        self._callbacks["mean_train"] = m
        mp = None if m is None else m.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
        assert self.lib.sparknet_set_train_data_callback_bytes(self.state, 0, self._callbacks["img_train"], 1 if signed else 0,
                                                               mp, 0 if m is None else m.size, scale) == 0
        assert self.lib.set_train_data_callback(self.state, 1, self._callbacks["lab_train"]) == 0

    def setTrainDataWidening(self, src: MinibatchSource, signed: bool = True) -> None:
        """The reference's own arrangement: byte pixels on the host, widened to float inside the fp32 callback."""
        self._callbacks["img_train"], self._callbacks["lab_train"] = self._widening_image_cb(src, signed), self._label_cb(src)
        assert self.lib.set_train_data_callback(self.state, 0, self._callbacks["img_train"]) == 0
        assert self.lib.set_train_data_callback(self.state, 1, self._callbacks["lab_train"]) == 0

    def setTestData(self, src: MinibatchSource, num_batches: int) -> None:   # Net.scala:90-96
        self.numTestBatches = num_batches
        self._callbacks["img_test"], self._callbacks["lab_test"] = self._image_cb(src), self._label_cb(src)
        assert self.lib.set_test_data_callback(self.state, 0, self._callbacks["img_test"]) == 0
        assert self.lib.set_test_data_callback(self.state, 1, self._callbacks["lab_test"]) == 0

    # -- compute ------------------------------------------------------------------------------
    def train(self, num_steps: int) -> None:                        # Net.scala:102-105
        self.lib.set_mode_gpu()
        self.lib.solver_step(self.state, num_steps)

    def test(self) -> np.ndarray:                                   # Net.scala:107-120
        self.lib.set_mode_gpu()
        assert self.numTestBatches is not None
        self.lib.solver_test(self.state, self.numTestBatches)
        n = self.lib.num_test_scores(self.state)
        return np.array([self.lib.get_test_score(self.state, i) for i in range(n)], dtype=np.float32)

    def forward(self) -> None:                                      # Net.scala:122-125
        self.lib.set_mode_gpu()
        self.lib.forward(self.state)

    def backward(self) -> None:                                     # Net.scala:127-130
        self.lib.set_mode_gpu()
        self.lib.backward(self.state)

    # -- weights / blobs ------------------------------------------------------------------------
    def _shape(self, blob) -> Tuple[int, ...]:                      # getShape, Net.scala:243-250
        return tuple(self.lib.get_axis_shape(blob, k) for k in range(self.lib.get_num_axes(blob)))

    def _view(self, blob, diff: bool = False) -> np.ndarray:
        shape = self._shape(blob)
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        ptr = (self.lib.get_diff if diff else self.lib.get_data)(blob)
        return np.ctypeslib.as_array(ptr, shape=(n,)).reshape(shape)

    def weight_blobs(self) -> Iterator[Tuple[int, int, int]]:
        for i in range(self.numLayers):
            for j in range(self.layerNumBlobs[i]):
                yield i, j, self.lib.get_weight_blob(self.state, i, j)

    def setWeights(self, wc: WeightCollection) -> None:             # Net.scala:132-150
        assert wc.numLayers == self.numLayers
        for i in range(self.numLayers):
            ws = wc.allWeights[self.layerNames[i]]
            assert len(ws) == self.layerNumBlobs[i]
            for j in range(self.layerNumBlobs[i]):
                blob = self.lib.get_weight_blob(self.state, i, j)
                view = self._view(blob)                             # get_data => host copy becomes authoritative
                assert view.shape == ws[j].shape
                view[...] = ws[j]

    def getWeights(self) -> WeightCollection:                       # Net.scala:152-172
        out: Dict[str, List[np.ndarray]] = {}
        for i in range(self.numLayers):
            out[self.layerNames[i]] = [self._view(self.lib.get_weight_blob(self.state, i, j)).copy()
                                       for j in range(self.layerNumBlobs[i])]
        return WeightCollection(out, list(self.layerNames))

    def getData(self, diff: bool = False) -> Dict[str, np.ndarray]:  # Net.scala:174-192
        out = {}
        for i in range(self.lib.num_data_blobs(self.state)):
            name = self.lib.get_data_blob_name(self.state, i).decode()
            out[name] = self._view(self.lib.get_data_blob(self.state, i), diff).copy()
        return out

    def getWeightDiffs(self) -> Dict[str, List[np.ndarray]]:
        return {self.layerNames[i]: [self._view(self.lib.get_weight_blob(self.state, i, j), True).copy()
                                     for j in range(self.layerNumBlobs[i])] for i in range(self.numLayers)}
