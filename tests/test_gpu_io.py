"""caffemodel / solverstate wire format: save_weights_to_file, load_weights_from_file,
restore_solver_from_file and the snapshot extension, checked against the Python codec of the reference schema
(sparknet_b200/caffe_proto.py, pinned byte-for-byte by tests/test_proto.py) and by a train / snapshot / restore /
continue round trip that must be bit-identical to the uninterrupted run.
Reference: libccaffe/ccaffe.cpp:261-273, caffe/src/caffe/net.cpp:805-840,911-923, caffe/src/caffe/blob.cpp:390-534,
caffe/src/caffe/solvers/sgd_solver.cpp:255-318."""
import os

import numpy as np
import pytest

from sparknet_b200 import zoo
from sparknet_b200.caffe_proto import Msg, decode, encode
from sparknet_b200.net import CaffeNet, MinibatchSource
from tests.harness import fptr, synthetic_batches

pytestmark = pytest.mark.gpu
F = np.float32


def make_net(product, name="lenet", batch=8, n_batches=8, first_batch=0):
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0)
    p = CaffeNet.from_solver(sol, product)
    product.sparknet_set_precision(p.state, 0)
    imgs, labs = synthetic_batches(name, n_batches, batch)
    p.setTrainData(MinibatchSource(imgs[first_batch:] + imgs[:first_batch], labs[first_batch:] + labs[:first_batch]))
    return p


def varint(v):
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def field(number, payload: bytes) -> bytes:            # length-delimited field
    return varint((number << 3) | 2) + varint(len(payload)) + payload


def test_save_weights_is_a_netparameter_with_every_layers_blobs(product, tmp_path):
    p = make_net(product)
    p.train(2)
    path = str(tmp_path / "lenet.caffemodel")
    product.save_weights_to_file(p.state, path.encode())
    net = decode("NetParameter", open(path, "rb").read())
    w = p.getWeights()
    names = [l.name for l in net.layer]
    assert names == w.layerNames                           # every layer of the TRAIN net, in order (net.cpp:911-923)
    for l in net.layer:
        want = w.allWeights[l.name]
        assert len(l.blobs) == len(want)
        for b, a in zip(l.blobs, want):
            assert list(b.shape.dim) == list(a.shape)      # Blob::ToProto writes the shape message, not the legacy fields
            assert np.array_equal(np.asarray(b.data, dtype=F).reshape(a.shape), a)
            assert len(b.diff) == 0                         # snapshot_diff defaults to false
    conv1 = [l for l in net.layer if l.name == "conv1"][0]  # the layer's own parameters travel with it (layer.hpp:507-514)
    assert conv1.type == "Convolution" and conv1.convolution_param.num_output == 20
    p.close()


def test_load_weights_by_layer_name_with_current_legacy_and_v1_encodings(product, tmp_path):
    p = make_net(product)
    w = p.getWeights()
    rng = np.random.default_rng(4)
    new = {n: [rng.standard_normal(a.shape).astype(F) for a in w.allWeights[n]] for n in ("conv1", "ip1", "ip2")}

    def blob(a, legacy=False):
        if legacy:                                          # deprecated 4-D fields, indexed from the END of the shape (blob.cpp:390-404)
            dims = [1] * (4 - a.ndim) + list(a.shape)
            return Msg("BlobProto", num=dims[0], channels=dims[1], height=dims[2], width=dims[3], data=a.ravel().tolist())
        return Msg("BlobProto", shape=Msg("BlobShape", dim=list(a.shape)), data=a.ravel().tolist())

    layers = [Msg("LayerParameter", name="conv1", type="Convolution", blobs=[blob(a) for a in new["conv1"]]),
              Msg("LayerParameter", name="ip1", type="InnerProduct", blobs=[blob(a, legacy=True) for a in new["ip1"]]),
              Msg("LayerParameter", name="not_in_this_net", type="InnerProduct", blobs=[blob(np.zeros((3, 3), F))])]
    raw = encode(Msg("NetParameter", name="source", layer=layers))
    # a V1LayerParameter entry (NetParameter.layers = 2: name = 4, blobs = 6), the encoding of the published caffemodels
    v1 = field(4, b"ip2") + b"".join(field(6, encode(blob(a, legacy=True))) for a in new["ip2"])
    raw += field(2, v1)
    path = str(tmp_path / "source.caffemodel")
    open(path, "wb").write(raw)
    product.load_weights_from_file(p.state, path.encode())
    got = p.getWeights()
    for n in new:
        for a, b in zip(new[n], got.allWeights[n]):
            assert np.array_equal(a, b), n
    for a, b in zip(w.allWeights["conv2"], got.allWeights["conv2"]):
        assert np.array_equal(a, b)                         # layers absent from the file keep their weights
    p.forward()                                             # and the device copies (incl. the tf32 shadow) follow the host write
    p.close()


def test_snapshot_restore_continue_is_bit_identical_to_the_uninterrupted_run(product, tmp_path):
    model, state = str(tmp_path / "it3.caffemodel"), str(tmp_path / "it3.solverstate")
    a = make_net(product)
    a.train(3)
    assert product.sparknet_snapshot(a.state, model.encode(), state.encode()) == 0
    a.train(2)
    st = decode("SolverState", open(state, "rb").read())
    assert st.iter == 3 and st.learned_net == model and len(st.history) == 8
    b = make_net(product, first_batch=3)                    # a fresh process would continue with the 4th minibatch
    product.restore_solver_from_file(b.state, state.encode())
    assert product.sparknet_get_iter(b.state) == 3
    b.train(2)                                              # LeNet's "inv" learning rate depends on iter: it must have been restored
    wa, wb = a.getWeights(), b.getWeights()
    for n in wa.layerNames:
        for x, y in zip(wa.allWeights[n], wb.allWeights[n]):
            assert np.array_equal(x, y), n
    blobs = [x for n in wa.layerNames for x in wa.allWeights[n]]    # learnable parameters in net order
    for i, x in enumerate(blobs):                                   # the momentum history went through the file too
        ha, hb = np.empty(x.size, F), np.empty(x.size, F)
        assert product.sparknet_get_history(a.state, i, fptr(ha), x.size) == 0
        assert product.sparknet_get_history(b.state, i, fptr(hb), x.size) == 0
        assert np.array_equal(ha, hb), i
    a.close(); b.close()
    assert os.path.getsize(model) > 431080 * 4
