"""The paths either side of the training loop that the unchanged Scala driver uses, on the device, against the CPU oracle:
solver_test with train<->test weight sharing, the cheap getWeights / setWeights path behind the
get_weight_blob handles (F2), the byte-pixel data callback, and recorded iterations surviving a workspace reallocation."""
import ctypes

import numpy as np
import pytest

from sparknet_b200 import zoo
from sparknet_b200.net import CaffeNet, MinibatchSource
from tests.harness import fptr, rel_err, synthetic_batches

pytestmark = pytest.mark.gpu
F = np.float32


def copy_counts(product, state):
    d, h = ctypes.c_long(0), ctypes.c_long(0)
    product.sparknet_param_copy_counts(state, ctypes.byref(d), ctypes.byref(h))
    return d.value, h.value


@pytest.mark.parametrize("name,batch,base_lr", [("lenet", 64, 0.01), ("cifar10_quick", 100, 0.01)])
def test_solver_test_after_solver_step_matches_the_oracle(oracle, product, name, batch, base_lr):
    """The driver's test round (CifarApp.scala:101-116 -> CaffeNet.test -> solver_test, libccaffe/ccaffe.cpp:235-243 ->
    Solver::TestAndStoreResult, solver.cpp:413-444): scores summed over the test batches, in output-blob order
    (accuracy, loss).  The TEST net must see the TRAIN net's weights as solver_step left them
    (Net::ShareTrainedLayersWith, net.cpp:737-766) and as setWeights wrote them."""
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0, base_lr=base_lr)
    o, p = CaffeNet.from_solver(sol, oracle), CaffeNet.from_solver(sol, product)
    assert product.sparknet_set_precision(p.state, 0) == 0
    imgs, labs = synthetic_batches(name, 4, batch)
    timgs, tlabs = synthetic_batches(name, 3, batch, seed=99)
    for n in (o, p):
        n.setTrainData(MinibatchSource(imgs, labs))
        n.setTestData(MinibatchSource(timgs, tlabs), 3)
    p.setWeights(o.getWeights())
    s0o, s0p = o.test(), p.test()
    assert s0o.shape == s0p.shape == (2,)
    assert s0o[0] == s0p[0]                                          # accuracy: summed 0/1 hits, exact
    assert abs(s0o[1] - s0p[1]) <= 1e-5 * abs(s0o[1])
    o.train(5); p.train(5)
    s1o, s1p = o.test(), p.test()                                    # weights updated on the device, never touched by the host
    assert abs(s1o[1] - s0o[1]) > 1e-5 * abs(s0o[1])                 # training moved the test loss ...
    # ... and the product's TEST net saw it move too: same direction, comparable amount (five SGD steps on this net
    # amplify summation-order noise by tens of percent; a TEST net that did not share the weights would
    # not move at all)
    assert 0.5 <= (s1p[1] - s0p[1]) / (s1o[1] - s0o[1]) <= 2.0
    assert abs(s1o[0] - s1p[0]) <= 5.0 / batch + 1e-6               # a few near-tie samples of 300 may flip
    # the oracle's trained weights written through the get_data pointers: strict again
    p.setWeights(o.getWeights())
    s2p = p.test()
    assert s2p[0] == s1o[0]
    assert abs(s2p[1] - s1o[1]) <= 1e-5 * abs(s1o[1])
    assert product.num_test_scores(p.state) == 2 == oracle.num_test_scores(o.state)
    o.close(); p.close()


def test_weights_cross_the_bus_once_each_way(product):
    """getWeights / setWeights as Net.scala:132-172 does them — one get_weight_blob + get_data per blob, then element
    reads / writes through the pointer — on CaffeNet's 16 blobs (61 M parameters): after device-side updates the first
    get_data moves the whole arena in ONE device-to-host copy and the other 15 copy nothing; after the host wrote all
    blobs the next device use is ONE host-to-device copy.  Values: exact round trip."""
    sol = zoo.solver_with_net("caffenet", train_bs=2, test_bs=2, random_seed=1701, display=0)
    p = CaffeNet.from_solver(sol, product)
    imgs, labs = synthetic_batches("caffenet", 2, 2)
    p.setTrainData(MinibatchSource(imgs, labs))
    p.train(1)                                                       # weights are now newest on the device
    n = int(product.sparknet_num_params(p.state))
    flat = np.empty(n, dtype=F)
    assert product.sparknet_get_weights_flat(p.state, fptr(flat), n) == 0   # device read that does not involve the mirror
    d0, h0 = copy_counts(product, p.state)
    w = p.getWeights()                                               # 16 x (get_weight_blob, get_data)
    d1, h1 = copy_counts(product, p.state)
    assert (d1 - d0, h1 - h0) == (1, 0)
    assert sum(len(v) for v in w.allWeights.values()) == 16
    assert np.array_equal(np.concatenate([a.ravel() for ln in w.layerNames for a in w.allWeights[ln]]), flat)
    # element access through the raw pointer, as the JVM does it (getFloat / setFloat at byte offsets)
    blob = product.get_weight_blob(p.state, p.layerNames.index("fc6"), 0)
    ptr = product.get_data(blob)
    fc6 = w.allWeights["fc6"][0].ravel()
    for i in (0, 1, 4095, 9216 * 17 + 5, fc6.size - 1):
        assert ptr[i] == fc6[i]
    for ln in w.layerNames:
        for a in w.allWeights[ln]:
            a *= F(0.5)
    p.setWeights(w)                                                  # 16 x get_data + element writes
    d2, h2 = copy_counts(product, p.state)
    assert (d2 - d1, h2 - h1) == (0, 0)                              # the host copies were already authoritative
    back = p.getWeights()
    assert all(np.array_equal(a, b) for ln in w.layerNames for a, b in zip(w.allWeights[ln], back.allWeights[ln]))
    assert copy_counts(product, p.state) == (d2, h2)
    assert product.sparknet_get_weights_flat(p.state, fptr(flat), n) == 0   # first device use of the host-written weights
    d3, h3 = copy_counts(product, p.state)
    assert (d3 - d2, h3 - h2) == (0, 1)
    assert np.array_equal(np.concatenate([a.ravel() for ln in w.layerNames for a in w.allWeights[ln]]), flat)
    p.train(1)
    assert copy_counts(product, p.state) == (d3, h3)                 # nothing left to move
    # a single blob written by the host between two steps: one copy down (all blobs were newer on the device), one copy up
    # (only that blob's range)
    v = p._view(blob)
    d4, h4 = copy_counts(product, p.state)
    assert (d4 - d3, h4 - h3) == (1, 0)
    v[...] = 0
    p.train(1)
    d5, h5 = copy_counts(product, p.state)
    assert (d5 - d4, h5 - h4) == (0, 1)
    p.close()


BYTE_CASES = [
    # net, batch, signed, mean kind, scale
    ("cifar10_quick", 100, True, "image", 1.0),         # raw Java bytes minus a mean image (ImageNetApp.scala:158-175)
    ("lenet", 64, False, "channel", 1.0 / 256.0),       # 0..255 pixels, per-channel mean, scaled (lenet_train_test.prototxt:10-12)
    ("caffenet", 3, True, None, 1.0),                   # odd element count: the unaligned tail / resident-slot paths
]


@pytest.mark.parametrize("resident", [False, True])
@pytest.mark.parametrize("name,batch,signed,mean_kind,scale", BYTE_CASES)
def test_byte_pixel_callback_is_bit_identical_to_the_float_callback(product, name, batch, signed, mean_kind, scale, resident):
    """sparknet_set_train_data_callback_bytes: the callback hands over bytes and the device computes
    ((float)b - mean) * scale — the two fp32 operations the reference's callback does on the host (ByteImage.java:86-104,
    ImageNetApp.scala:158-175).  Against the fp32 callback fed those host-computed floats: same data blob, and after four
    solver iterations (eager and replayed) the same weights and losses, bit for bit."""
    c, h, w, classes, _, _ = zoo.GEOMETRY[name]
    rng = np.random.default_rng(5)
    dt = np.int8 if signed else np.uint8
    lo, hi = (-128, 128) if signed else (0, 256)
    raw = [rng.integers(lo, hi, size=(batch, c, h, w)).astype(dt) for _ in range(3)]
    labs = [rng.integers(0, classes, size=(batch,)).astype(F) for _ in range(3)]
    mean = None
    floats = [r.astype(F) for r in raw]
    if mean_kind == "image":
        mean = (rng.random((c, h, w), dtype=F) * F(20.0)).astype(F)
        floats = [f - mean[None] for f in floats]
    elif mean_kind == "channel":
        mean = (rng.random((c,), dtype=F) * F(100.0)).astype(F)
        floats = [f - mean[None, :, None, None] for f in floats]
    if scale != 1.0:
        floats = [f * F(scale) for f in floats]
    floats = [np.ascontiguousarray(f, dtype=F) for f in floats]
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0)
    a, b = CaffeNet.from_solver(sol, product), CaffeNet.from_solver(sol, product)
    for n in (a, b):
        product.sparknet_set_precision(n.state, 1)
    a.setTrainData(MinibatchSource(floats, labs))
    b.setTrainDataBytes(MinibatchSource(raw, labs), signed=signed, mean=mean, scale=scale)
    b.setWeights(a.getWeights())
    if resident:
        assert product.sparknet_set_resident_inputs(a.state, 3) == 0
        assert product.sparknet_set_resident_inputs(b.state, 3) == 0
    a.forward(); b.forward()
    da, db = a.getData()["data"], b.getData()["data"]
    assert np.array_equal(da, floats[0]) and np.array_equal(db, floats[0])
    a.train(4); b.train(4)
    if not resident:                                                 # the blob holds the last minibatch of the call on both paths
        assert np.array_equal(a.getData()["data"], floats[1]) and np.array_equal(b.getData()["data"], floats[1])
    la, lb = np.empty(4, dtype=F), np.empty(4, dtype=F)
    assert product.sparknet_get_loss_history(a.state, fptr(la), 4) == 4 == product.sparknet_get_loss_history(b.state, fptr(lb), 4)
    assert np.isfinite(la).all() and la.tobytes() == lb.tobytes()
    n = int(product.sparknet_num_params(a.state))
    fa, fb = np.empty(n, dtype=F), np.empty(n, dtype=F)
    assert product.sparknet_get_weights_flat(a.state, fptr(fa), n) == 0 == product.sparknet_get_weights_flat(b.state, fptr(fb), n)
    assert fa.tobytes() == fb.tobytes()
    if not resident:
        st = (ctypes.c_double * 3)()
        assert product.sparknet_input_stats(b.state, st) == 0
        assert st[0] >= 8                                            # (data, label) x 4 iterations at least
    # switching the layer back to a float callback works
    b.setTrainData(MinibatchSource(floats, labs))
    if resident:
        assert product.sparknet_set_resident_inputs(b.state, 0) == 0
    b.forward()
    assert np.array_equal(b.getData()["data"], floats[0])
    a.close(); b.close()


def test_recorded_iterations_survive_a_workspace_reallocation(product):
    """Recorded iterations (CUDA graphs) hold the scratch workspace's address in their split-reduction and bias-partial
    launches.  Anything that grows the workspace later frees that buffer; the recordings must then be dropped and made
    again, not replayed into freed memory.  Same sequence with replay on and off: bit-identical weights and losses."""
    name, batch = "cifar10_quick", 100
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0)
    imgs, labs = synthetic_batches(name, 4, batch)
    runs = []
    for mode in (0, 1):
        p = CaffeNet.from_solver(sol, product)
        product.sparknet_set_precision(p.state, 1)
        product.sparknet_set_graph_mode(p.state, mode)
        p.setTrainData(MinibatchSource(imgs, labs))
        if mode == 0:
            w0 = p.getWeights()
        p.setWeights(w0)
        p.train(6)
        r0 = int(product.sparknet_graph_replays(p.state))
        have = int(product.sparknet_reserve_workspace(0))
        assert int(product.sparknet_reserve_workspace(2 * have + (1 << 20))) >= 2 * have + (1 << 20)   # reallocates
        p.train(6)
        r1 = int(product.sparknet_graph_replays(p.state))
        n = int(product.sparknet_num_params(p.state))
        flat = np.empty(n, dtype=F)
        assert product.sparknet_get_weights_flat(p.state, fptr(flat), n) == 0
        losses = np.empty(12, dtype=F)
        assert product.sparknet_get_loss_history(p.state, fptr(losses), 12) == 12
        runs.append((flat, losses, r0, r1))
        p.close()
    (w_e, l_e, e0, e1), (w_g, l_g, g0, g1) = runs
    assert e0 == e1 == 0
    assert g0 == 4                                                   # two eager iterations, then replays
    assert g1 - g0 == 4                                              # recorded again after the reallocation (two eager iterations first)
    assert np.isfinite(l_g).all() and l_g.tobytes() == l_e.tobytes()
    assert w_g.tobytes() == w_e.tobytes()
