"""GPU parity at net / solver level through the C ABI (product vs CPU oracle), plus the boundary
contracts of SURVEY.md §8(b): get_data semantics, callback pairing, weight averaging."""
import ctypes

import numpy as np
import pytest

from sparknet_b200 import zoo
from sparknet_b200.net import CaffeNet, MinibatchSource, WeightCollection
from tests.harness import assert_mostly_close, fptr, rel_err, synthetic_batches

pytestmark = pytest.mark.gpu
F = np.float32


def make_pair(oracle, product, name, batch, n_batches, precision=0, **over):
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0, **over)
    o = CaffeNet.from_solver(sol, oracle)
    p = CaffeNet.from_solver(sol, product)
    product.sparknet_set_precision(p.state, precision)
    imgs, labs = synthetic_batches(name, n_batches, batch)
    o.setTrainData(MinibatchSource(imgs, labs))
    p.setTrainData(MinibatchSource(imgs, labs))
    p.setWeights(o.getWeights())                       # fillers are unpinned: copy the oracle's draw
    return o, p, imgs, labs


def inject_same_masks(oracle, product, o, p, seed=0):
    rng = np.random.default_rng(seed)
    for idx, name in enumerate(o.layerNames):
        if name.startswith("drop"):
            blob = [b for b in o.getData().items() if b[0] == {"drop6": "fc6", "drop7": "fc7"}[name]][0][1]
            m = (rng.random(blob.size) > 0.5).astype(np.uint32)
            mp = m.ctypes.data_as(ctypes.POINTER(ctypes.c_uint))
            assert oracle.sparknet_inject_dropout_mask(o.state, idx, mp, m.size) == 0
            assert product.sparknet_inject_dropout_mask(p.state, idx, mp, m.size) == 0


@pytest.mark.parametrize("name,batch", [("lenet", 16), ("cifar10_quick", 10), ("cifar10_full", 10)])
def test_forward_backward_every_blob(oracle, product, name, batch):
    o, p, *_ = make_pair(oracle, product, name, batch, 1)
    o.forward(); p.forward()
    do, dp = o.getData(), p.getData()
    assert list(do) == list(dp)
    for k in do:
        assert do[k].shape == dp[k].shape
        assert rel_err(dp[k], do[k]) <= 1e-5, (k, rel_err(dp[k], do[k]))
    o.backward(); p.backward()
    go, gp = o.getData(diff=True), p.getData(diff=True)
    for k in go:
        if k in ("data", "label", "loss"):
            continue
        assert_mostly_close(gp[k], go[k], 1e-5, k)
    wo, wp = o.getWeightDiffs(), p.getWeightDiffs()
    for lname in wo:
        for j, (a, b) in enumerate(zip(wo[lname], wp[lname])):
            assert_mostly_close(b, a, 2e-5, f"{lname}[{j}]")
    o.close(); p.close()


def test_caffenet_forward_backward_small_batch(oracle, product):
    o, p, *_ = make_pair(oracle, product, "caffenet", 2, 1)
    inject_same_masks(oracle, product, o, p)
    o.forward(); p.forward()
    do, dp = o.getData(), p.getData()
    assert list(do) == list(dp)
    for k in do:
        assert rel_err(dp[k], do[k]) <= 1e-5, (k, rel_err(dp[k], do[k]))
    o.backward(); p.backward()
    wo, wp = o.getWeightDiffs(), p.getWeightDiffs()
    for lname in wo:
        for j, (a, b) in enumerate(zip(wo[lname], wp[lname])):
            assert_mostly_close(b, a, 5e-5, f"{lname}[{j}]")
    o.close(); p.close()


@pytest.mark.parametrize("name,batch,iters", [("lenet", 16, 5), ("cifar10_quick", 10, 4), ("cifar10_full", 10, 4)])
def test_solver_steps_match_weights_and_history(oracle, product, name, batch, iters):
    o, p, *_ = make_pair(oracle, product, name, batch, 3)
    o.train(iters); p.train(iters)
    assert product.sparknet_get_iter(p.state) == iters == oracle.sparknet_get_iter(o.state)
    wo, wp = o.getWeights(), p.getWeights()
    for lname in wo.layerNames:
        for j, (a, b) in enumerate(zip(wo.allWeights[lname], wp.allWeights[lname])):
            assert rel_err(b, a) <= 1e-5 * iters, (lname, j, rel_err(b, a))
    k = 0
    for lname in wo.layerNames:
        for a in wo.allWeights[lname]:
            ho = np.ctypeslib.as_array(oracle.oracle_history(o.state, k), shape=(a.size,)).copy()
            hp = np.empty(a.size, dtype=F)
            assert product.sparknet_get_history(p.state, k, fptr(hp), hp.size) == 0
            assert rel_err(hp, ho) <= 2e-5 * iters, (lname, k)
            k += 1
    o.close(); p.close()


def test_lenet_loss_trajectory_200_iterations_within_1_percent(oracle, product):
    o, p, *_ = make_pair(oracle, product, "lenet", 64, 8)
    lo = []
    for _ in range(200):
        o.train(1)
        lo.append(oracle.sparknet_last_loss(o.state))
    p.train(200)                                               # one call, no per-iteration sync
    lp = np.empty(200, dtype=F)
    assert product.sparknet_get_loss_history(p.state, fptr(lp), 200) == 200
    lo = np.array(lo)
    assert lo[-1] < 0.5 * lo[0]                                # it actually trains
    assert np.max(np.abs(lp - lo) / np.maximum(np.abs(lo), 1e-3)) < 0.01
    o.close(); p.close()


def test_callbacks_are_paired_and_counted_exactly(product):
    sol = zoo.solver_with_net("lenet", train_bs=8, random_seed=1, display=0)
    p = CaffeNet.from_solver(sol, product)
    imgs, labs = synthetic_batches("lenet", 3, 8)
    src = MinibatchSource(imgs, labs)
    order = []
    im, lb = src.next_image_minibatch, src.next_label_minibatch
    src.next_image_minibatch = lambda: (order.append("d"), im())[1]
    src.next_label_minibatch = lambda: (order.append("l"), lb())[1]
    p.setTrainData(src)
    p.train(7)
    product.sparknet_sync(p.state)
    assert order == ["d", "l"] * 7                              # MinibatchSampler contract
    assert product.set_train_data_callback(p.state, 2, p._callbacks["img_train"]) == -2
    assert product.set_train_data_callback(p.state, 99, p._callbacks["img_train"]) == -1
    p.close()


def test_get_data_pointer_is_authoritative_after_the_call(oracle, product):
    o, p, *_ = make_pair(oracle, product, "lenet", 8, 2)
    blob = product.get_weight_blob(p.state, 2, 0)
    ptr1 = ctypes.addressof(product.get_data(blob).contents)
    p.train(1)
    ptr2 = ctypes.addressof(product.get_data(blob).contents)     # shows the trained values, same address
    assert ptr1 == ptr2
    w = p.getWeights()
    # write through the pointer (what Net.scala:141-147 does), then train: the new values must be used
    for lname in w.layerNames:
        for a in w.allWeights[lname]:
            a *= 0
    p.setWeights(w)
    p.train(1)
    z = p.getWeights()
    assert all(np.all(np.abs(a) < 1e-2) for ws in z.allWeights.values() for a in ws)
    o.close(); p.close()


def test_resident_inputs_match_callback_mode(product):
    sol = zoo.solver_with_net("cifar10_quick", train_bs=10, random_seed=3, display=0)
    imgs, labs = synthetic_batches("cifar10_quick", 3, 10)
    a = CaffeNet.from_solver(sol, product)
    b = CaffeNet.from_solver(sol, product)
    a.setTrainData(MinibatchSource(imgs, labs)); b.setTrainData(MinibatchSource(imgs, labs))
    b.setWeights(a.getWeights())
    assert product.sparknet_set_resident_inputs(b.state, 3) == 0
    a.train(6); b.train(6)
    wa, wb = a.getWeights(), b.getWeights()
    for lname in wa.layerNames:
        for x, y in zip(wa.allWeights[lname], wb.allWeights[lname]):
            assert np.array_equal(x, y)
    a.close(); b.close()


@pytest.mark.parametrize("nrep", [1, 2, 3, 4, 8])
def test_averaging_kernel_is_bit_exact_rank_ordered(oracle, product, nrep):
    import torch
    rng = np.random.default_rng(nrep)
    n = 1_000_003
    reps = [(rng.standard_normal(n) * 10.0 ** rng.integers(-30, 30, n)).astype(F) for _ in range(nrep)]
    reps[0][:7] = [0.0, -0.0, 1e-45, -1e-45, 3.4e38, 1.0, 1e-38]     # zeros, denormals, near-overflow
    ptrs = (ctypes.POINTER(ctypes.c_float) * nrep)(*[fptr(a) for a in reps])
    want = np.empty(n, dtype=F)
    oracle.oracle_average(ptrs, nrep, n, fptr(want))
    dev = [torch.from_numpy(a).cuda() for a in reps]
    out = torch.empty(n, dtype=torch.float32, device="cuda")
    dptrs = (ctypes.c_void_p * nrep)(*[d.data_ptr() for d in dev])
    torch.cuda.synchronize()
    assert product.sparknet_average_replicas_device(dptrs, nrep, n, ctypes.c_void_p(out.data_ptr())) == 0
    got = out.cpu().numpy()
    assert got.tobytes() == want.tobytes()                           # memcmp, incl. -0.0 and denormals


def test_two_replicas_two_rounds_momentum_is_not_averaged(oracle, product):
    """The SparkNet round loop (src/main/scala/apps/CifarApp.scala:95-136) on two replicas: local
    steps, average weights only, continue with each replica's own momentum history."""
    tau, rounds, name, batch = 3, 2, "lenet", 8
    sol = zoo.solver_with_net(name, train_bs=batch, random_seed=5, display=0)
    def replicas(lib):
        nets = []
        for r in range(2):
            n = CaffeNet.from_solver(sol, lib)
            imgs, labs = synthetic_batches(name, tau * rounds, batch, seed=1701 + r)
            n.setTrainData(MinibatchSource(imgs, labs))
            nets.append(n)
        return nets
    O, P = replicas(oracle), replicas(product)
    w0 = O[0].getWeights()
    for n in O + P:
        n.setWeights(w0)
    for _ in range(rounds):
        for nets in (O, P):
            for n in nets:
                n.train(tau)
            tot = WeightCollection.add(nets[0].getWeights(), nets[1].getWeights())
            tot.scalarDivide(2.0)
            for n in nets:
                n.setWeights(tot)
    wo, wp = O[0].getWeights(), P[0].getWeights()
    for lname in wo.layerNames:
        for a, b in zip(wo.allWeights[lname], wp.allWeights[lname]):
            assert rel_err(b, a) <= 1e-4
    for n in O + P:
        n.close()
