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
    # ReLU gates (y > 0) can legitimately differ where a pre-activation is within summation-order noise
    # of zero; one flipped gate changes the gradient of a whole neighbourhood upstream.  Strict 1e-5
    # when all gates agree (layer-level tests are always strict), wiring-level 5e-2 otherwise.
    gates_agree = all(np.array_equal(do[k] > 0, dp[k] > 0) for k in do)
    tol = 1e-5 if gates_agree else 5e-2
    o.backward(); p.backward()
    go, gp = o.getData(diff=True), p.getData(diff=True)
    for k in go:
        if k in ("data", "label", "loss"):
            continue
        assert_mostly_close(gp[k], go[k], tol, k, max_bad_frac=5e-4 if gates_agree else 0.2)
    wo, wp = o.getWeightDiffs(), p.getWeightDiffs()
    for lname in wo:
        for j, (a, b) in enumerate(zip(wo[lname], wp[lname])):
            assert_mostly_close(b, a, 2 * tol, f"{lname}[{j}]", max_bad_frac=5e-4 if gates_agree else 0.2)
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


def oracle_self_sensitivity(oracle, name, batch, iters, n_batches, eps=1e-6):
    """How far the ORACLE moves from itself when its initial weights are perturbed by `eps` relative
    (about the size of fp32 summation-order noise): per-blob rel err of weights after `iters` steps.
    Some configs are intrinsically unstable (cifar10_quick: 1e-4-std conv1 filters on +-128 inputs,
    ReLU/max-pool gate flips) — the oracle differs from itself by 5e-2 there — so multi-iteration
    parity is judged against max(strict tolerance, 3 x this sensitivity)."""
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0)
    out = []
    for perturb in (0.0, eps):
        n = CaffeNet.from_solver(sol, oracle)
        imgs, labs = synthetic_batches(name, n_batches, batch)
        n.setTrainData(MinibatchSource(imgs, labs))
        if perturb:
            w = n.getWeights()
            rng = np.random.default_rng(0)
            for ln in w.layerNames:
                for a in w.allWeights[ln]:
                    a *= (1 + perturb * rng.standard_normal(a.shape)).astype(F)
            n.setWeights(w)
        n.train(iters)
        w = n.getWeights()
        hist, k = {}, 0
        for ln in w.layerNames:
            for j, a in enumerate(w.allWeights[ln]):
                hist[(ln, j)] = np.ctypeslib.as_array(oracle.oracle_history(n.state, k), shape=(a.size,)).copy()
                k += 1
        out.append((w, hist))
        n.close()
    (a, ha), (b, hb) = out
    sens_w = {(ln, j): rel_err(y, x) for ln in a.layerNames for j, (x, y) in enumerate(zip(a.allWeights[ln], b.allWeights[ln]))}
    sens_h = {k: rel_err(hb[k], ha[k]) for k in ha}
    return sens_w, sens_h


@pytest.mark.parametrize("name,batch,iters", [("lenet", 16, 5), ("cifar10_quick", 10, 4), ("cifar10_full", 10, 4)])
def test_solver_steps_match_weights_and_history(oracle, product, name, batch, iters):
    o, p, *_ = make_pair(oracle, product, name, batch, 3)
    sens, sens_h = oracle_self_sensitivity(oracle, name, batch, iters, 3)
    o.train(iters); p.train(iters)
    assert product.sparknet_get_iter(p.state) == iters == oracle.sparknet_get_iter(o.state)
    wo, wp = o.getWeights(), p.getWeights()
    for lname in wo.layerNames:
        for j, (a, b) in enumerate(zip(wo.allWeights[lname], wp.allWeights[lname])):
            tol = max(1e-5 * iters, 3 * sens[(lname, j)])
            assert rel_err(b, a) <= tol, (lname, j, rel_err(b, a), tol)
    k = 0
    for lname in wo.layerNames:
        for j, a in enumerate(wo.allWeights[lname]):
            ho = np.ctypeslib.as_array(oracle.oracle_history(o.state, k), shape=(a.size,)).copy()
            hp = np.empty(a.size, dtype=F)
            assert product.sparknet_get_history(p.state, k, fptr(hp), hp.size) == 0
            tol = max(2e-5 * iters, 3 * sens_h[(lname, j)])
            assert rel_err(hp, ho) <= tol, (lname, k, rel_err(hp, ho), tol)
            k += 1
    o.close(); p.close()


def learnable_batches(n_batches, batch, seed=1701):
    """LeNet-shaped synthetic task that can actually be learned: ten fixed prototypes plus noise."""
    rng = np.random.default_rng(seed)
    proto = rng.random((10, 1, 28, 28), dtype=F)
    imgs, labs = [], []
    for _ in range(n_batches):
        lab = rng.integers(0, 10, size=batch)
        imgs.append(np.ascontiguousarray(0.6 * proto[lab] + 0.4 * rng.random((batch, 1, 28, 28), dtype=F), dtype=F))
        labs.append(lab.astype(F))
    return imgs, labs


def test_lenet_loss_trajectory_200_iterations_within_1_percent(oracle, product):
    """north_star: 'the training-loss trajectory over 200 iterations must stay within 1%'.
    SGD amplifies fp32 summation-order noise: the oracle run twice with a 1e-6 relative weight
    perturbation already differs from itself by up to a few percent of the *current* loss once the
    loss has collapsed (measured: DESIGN.md).  So: (a) every iteration within 1% of the loss scale
    (initial loss), and (b) within max(1%, 3x the oracle's own sensitivity) of the current loss."""
    sol = zoo.solver_with_net("lenet", train_bs=64, test_bs=64, random_seed=1701, display=0)
    imgs, labs = learnable_batches(16, 64)
    runs = {}
    for key, lib, perturb in (("oracle", oracle, 0.0), ("oracle_eps", oracle, 1e-6), ("product", product, 0.0)):
        n = CaffeNet.from_solver(sol, lib)
        n.setTrainData(MinibatchSource(imgs, labs))
        if key == "oracle":
            w0 = n.getWeights()
        else:
            w = WeightCollection({k: [a.copy() for a in v] for k, v in w0.allWeights.items()}, w0.layerNames)
            if perturb:
                rng = np.random.default_rng(0)
                for ln in w.layerNames:
                    for a in w.allWeights[ln]:
                        a *= (1 + perturb * rng.standard_normal(a.shape)).astype(F)
            n.setWeights(w)
        if lib is product:
            product.sparknet_set_precision(n.state, 0)
            n.train(200)                                           # one call, no per-iteration sync
            l = np.empty(200, dtype=F)
            assert product.sparknet_get_loss_history(n.state, fptr(l), 200) == 200
        else:
            l = []
            for _ in range(200):
                n.train(1)
                l.append(lib.sparknet_last_loss(n.state))
            l = np.array(l, dtype=F)
        runs[key] = l.astype(np.float64)
        n.close()
    lo, le, lp = runs["oracle"], runs["oracle_eps"], runs["product"]
    assert lo[-1] < 0.05 * lo[0]                                   # it actually trains
    assert np.max(np.abs(lp - lo)) / lo[0] < 0.01                  # (a)
    rel = lambda x: np.max(np.abs(x - lo) / np.maximum(np.abs(lo), 1e-3))
    assert rel(lp) < max(0.01, 3 * rel(le)), (rel(lp), rel(le))    # (b)


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


@pytest.mark.parametrize("name,batch,precision", [("lenet", 16, 0), ("cifar10_quick", 10, 1), ("cifar10_full", 10, 1),
                                                  ("caffenet", 2, 1)])
def test_graph_replay_is_bit_identical_to_eager_iterations(product, name, batch, precision):
    """solver_step replays a recorded CUDA graph for the steady-state iterations (no per-iteration host
    work beyond the data callbacks).  Same kernels, same order, same arguments: weights, momentum history and the loss
    of every iteration must be bit-identical to eager launches — across several solver_step calls, with host writes to
    the weights in between (setWeights through get_data pointers), an lr policy that changes every iteration (LeNet
    'inv') and dropout counters (CaffeNet)."""
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0)
    imgs, labs = synthetic_batches(name, 4, batch)
    runs = []
    for mode in (0, 1):
        p = CaffeNet.from_solver(sol, product)
        product.sparknet_set_precision(p.state, precision)
        assert product.sparknet_set_graph_mode(p.state, mode) == -1
        p.setTrainData(MinibatchSource(imgs, labs))
        if mode == 0:
            w0 = p.getWeights()
        p.setWeights(w0)
        p.train(5)
        w = p.getWeights()                                   # host reads ...
        for ln in w.layerNames:
            for a in w.allWeights[ln]:
                a *= F(0.999)
        p.setWeights(w)                                      # ... and writes between two calls
        p.train(6)
        p.train(1)
        p.train(4)
        n = int(product.sparknet_num_params(p.state))
        flat = np.empty(n, dtype=F)
        assert product.sparknet_get_weights_flat(p.state, fptr(flat), n) == 0
        losses = np.empty(16, dtype=F)
        assert product.sparknet_get_loss_history(p.state, fptr(losses), 16) == 16
        hist, k, w = [], 0, p.getWeights()
        for ln in w.layerNames:
            for a in w.allWeights[ln]:
                h = np.empty(a.size, dtype=F)
                assert product.sparknet_get_history(p.state, k, fptr(h), a.size) == 0
                hist.append(h)
                k += 1
        replays = int(product.sparknet_graph_replays(p.state))
        runs.append((flat, losses, hist, replays))
        p.close()
    (w_e, l_e, h_e, r_e), (w_g, l_g, h_g, r_g) = runs
    assert r_e == 0
    # first call: two eager iterations before the first recording; second call: its first iteration is eager because
    # the host touched the weights in between; the last two calls found nothing touched and replay from iteration one
    assert r_g == (5 - 2) + (6 - 1) + 1 + 4
    assert np.isfinite(l_g).all()
    assert l_g.tobytes() == l_e.tobytes()
    assert w_g.tobytes() == w_e.tobytes()
    assert len(h_e) == len(h_g) and all(a.tobytes() == b.tobytes() for a, b in zip(h_e, h_g))


@pytest.mark.parametrize("nrep", [1, 2, 3, 5, 8])
def test_peer_memory_exchange_kernel_is_bit_exact_rank_ordered(oracle, product, nrep):
    """The one-kernel exchange of the multi-GPU driver (kernels.cu average_peers_kernel: rank-ordered fold of the own
    slice from every replica, result stored into every replica, ready/done flags) run with nrep replicas on ONE GPU;
    two rounds, every replica's arena against the oracle's sequential fold (NDArraySpec.scala:34-56 arithmetic)."""
    rng = np.random.default_rng(100 + nrep)
    n = 1024 * 37 + (3 if nrep % 2 else 0)                              # ragged tail when odd
    reps = np.stack([(rng.standard_normal(n) * 10.0 ** rng.integers(-30, 30, n)).astype(F) for _ in range(nrep)])
    reps[0][:7] = [0.0, -0.0, 1e-45, -1e-45, 3.4e38, 1.0, 1e-38]
    want = reps.copy()
    for _ in range(2):
        ptrs = (ctypes.POINTER(ctypes.c_float) * nrep)(*[fptr(want[r]) for r in range(nrep)])
        avg = np.empty(n, dtype=F)
        oracle.oracle_average(ptrs, nrep, n, fptr(avg))
        want = np.stack([avg] * nrep)
    got = np.empty_like(reps)
    assert product.sparknet_test_average_peers(nrep, n, fptr(reps), fptr(got), 2) == 0
    for r in range(nrep):
        assert got[r].tobytes() == want[r].tobytes(), r


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
