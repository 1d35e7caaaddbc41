"""Net- and solver-level pins of the CPU oracle (CPU only).

Reference anchors: blob naming / shapes at the ABI (src/test/scala/libs/CifarFeaturizationSpec.scala:83-103),
chance accuracy of a random-init net (src/test/scala/libs/CifarSpec.scala:81-92), callback pairing
(src/test/scala/libs/MinibatchSamplerSpec.scala:11-43), the SGD update law
(caffe/src/caffe/test/test_gradient_based_solver.cpp:225-399), learning-rate policies
(caffe/src/caffe/solvers/sgd_solver.cpp:27-63), test-score ordering (caffe/src/caffe/net.cpp:266-274)."""
import numpy as np
import pytest

from sparknet_b200 import zoo
from sparknet_b200.net import CaffeNet, MinibatchSource
from tests.harness import synthetic_batches

F = np.float32


def make(oracle, name, batch, n_batches=4, **over):
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=batch, random_seed=1701, display=0, **over)
    net = CaffeNet.from_solver(sol, oracle)
    imgs, labs = synthetic_batches(name, n_batches, batch)
    src = MinibatchSource(imgs, labs)
    net.setTrainData(src)
    return net, src, imgs, labs


def test_cifar_blob_names_and_shapes_at_the_abi(oracle):
    net, *_ = make(oracle, "cifar10_full", 100, 1)
    net.forward()
    data = net.getData()
    assert sorted(data) == sorted(["data", "label", "conv1", "pool1", "norm1", "conv2", "pool2", "norm2", "conv3",
                                   "pool3", "ip1", "loss"])
    assert data["conv1"].shape == (100, 32, 32, 32)        # CifarFeaturizationSpec.scala:101
    assert data["pool3"].shape == (100, 64, 4, 4) and data["ip1"].shape == (100, 10) and data["loss"].shape == ()
    assert net.layerNames[:3] == ["data", "label", "conv1"] and net.layerNames[-1] == "loss"
    assert net.layerNumBlobs[:3] == [0, 0, 2]
    w = net.getWeights()
    assert w.allWeights["conv1"][0].shape == (32, 3, 5, 5) and w.allWeights["ip1"][0].shape == (10, 1024)
    assert sum(a.size for ws in w.allWeights.values() for a in ws) == 89578
    net.close()


@pytest.mark.parametrize("name,params", [("lenet", 431080), ("cifar10_quick", 145578)])
def test_param_counts(oracle, name, params):
    net, *_ = make(oracle, name, 4, 1)
    assert sum(a.size for ws in net.getWeights().allWeights.values() for a in ws) == params
    net.close()


def test_callbacks_come_in_data_label_pairs_exactly_step_times(oracle):
    net, src, *_ = make(oracle, "lenet", 8, 3)
    order = []
    src_img, src_lab = src.next_image_minibatch, src.next_label_minibatch
    src.next_image_minibatch = lambda: (order.append("d"), src_img())[1]
    src.next_label_minibatch = lambda: (order.append("l"), src_lab())[1]
    net.train(5)
    assert order == ["d", "l"] * 5
    assert oracle.sparknet_get_iter(net.state) == 5
    net.close()


def test_sgd_update_law_with_momentum_and_decay(oracle):
    """w' = w - (lr*lr_mult*(g + wd*decay_mult*w) + mu*h), h' = that update (sgd_solver.cpp:145-239)."""
    net, src, imgs, labs = make(oracle, "lenet", 8, 2)
    names = net.layerNames
    w0 = net.getWeights()
    # gradient at w0 on batch 0, measured through forward/backward on a twin net with the same weights
    twin, tsrc, *_ = make(oracle, "lenet", 8, 2)
    twin.setWeights(w0)
    twin.forward(); twin.backward()
    g0 = twin.getWeightDiffs()
    net.train(1)
    w1 = net.getWeights()
    lr, mu, wd = F(0.01), F(0.9), F(0.0005)               # iter 0: inv policy => base_lr * 1^-0.75
    h = {}
    for n in names:
        for j, w in enumerate(w0.allWeights[n]):
            lr_mult = F(1.0 if j == 0 else 2.0)
            g = g0[n][j] + (wd * F(1.0)) * w
            upd = (lr * lr_mult) * g
            h[(n, j)] = upd
            assert np.allclose(w1.allWeights[n][j], w - upd, rtol=1e-5, atol=1e-8), (n, j)
    # second step: history participates
    twin.setWeights(w1)
    for n in names:                                        # raw backward() accumulates: clear first
        for j in range(len(g0[n])):
            twin._view(twin.lib.get_weight_blob(twin.state, names.index(n), j), diff=True)[...] = 0
    twin.forward(); twin.backward()
    g1 = twin.getWeightDiffs()
    net.train(1)
    w2 = net.getWeights()
    lr2 = F(0.01) * F(np.power(F(1.0) + F(0.0001) * F(1), F(-0.75)))
    for n in names:
        for j, w in enumerate(w1.allWeights[n]):
            lr_mult = F(1.0 if j == 0 else 2.0)
            upd = (lr2 * lr_mult) * (g1[n][j] + wd * w) + mu * h[(n, j)]
            assert np.allclose(w2.allWeights[n][j], w - upd, rtol=2e-5, atol=1e-8), (n, j)
    net.close(); twin.close()


def test_training_reduces_loss_and_is_reproducible(oracle):
    net, *_ = make(oracle, "lenet", 16, 2)
    w0 = net.getWeights()
    losses = []
    for _ in range(30):
        net.train(1)
        losses.append(oracle.sparknet_last_loss(net.state))
    assert abs(losses[0] - np.log(10)) < 0.5 and losses[-1] < 0.5 * losses[0]
    net2, *_ = make(oracle, "lenet", 16, 2)
    net2.setWeights(w0)
    net2.train(30)
    assert oracle.sparknet_last_loss(net2.state) == losses[-1]
    net.close(); net2.close()


def test_test_net_scores_are_sums_in_name_order_and_chance_accuracy(oracle):
    net, src, imgs, labs = make(oracle, "cifar10_quick", 20, 5)
    net.setTestData(MinibatchSource(imgs, labs), 5)
    scores = net.test()                                    # sums over 5 batches: [accuracy, loss]
    assert scores.shape == (2,)
    assert 0.0 <= scores[0] / 5 <= 0.35                    # random init ~ 10% (CifarSpec.scala:81-92)
    assert abs(scores[1] / 5 - np.log(10)) < 0.3
    # TEST net has the auto-inserted split layers after their producers (indices 0/1 still data/label)
    net.close()


def test_caffenet_structure(oracle):
    sol = zoo.solver_with_net("caffenet", train_bs=1, test_bs=1, random_seed=1, display=0)
    net = CaffeNet.from_solver(sol, oracle)
    assert net.numLayers == 25
    assert net.layerNames[2] == "conv1" and net.layerNames[-1] == "loss"
    shapes = {n: [a.shape for a in ws] for n, ws in net.getWeights().allWeights.items() if ws}
    assert shapes["conv2"] == [(256, 48, 5, 5), (256,)] and shapes["fc6"] == [(4096, 9216), (4096,)]
    assert sum(int(np.prod(s)) for ss in shapes.values() for s in ss) == 60965224
    net.close()
