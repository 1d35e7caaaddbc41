"""World-size-2 test of the averaging round on CPU (gloo): the multi-process driver logic
(rendezvous, rank-ordered fold, momentum stays local) with the CPU oracle standing in for the
per-GPU replica.  The arithmetic is the one the CUDA path uses (tests/test_gpu_nets.py checks the
kernel bit-exactly against it)."""
import os
import socket

import numpy as np
import pytest

from sparknet_b200 import zoo
from sparknet_b200.net import CaffeNet, MinibatchSource, WeightCollection
from sparknet_b200.parallel import HostAverager, run_rounds, sequential_average

TAU, ROUNDS, BATCH = 2, 2, 8


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_replica(rank):
    from tests.harness import load_oracle, synthetic_batches
    oracle = load_oracle()
    sol = zoo.solver_with_net("lenet", train_bs=BATCH, random_seed=5, display=0)
    net = CaffeNet.from_solver(sol, oracle)
    imgs, labs = synthetic_batches("lenet", TAU * ROUNDS, BATCH, seed=1701 + rank)
    net.setTrainData(MinibatchSource(imgs, labs))
    return net


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    net = _make_replica(rank)
    avg = HostAverager(net, rank, world)
    avg.broadcast_from_rank0()
    run_rounds(net, avg, TAU, ROUNDS)
    w = net.getWeights()
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), *[a for n in w.layerNames for a in w.allWeights[n]])
    dist.destroy_process_group()


def test_two_process_rounds_match_single_process_emulation(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = np.load(tmp_path / "rank0.npz")
    r1 = np.load(tmp_path / "rank1.npz")
    for k in r0.files:
        assert r0[k].tobytes() == r1[k].tobytes()          # every replica holds the same averaged weights
    # single-process emulation of the same two workers
    nets = [_make_replica(0), _make_replica(1)]
    nets[1].setWeights(nets[0].getWeights())                # broadcast from rank 0
    for _ in range(ROUNDS):
        for n in nets:
            n.train(TAU)
        tot = WeightCollection.add(nets[0].getWeights(), nets[1].getWeights())
        tot.scalarDivide(2.0)
        for n in nets:
            n.setWeights(tot)
    w = nets[0].getWeights()
    flat = [a for n in w.layerNames for a in w.allWeights[n]]
    for k, a in zip(r0.files, flat):
        assert r0[k].tobytes() == a.tobytes()               # bit-exact: rank order == WeightCollection.add order


def test_sequential_average_is_rank_ordered():
    rng = np.random.default_rng(0)
    reps = [[(rng.standard_normal(1000) * 10.0 ** rng.integers(-6, 6, 1000)).astype(np.float32)] for _ in range(5)]
    got = sequential_average(reps)[0]
    acc = reps[0][0].copy()
    for r in reps[1:]:
        acc = (acc + r[0]).astype(np.float32)
    assert got.tobytes() == (acc / np.float32(5)).astype(np.float32).tobytes()
