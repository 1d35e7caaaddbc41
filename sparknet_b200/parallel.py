"""Weight averaging across the replicas of one box — the replacement of SparkNet's
collect / reduce(WeightCollection.add) / scalarDivide / broadcast / setWeights round
(src/main/scala/apps/CifarApp.scala:92-99,133-134; src/main/scala/libs/Net.scala:14-47,132-172).

One process per GPU.  Every rank contributes its whole weight arena; the result on every rank is
    (((w_0 + w_1) + w_2) + ... + w_{N-1}) / (float)N
in IEEE fp32, summed in RANK ORDER — bit-exact against the reference's sequential JVM sum.

`PeerAverager` (the default): ONE kernel per GPU does the whole exchange over NVLink peer memory
(csrc/kernels.cu average_peers_kernel): every GPU folds its slice of the arena from all replicas in rank
order and stores the result into every replica — 2(N-1)/N * P bytes per GPU.  torch.distributed only
carries the 64-byte IPC handles at set-up.
`Averager` (SPARKNET_AVG=nccl, kept for comparison): NCCL all-gather of the whole arena as the transport,
then this library's fold kernel (average_replicas) over the gathered copies, (N-1) * P bytes per GPU.
NCCL's own all-reduce is not used because its ring/tree order is not the reference's order.

torch / torch.distributed are plumbing here (process group, device buffer for the gather)."""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np


class _DevArray:
    """Minimal __cuda_array_interface__ carrier so torch can wrap the library's weight arena."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3,
                                         "strides": None}


class Averager:
    def __init__(self, lib, net, rank: int, world: int):
        self.lib, self.net, self.rank, self.world = lib, net, rank, world
        self.method = "identity (1 replica)" if world == 1 else \
            "NCCL all-gather over NVLink + rank-ordered fp32 reduce kernel on the weight arena"
        if world == 1:
            return
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        ptr, cnt = ctypes.c_void_p(), ctypes.c_long()
        assert lib.sparknet_param_arena(net.state, ctypes.byref(ptr), ctypes.byref(cnt)) == 0
        self.n = int(cnt.value)
        self.arena = torch.as_tensor(_DevArray(ptr.value, self.n), device="cuda")
        assert self.arena.data_ptr() == ptr.value
        self.gathered = torch.empty((world, self.n), dtype=torch.float32, device="cuda")
        self.stream = torch.cuda.ExternalStream(lib.sparknet_compute_stream(net.state))
        self.ptrs = (ctypes.c_void_p * world)(*[self.gathered[r].data_ptr() for r in range(world)])
        self.ev0 = torch.cuda.Event(enable_timing=True)
        self.ev1 = torch.cuda.Event(enable_timing=True)

    def _touch(self):
        ptr, cnt = ctypes.c_void_p(), ctypes.c_long()
        self.lib.sparknet_param_arena(self.net.state, ctypes.byref(ptr), ctypes.byref(cnt))   # heads -> device

    def broadcast_from_rank0(self) -> None:
        """All replicas start from rank 0's weights (the driver's first broadcast, CifarApp.scala:92-99)."""
        if self.world == 1:
            return
        self._touch()
        with self.torch.cuda.stream(self.stream):
            self.dist.broadcast(self.arena, src=0)
        self.stream.synchronize()

    def drain(self) -> float:
        return 0.0

    def average(self, sync: bool = True) -> float:
        """Average the weights of all replicas in place; returns the device time in ms (always synchronous)."""
        if self.world == 1:
            return 0.0
        self._touch()
        with self.torch.cuda.stream(self.stream):
            self.ev0.record(self.stream)
            self.dist.all_gather_into_tensor(self.gathered.view(-1), self.arena)
        # same stream: the reduce kernel runs after the gather without a host sync
        assert self.lib.sparknet_average_replicas_async(self.ptrs, self.world, self.n,
                                                        ctypes.c_void_p(self.arena.data_ptr())) == 0
        self.ev1.record(self.stream)
        self.ev1.synchronize()
        return float(self.ev0.elapsed_time(self.ev1))


class PeerAverager:
    """Weight averaging in one kernel per GPU over peer memory (include/ccaffe.h sparknet_peer_*)."""

    def __init__(self, lib, net, rank: int, world: int):
        self.lib, self.net, self.rank, self.world = lib, net, rank, world
        self.method = "identity (1 replica)" if world == 1 else \
            "one fused kernel per GPU over NVLink peer memory: rank-ordered fp32 fold of the own slice + store to all replicas"
        if world == 1:
            return
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        hb = lib.sparknet_peer_handle_bytes()
        mine = np.zeros(2 * hb, dtype=np.uint8)
        rc = lib.sparknet_peer_export(net.state, ctypes.c_void_p(mine.ctypes.data), ctypes.c_void_p(mine.ctypes.data + hb))
        assert rc == 0, rc
        everyone = [torch.empty(2 * hb, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(everyone, torch.from_numpy(mine).cuda())
        allh = np.stack([t.cpu().numpy() for t in everyone])                      # [world][2 * hb]
        arena_h = np.ascontiguousarray(allh[:, :hb])
        flags_h = np.ascontiguousarray(allh[:, hb:])
        rc = lib.sparknet_peer_attach(net.state, rank, world, ctypes.c_void_p(arena_h.ctypes.data),
                                      ctypes.c_void_p(flags_h.ctypes.data))
        ok = torch.tensor([1 if rc == 0 else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)              # every replica must have mapped every other one
        self.attached = int(ok) == 1
        if not self.attached:
            return
        ptr, cnt = ctypes.c_void_p(), ctypes.c_long()
        assert lib.sparknet_param_arena(net.state, ctypes.byref(ptr), ctypes.byref(cnt)) == 0
        self.n = int(cnt.value)
        self.arena = torch.as_tensor(_DevArray(ptr.value, self.n), device="cuda")
        self.stream = torch.cuda.ExternalStream(lib.sparknet_compute_stream(net.state))
        self.ev0 = torch.cuda.Event(enable_timing=True)
        self.ev1 = torch.cuda.Event(enable_timing=True)
        self._pending, self._spare = [], []

    def broadcast_from_rank0(self) -> None:
        """All replicas start from rank 0's weights (the driver's first broadcast, CifarApp.scala:92-99)."""
        if self.world == 1:
            return
        ptr, cnt = ctypes.c_void_p(), ctypes.c_long()
        self.lib.sparknet_param_arena(self.net.state, ctypes.byref(ptr), ctypes.byref(cnt))   # heads -> device
        with self.torch.cuda.stream(self.stream):
            self.dist.broadcast(self.arena, src=0)
        self.stream.synchronize()

    def average(self, sync: bool = True) -> float:
        """Average the weights of all replicas in place.  sync=True (default): waits for the exchange, checks that no
        replica timed out and returns the device time in ms (it includes waiting for the slowest replica to arrive).
        sync=False: only enqueues the exchange — the host goes on to the next round's first data callback while it
        runs; the time is collected by `drain()`."""
        if self.world == 1:
            return 0.0
        if not sync:
            ev0 = self._spare.pop() if self._spare else self.torch.cuda.Event(enable_timing=True)
            ev1 = self._spare.pop() if self._spare else self.torch.cuda.Event(enable_timing=True)
            ev0.record(self.stream)
            rc = self.lib.sparknet_average_peers(self.net.state)
            assert rc == 0, rc
            ev1.record(self.stream)
            self._pending.append((ev0, ev1))
            return 0.0
        self.ev0.record(self.stream)
        rc = self.lib.sparknet_average_peers(self.net.state)
        assert rc == 0, rc
        self.ev1.record(self.stream)
        self.ev1.synchronize()
        self.check()                                   # weights are undefined after a time-out: never train on
        return float(self.ev0.elapsed_time(self.ev1))

    def drain(self) -> float:
        """Wait for the exchanges enqueued with sync=False, check their status, return their summed device time in ms."""
        if self.world == 1:
            return 0.0
        total = 0.0
        for ev0, ev1 in self._pending:
            ev1.synchronize()
            total += float(ev0.elapsed_time(ev1))
            self._spare += [ev0, ev1]
        self._pending = []
        self.check()
        return total

    def timing(self):
        """(rounds, mean ms waiting for the slowest replica, mean ms of the exchange proper) since the last call,
        from the kernel's own clock (include/ccaffe.h sparknet_peer_timing)."""
        if self.world == 1:
            return 0, 0.0, 0.0
        w, x = ctypes.c_float(0), ctypes.c_float(0)
        n = self.lib.sparknet_peer_timing(self.net.state, ctypes.byref(w), ctypes.byref(x))
        return n, float(w.value), float(x.value)

    def check(self) -> None:
        assert self.world == 1 or self.lib.sparknet_peer_status(self.net.state) == 0, \
            "a replica did not arrive at an averaging round within the time-out"


def make_averager(lib, net, rank: int, world: int):
    """The multi-GPU driver's choice: peer-memory kernel unless SPARKNET_AVG=nccl."""
    import os
    import sys
    if os.environ.get("SPARKNET_AVG", "peer") == "nccl":
        return Averager(lib, net, rank, world)
    avg = PeerAverager(lib, net, rank, world)
    if world > 1 and not avg.attached:
        # GPUs without peer access to each other (or no CUDA IPC in this container): the exchange kernel cannot
        # run; the NCCL all-gather + fold kernel computes the same bits.  Said loudly, and recorded in `method`.
        if rank == 0:
            print("sparknet: peer-memory mapping failed on at least one replica; averaging through NCCL all-gather + "
                  "fold kernel instead", file=sys.stderr, flush=True)
        avg = Averager(lib, net, rank, world)
        avg.method += " (peer-memory mapping unavailable on this box)"
    return avg


def sequential_average(weight_lists):
    """Host restatement of the same arithmetic for N lists of numpy arrays (used by the gloo tests)."""
    n = len(weight_lists)
    out = []
    for parts in zip(*weight_lists):
        acc = parts[0].astype(np.float32, copy=True)
        for p in parts[1:]:
            acc = (acc + p).astype(np.float32)
        out.append((acc / np.float32(n)).astype(np.float32))
    return out


class HostAverager:
    """The same exchange on host memory through the C ABI's weight accessors and any torch.distributed
    backend (gloo in the CPU tests): every rank gathers all replicas' weights and folds them in rank order.
    Same arithmetic as `Averager`; used where no GPU exists and as the reference for it."""

    def __init__(self, net, rank: int, world: int):
        self.net, self.rank, self.world = net, rank, world
        self.method = "host all_gather + rank-ordered fp32 fold"

    def broadcast_from_rank0(self) -> None:
        if self.world == 1:
            return
        import torch
        import torch.distributed as dist
        w = self.net.getWeights()
        for name in w.layerNames:
            for a in w.allWeights[name]:
                t = torch.from_numpy(a)
                dist.broadcast(t, src=0)
        self.net.setWeights(w)

    def drain(self) -> float:
        return 0.0

    def average(self, sync: bool = True) -> float:
        """Same call shape as PeerAverager.average; the host path is synchronous either way."""
        if self.world == 1:
            return 0.0
        import torch
        import torch.distributed as dist
        w = self.net.getWeights()
        for name in w.layerNames:
            for j, a in enumerate(w.allWeights[name]):
                parts = [torch.empty_like(torch.from_numpy(a)) for _ in range(self.world)]
                dist.all_gather(parts, torch.from_numpy(a))
                w.allWeights[name][j] = sequential_average([[p.numpy()] for p in parts])[0]
        self.net.setWeights(w)
        return 0.0


def run_rounds(net, averager, tau: int, rounds: int) -> None:
    """The driver loop of src/main/scala/apps/CifarApp.scala:95-136 without the test phase:
    (broadcast/set weights) -> train(tau) -> average, `rounds` times."""
    for _ in range(rounds):
        net.train(tau)
        averager.average(sync=False)      # enqueue only where the averager can: the next round's first callback runs under it
    averager.drain()
    if hasattr(averager, "check"):
        averager.check()
