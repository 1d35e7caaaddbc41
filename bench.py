#!/usr/bin/env python
"""bench.py — the hot path of BASELINE.json measured on B200.

Metric: CaffeNet (bvlc_reference_caffenet, batch 256 per GPU) training images/sec with SparkNet's
parallel-SGD rounds (tau = 50 local solver iterations per worker, then weight averaging), on
synthetic data of the config's shape.  One "step" = one solver iteration over one batch; the timed
region runs exactly K steps as rounds of tau with the averaging inside it.

    python bench.py --gpus N --steps K --warmup W             # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W     # the reference's CPU path (oracle port)

Prints ONE JSON line (see the task contract):
  e2e      the headline: solver_step through the reference's own C ABI with HOST buffers — the fp32 data callback
           (set_train_data_callback) widens the byte pixels the driver holds into the library's buffer, as the
           reference's makeImageCallback does, then H2D + every iteration's loss D2H, all inside the timed region;
  e2e_u8   the same through the byte-pixel callback of the ABI superset (the device widens: a quarter of the host
           traffic and of the H2D bytes);
  value    the same steps with the minibatch pool resident in HBM (explains e2e: what the kernels alone do);
  roofline (tensor group), roofline_hbm (per-kernel table of everything else), parity (one iteration against the CPU
  oracle at the benchmark batch), cpu_baseline (the oracle on the box's host cores).
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from sparknet_b200 import zoo  # noqa: E402
from sparknet_b200.layers import TRAIN  # noqa: E402
from sparknet_b200.net import CaffeNet, MinibatchSource  # noqa: E402

WORKLOADS = {
    # name: (zoo net, batch, tau)
    "caffenet": ("caffenet", 256, 50),
    "cifar10_quick": ("cifar10_quick", 100, 10),
    "cifar10_full": ("cifar10_full", 100, 50),
    "lenet": ("lenet", 64, 1),
}
CONFIG_INDEX = {"caffenet": 3, "cifar10_quick": 1, "cifar10_full": 2, "lenet": 0}
# Algorithmic work per iteration, SURVEY.md §8(d) (FLOPs = 2*MACs of conv + inner-product fwd+bwd)
GEMM_GFLOP_PER_ITER = {"caffenet": 1058.7, "cifar10_quick": 6.921, "cifar10_full": 6.887, "lenet": 0.844}
ELTWISE_MB_PER_ITER = {"caffenet": 5915.6, "cifar10_quick": 92.6, "cifar10_full": 125.4, "lenet": 11.9}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def synthetic_pool(name, batch, n, seed):
    """pixels uniform over the byte range (the apps feed raw byte values: a signed Java byte widened to
    float, ByteImage.java:86-104), uniform [0,1) floats for LeNet; labels uniform ints stored as float.  cifar / CaffeNet
    pools are held as int8 — what the reference's driver holds — and widened on the way in."""
    c, h, w, classes, _, _ = zoo.GEOMETRY[name]
    rng = np.random.default_rng(seed)
    imgs, labs = [], []
    for _ in range(n):
        if name == "lenet":
            imgs.append(np.ascontiguousarray(rng.random((batch, c, h, w), dtype=np.float32)))
        else:
            imgs.append(rng.integers(-128, 128, size=(batch, c, h, w), dtype=np.int8))
        labs.append(rng.integers(0, classes, size=(batch,)).astype(np.float32))
    return imgs, labs


def as_float(imgs):
    return [np.ascontiguousarray(x, dtype=np.float32) for x in imgs]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms from the warm-up on (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,"
         "utilization.gpu")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-i", str(gpu_index), "-lms", "50"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [l.strip().split(", ") for l in open(self.f.name) if l.strip()]
        os.unlink(self.f.name)
        rows = [r for r in rows if len(r) >= 10]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            for name, v in zip(names, r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)

        def num(x):
            try:
                return float(x)
            except ValueError:
                return 0.0
        load = [num(r[1]) for r in rows if num(r[9]) > 0] or [num(r[1]) for r in rows]
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": num(rows[0][2]), "samples": len(rows),
                "samples_under_load": sum(1 for r in rows if num(r[9]) > 0), "period_ms": 50,
                "power_w_max": max(num(r[3]) for r in rows), "reasons": sorted(reasons)}


def fast_oracle():
    """The CPU oracle with its sgemm routed to the OpenBLAS of the scipy wheel and every host core in use; the thread
    count is set explicitly so that the arm is the same under torch.distributed.run (which exports OMP_NUM_THREADS=1)."""
    from tests.harness import find_openblas, load_oracle
    cores = host_cores()
    for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[k] = str(cores)
    oracle = load_oracle()
    oracle.oracle_set_threads(cores)
    blas_path, blas_sym = find_openblas()
    blas = "naive row-parallel sgemm (std::thread)"
    if blas_path and blas_sym == "scipy_cblas_sgemm":      # 32-bit-int interface only
        if oracle.oracle_use_blas(blas_path.encode(), blas_sym.encode()) == 0:
            blas = "OpenBLAS " + os.path.basename(blas_path)
            try:
                h = ctypes.CDLL(blas_path, mode=ctypes.RTLD_GLOBAL)
                for sym in ("scipy_openblas_set_num_threads", "openblas_set_num_threads"):
                    if hasattr(h, sym):
                        getattr(h, sym)(ctypes.c_int(cores))
                        break
            except OSError:
                pass
    oracle.set_mode_cpu()
    return oracle, cores, blas


def cpu_reference_run(workload, steps, warmup, budget_s):
    """The reference's CPU path for this workload (solver_step -> forward/backward/update through the
    raw C ABI with set_mode_cpu), here the oracle port because the reference cannot be compiled in
    this image.  Every step is one solver iteration; on the full batch when steps + warm-up fit the time
    budget, else on a bounded sample batch — which the line then states."""
    name, batch, tau = WORKLOADS[workload]
    oracle, cores, blas = fast_oracle()

    def run(sample_batch, n_warm, n_steps):
        sol = zoo.solver_with_net(name, train_bs=sample_batch, test_bs=1, random_seed=1701, display=0)
        net = CaffeNet.from_solver(sol, oracle)
        imgs, labs = synthetic_pool(name, sample_batch, 2, 1701)
        net.setTrainData(MinibatchSource(as_float(imgs), labs))
        if n_warm:
            oracle.solver_step(net.state, n_warm)
        t0 = time.perf_counter()
        oracle.solver_step(net.state, n_steps)
        dt = time.perf_counter() - t0
        net.close()
        return dt

    # calibration: seconds per iteration = a + b * images (the SGD update over the parameters does not depend on the batch)
    b1, b2 = min(batch, 4), min(batch, 16)
    t1, t2 = run(b1, 1, 1), run(b2, 1, 1)
    per_img = max((t2 - t1) / max(b2 - b1, 1), 1e-9) if b2 > b1 else t1 / b1
    fixed = max(t1 - per_img * b1, 0.0)
    total_steps = max(1, steps + warmup)
    sample = int(max(1, min(batch, (budget_s / total_steps - fixed) / per_img)))
    dt = run(sample, warmup, steps)
    full = sample == batch
    return {"value": sample * steps / dt, "ms_per_step": dt / steps * 1e3, "cores": cores, "sample_batch": sample, "blas": blas,
            "sample": f"{steps} solver iteration(s) of {name} after {warmup} warm-up iteration(s) on "
                      + (f"the full batch of {batch} images" if full else f"a batch of {sample} (of {batch}) images — the full "
                         f"batch does not fit the {budget_s:.0f} s budget of this arm")
                      + f", CPU oracle port, {blas}, {cores} threads"}


def measure_tf32_peak():
    """Tensor-core denominator for TF32 (MEASURED_PEAKS.json only has bf16): torch.matmul with TF32
    allowed on 8192^3, best of 10, the same way the driver measured the bf16 figure."""
    import torch
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn(n, n, device="cuda")
        b = torch.randn(n, n, device="cuda")
        for _ in range(3):
            a @ b
        best = 1e9
        for _ in range(10):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            a @ b
            e.record()
            torch.cuda.synchronize()
            best = min(best, s.elapsed_time(e))
        return 2 * n ** 3 / (best * 1e-3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def ncu_traffic(name):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of the tensor-core launches of one iteration, from the
    committed ncu capture of this command (profiles/ncu_gemm_traffic.json names the capture it came from)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_gemm_traffic.json")))
        return t.get(name, {}).get("bytes"), t.get(name, {}).get("source")
    except (OSError, ValueError):
        return None, None


def train_layers(sol):
    out = []
    for l in sol.net_param.layer:
        inc = getattr(l, "include", None)
        if inc and all(r.phase != TRAIN for r in inc):
            continue
        out.append(l)
    return out


def blob_sizes(lib, st):
    out = {}
    for i in range(lib.num_data_blobs(st)):
        b = lib.get_data_blob(st, i)
        n = 1
        for k in range(lib.get_num_axes(b)):
            n *= lib.get_axis_shape(b, k)
        out[lib.get_data_blob_name(st, i).decode()] = n
    return out


def moved_bytes(layers, sizes, lf, lb):
    """Bytes the bandwidth-bound kernels of one iteration actually move (fp32 4 B, int32 mask 4 B, dropout mask 1 B), per
    layer and direction; a layer that launched nothing (its work rides in a neighbour's kernel) moves nothing."""
    out = {}
    for i, l in enumerate(layers):
        t = l.type
        if t in ("Convolution", "InnerProduct", "JavaData", "Accuracy", "Split"):
            continue
        n_in = sizes[l.bottom[0]]
        n_out = sizes[l.top[0]]
        f = b = 0
        if t == "Pooling":
            is_max = l.pooling_param.pool == 0
            f = 4 * n_in + (8 if is_max else 4) * n_out
            b = 4 * n_in + (8 if is_max else 4) * n_out
            below = layers[i - 1] if i else None
            if is_max and below is not None and below.type == "ReLU" and below.top[0] == l.bottom[0] and lb[i - 1] == 0:
                b += 4 * n_out                                   # the pooled values gate the gradient (fused ReLU backward)
        elif t == "ReLU":
            f, b = 8 * n_in, 12 * n_in
        elif t == "LRN":
            f, b = 8 * n_in, 12 * n_in                           # no scale tensor is kept
            nxt = [m for m in layers[i + 1:] if l.top[0] in m.bottom]
            if nxt and nxt[0].type == "Convolution" and getattr(l.lrn_param, "norm_region", 0) in (0, None):
                f += 4 * n_in                                    # ... but the next convolution's packed input is written too
        elif t == "Dropout":
            f, b = 9 * n_in, 9 * n_in
        elif t == "SoftmaxWithLoss":
            f, b = 16 * n_in, 8 * n_in
        out[l.name] = (f if lf[i] else 0, b if lb[i] else 0)
    return out


def parity_check(lib, name, batch, precision):
    """One forward + backward of the benchmark's own configuration (same batch, same mode) on the CUDA path and on the CPU
    oracle: identical inputs, weights and dropout masks; the loss and the last inner product's weight gradient."""
    oracle, _, _ = fast_oracle()
    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=1, random_seed=1701, display=0)
    o, p = CaffeNet.from_solver(sol, oracle), CaffeNet.from_solver(sol, lib)
    lib.sparknet_set_precision(p.state, precision)
    imgs, labs = synthetic_pool(name, batch, 1, 4242)
    imgs = as_float(imgs)
    o.setTrainData(MinibatchSource(imgs, labs))
    p.setTrainData(MinibatchSource(imgs, labs))
    p.setWeights(o.getWeights())
    layers = train_layers(sol)
    sizes = blob_sizes(oracle, o.state)
    rng = np.random.default_rng(0)
    for idx, l in enumerate(layers):
        if l.type == "Dropout":
            m = (rng.random(sizes[l.top[0]]) > 0.5).astype(np.uint32)
            mp = m.ctypes.data_as(ctypes.POINTER(ctypes.c_uint))
            assert oracle.sparknet_inject_dropout_mask(o.state, idx, mp, m.size) == 0
            assert lib.sparknet_inject_dropout_mask(p.state, idx, mp, m.size) == 0
    t0 = time.perf_counter()
    o.forward(); o.backward()
    t_cpu = time.perf_counter() - t0
    p.forward(); p.backward()

    def loss_of(net):
        for i in range(net.lib.num_data_blobs(net.state)):
            if net.lib.get_data_blob_name(net.state, i).decode() == "loss":
                return float(net._view(net.lib.get_data_blob(net.state, i)).ravel()[0])
        return float("nan")
    lo, lp = loss_of(o), loss_of(p)
    last = max(i for i, l in enumerate(layers) if l.type == "InnerProduct")
    go = o._view(oracle.get_weight_blob(o.state, last, 0), True).astype(np.float64)
    gp = p._view(lib.get_weight_blob(p.state, last, 0), True).astype(np.float64)
    norm = float(np.max(np.abs(gp - go)) / max(float(np.max(np.abs(go))), 1e-30))
    l2 = float(np.linalg.norm(gp - go) / max(float(np.linalg.norm(go)), 1e-30))
    # free-running pass: the errors of all layers compound (tests/test_gpu_baseline_sizes.py holds every layer to 1e-3 on the
    # oracle's input and the whole net to 3e-3 on its own)
    tol = 3e-3 if precision == 1 else 1e-4
    o.close(); p.close()
    return {"config": f"{name} batch {batch}, one forward + backward, {'TF32' if precision == 1 else 'fp32 FFMA'} vs CPU oracle",
            "loss_cuda": lp, "loss_oracle": lo, "loss_rel_err": abs(lp - lo) / max(abs(lo), 1e-30),
            "layer": layers[last].name, "weight_grad_rel_err_normwise": norm, "weight_grad_rel_l2": l2, "tolerance": tol,
            "ok": bool(np.isfinite(lp) and abs(lp - lo) <= 1e-3 * max(abs(lo), 1.0) and norm <= tol),
            "oracle_seconds": t_cpu}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="caffenet", choices=list(WORKLOADS))
    ap.add_argument("--tau", type=int, default=None)
    ap.add_argument("--precision", default=None, choices=[None, "ffma", "tf32"])
    ap.add_argument("--pool", type=int, default=None,
                    help="distinct synthetic minibatches per rank, replayed cyclically (default: tau, SURVEY.md 8d; at most 64)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--tau-sweep", default=None,
                    help="comma list of tau values (BASELINE.json configs[4]: 1,10,50,200): adds averaging cost vs compute per "
                         "round for each, measured after the main run with the inputs resident in HBM")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    name, batch, tau = WORKLOADS[args.workload]
    tau = args.tau or tau
    steps = args.steps
    warmup = args.warmup if args.impl == "reference" else max(args.warmup, 3)   # timing rule: W >= 3 on the device
    pool_n = args.pool or max(2, min(tau, 64))
    config = {"workload": f"{name} batch {batch}/GPU, tau={tau} local SGD iterations per averaging round "
                          f"(BASELINE.json configs[{CONFIG_INDEX[name]}])",
              "net": name, "batch_per_gpu": batch, "tau": tau, "global_batch": batch * args.gpus,
              "parallelism": f"dp{args.gpus} (one replica per GPU, weights averaged every tau steps)",
              "l2_policy": "inputs larger than L2 (a pool of distinct minibatches per rank is cycled; activations + parameters "
                           "per step far exceed 126 MB)" if name == "caffenet" else "flush between steps: a pool of distinct minibatches "
                           "per rank is cycled and every step rewrites all activations and parameters"}

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_reference_run(args.workload, steps, warmup, budget_s=240.0)
        line = {"impl": "reference", "metric": f"{name}_train_images_per_sec", "value": r["value"], "unit": "images/s",
                "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config, "reference_sample_batch": r["sample_batch"],
                "cpu_baseline": {"value": r["value"], "unit": "images/s", "cores": r["cores"], "kind": "port",
                                 "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    # Keep stdout for the ONE JSON line: NCCL / torch may print banners ("NCCL version ...") to fd 1.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from sparknet_b200.library import load_library
    from sparknet_b200.parallel import make_averager
    lib = load_library()
    lib.set_device(local_rank)
    lib.set_mode_gpu()

    sol = zoo.solver_with_net(name, train_bs=batch, test_bs=1, random_seed=1701 + rank, display=0)
    net = CaffeNet.from_solver(sol, lib)
    st = net.state
    if args.precision:
        lib.sparknet_set_precision(st, {"ffma": 0, "tf32": 1}[args.precision])
    precision = lib.sparknet_get_precision(st)
    imgs, labs = synthetic_pool(name, batch, pool_n, 1701 + rank)
    bytes_pool = imgs[0].dtype == np.int8
    # host threads of the data callback's copy / widening
    net.copy_threads = int(os.environ.get("SPARKNET_COPY_THREADS", 0)) or max(1, min(16, host_cores() // max(1, world)))
    numa = ctypes.create_string_buffer(128)
    lib.sparknet_numa_info(numa, 128)
    averager = make_averager(lib, net, rank, world)
    averager.broadcast_from_rank0()                     # all replicas start from the same weights (CifarApp.scala:92-99)
    n_params = int(lib.sparknet_num_params(st))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        lib.sparknet_sync(st)

    def run_steps(n):
        """n solver iterations as rounds of tau, averaging after every round (CifarApp.scala:118-134).  The exchange is
        only enqueued: the next round's first data callback runs on the host while it is in flight."""
        done = 0
        while done < n:
            k = min(tau, n - done)
            lib.solver_step(st, k)
            averager.average(sync=False)
            done += k
        return averager.drain()

    exchange = {}                                       # device-side clock of the last timed region's exchanges

    def peer_clock():
        return averager.timing() if hasattr(averager, "timing") else (0, 0.0, 0.0)

    def timed(n):
        barrier()
        peer_clock()                                    # reset the kernel's own clock
        stats = (ctypes.c_double * 3)()
        lib.sparknet_input_stats(st, stats)             # reset
        l0 = lib.sparknet_kernel_launches(st)
        lib.sparknet_timer_start(st)
        avg_ms = run_steps(n)
        ms = lib.sparknet_timer_stop_ms(st)
        barrier()
        lib.sparknet_input_stats(st, stats)
        _, wait_ms, xchg_ms = peer_clock()
        t = torch.tensor([ms, avg_ms, wait_ms, xchg_ms], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        exchange.update(wait_ms=float(t[2]), exchange_ms=float(t[3]))
        return float(t[0]), float(t[1]), lib.sparknet_kernel_launches(st) - l0, {"callbacks": int(stats[0]), "callback_ms": stats[1],
                                                                                 "wait_for_staging_buffer_ms": stats[2]}

    # ---- (0) once, in warm-up: the exchange against a host fold of the gathered pre-average weights, bit for bit -------
    clocks = ClockSampler(local_rank) if rank == 0 else None
    bit_exact = None
    src = MinibatchSource(imgs, labs)
    if bytes_pool:
        net.setTrainDataWidening(src, signed=True)      # the reference's arrangement: bytes on the host, floats into the buffer
    else:
        net.setTrainData(src)
    if world > 1:
        lib.solver_step(st, 2)
        mine = np.empty(n_params, dtype=np.float32)
        lib.sparknet_get_weights_flat(st, mine.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), n_params)
        allw = [torch.empty(n_params, dtype=torch.float32, device="cuda") for _ in range(world)]
        dist.all_gather(allw, torch.from_numpy(mine).cuda())
        acc = allw[0].cpu().numpy().copy()
        for r in range(1, world):
            acc = (acc + allw[r].cpu().numpy()).astype(np.float32)       # rank order, fp32 (JavaNDArray.java:84-117)
        want = (acc / np.float32(world)).astype(np.float32)
        differed = not np.array_equal(allw[0].cpu().numpy(), allw[-1].cpu().numpy())
        del allw
        averager.average()
        lib.sparknet_get_weights_flat(st, mine.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), n_params)
        flag = torch.tensor([1 if mine.tobytes() == want.tobytes() else 0, 1 if differed else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        bit_exact = {"bit_exact": bool(int(flag[0]) == 1), "replicas_differed_before": bool(int(flag[1]) == 1),
                     "checked": "every replica's arena after one exchange in warm-up == host fold (((w0+w1)+...)/N, fp32, rank "
                                "order) of the gathered pre-average arenas, memcmp"}
        del acc, want, mine

    # ---- (1) e2e: host buffers through the reference's fp32 callback ABI, H2D inside the timed region ----------
    run_steps(warmup)
    e2e_ms, e2e_avg_ms, _, e2e_stats = timed(steps)
    h2d = int(imgs[0].size * 4 + labs[0].nbytes)
    # ---- (1b) e2e_u8: the same through the byte-pixel callback (ABI superset) -----------------------------------
    e2e_u8 = None
    if bytes_pool:
        net.setTrainDataBytes(MinibatchSource(imgs, labs), signed=True)
        run_steps(warmup)
        u8_ms, _, _, u8_stats = timed(steps)
        e2e_u8 = {"value": args.gpus * batch * steps / (u8_ms * 1e-3), "unit": "images/s", "ms_per_step": u8_ms / steps,
                  "h2d_bytes_per_step": int(imgs[0].nbytes + labs[0].nbytes), "d2h_bytes_per_step": 4, "input_path": u8_stats,
                  "note": "sparknet_set_train_data_callback_bytes: the callback copies the int8 NCHW pixels it holds into the "
                          "library's pinned buffer, the device widens them (bit-identical blob)"}
    # ---- (2) value: the same steps with minibatches resident in HBM --------------------
    # (the first `resident_n` of the pool: solver_step records one CUDA graph per input buffer, so the resident set is
    # kept small and walked once as set-up — recording is one-time work like the plan enumeration, not a training step)
    resident_n = min(pool_n, 8)
    lib.sparknet_set_resident_inputs(st, resident_n)
    lib.solver_step(st, resident_n + 2)
    run_steps(warmup)
    ms, avg_ms, launches, _ = timed(steps)
    exchange_value = dict(exchange)
    clk = clocks.stop() if clocks else None
    n_rounds = (steps + tau - 1) // tau
    value = args.gpus * batch * steps / (ms * 1e-3)
    e2e = args.gpus * batch * steps / (e2e_ms * 1e-3)

    # ---- (2b) tau sweep: averaging cost against compute per round (BASELINE.json configs[4]) ----------
    sweep = None
    if args.tau_sweep:
        sweep = []
        for t in [int(x) for x in args.tau_sweep.split(",")]:
            rounds = max(2, min(20, 200 // t))
            def run_rounds(nr, t=t):
                for _ in range(nr):
                    lib.solver_step(st, t)
                    averager.average(sync=False)
                return averager.drain()
            run_rounds(1)
            barrier()
            peer_clock()
            lib.sparknet_timer_start(st)
            a_ms = run_rounds(rounds)
            t_ms = lib.sparknet_timer_stop_ms(st)
            barrier()
            _, w_ms, x_ms = peer_clock()
            v = torch.tensor([t_ms, a_ms, w_ms, x_ms], device="cuda", dtype=torch.float64)
            if world > 1:
                dist.all_reduce(v, op=dist.ReduceOp.MAX)
            sweep.append({"tau": t, "rounds": rounds, "ms_per_round": float(v[0]) / rounds,
                          "averaging_ms_per_round": float(v[1]) / rounds, "averaging_share": float(v[1]) / float(v[0]),
                          "wait_for_slowest_replica_ms": float(v[2]), "exchange_ms": float(v[3]),
                          "exchange_share": float(v[3]) * rounds / float(v[0]),
                          "images_per_sec": args.gpus * batch * t * rounds / (float(v[0]) * 1e-3)})
    if hasattr(averager, "check"):
        averager.check()                                # every exchange completed (no replica timed out)

    # ---- (3) per-layer device timing of one iteration -> roofline of the dominant kernels ---------
    L = net.numLayers
    f = (ctypes.c_float * L)()
    b = (ctypes.c_float * L)()
    u = ctypes.c_float(0)
    for _ in range(2):
        lib.sparknet_profile_step(st, f, b, ctypes.byref(u), L)
    lf, lb = (ctypes.c_int * L)(), (ctypes.c_int * L)()
    lib.sparknet_profile_launches(st, lf, lb, L)
    layer_ms = {net.layerNames[i]: (round(f[i], 4), round(b[i], 4)) for i in range(L)}
    gemm_layers = [i for i in range(L) if net.layerNumBlobs[i] > 0]
    gemm_ms = sum(f[i] + b[i] for i in gemm_layers)
    other_ms = sum(f[i] + b[i] for i in range(L) if i not in gemm_layers)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    tf32_peak = measure_tf32_peak() if rank == 0 else 0.0
    achieved_tflops = GEMM_GFLOP_PER_ITER[name] / gemm_ms if gemm_ms > 0 else 0.0     # GFLOP / ms = TFLOP/s
    traffic, traffic_src = ncu_traffic(name)
    roofline = {"bound": "tensor", "kernel": "conv + inner-product implicit GEMMs (fwd, dgrad, wgrad) of one iteration",
                "achieved": achieved_tflops, "peak": tf32_peak, "unit": "TFLOP/s",
                "frac": achieved_tflops / tf32_peak if tf32_peak else None,
                # dram__bytes_read.sum + dram__bytes_write.sum summed over the tensor-core launches of ONE iteration
                "traffic": traffic, "traffic_unit": "bytes per iteration (all launches of the group)", "traffic_source": traffic_src,
                "peak_source": "torch.matmul TF32 8192^3 best of 10, measured in this run (MEASURED_PEAKS.json has bf16 only)",
                "algorithmic_gflop_per_launch_group": GEMM_GFLOP_PER_ITER[name], "ms": gemm_ms,
                "share_of_step": gemm_ms / (gemm_ms + other_ms + u.value) if gemm_ms else None}
    # bandwidth-bound kernels: bytes actually moved by the kernels that were launched, per layer and direction
    layers = train_layers(sol)
    sizes = blob_sizes(lib, st)
    table, tot_b, tot_ms = [], 0.0, 0.0
    if [l.name for l in layers] == net.layerNames:
        mv = moved_bytes(layers, sizes, list(lf), list(lb))
        for i, l in enumerate(layers):
            if l.name not in mv:
                continue
            for d, nbytes, t_ms in (("fwd", mv[l.name][0], f[i]), ("bwd", mv[l.name][1], b[i])):
                if nbytes <= 0 or t_ms <= 0:
                    continue
                gbs = nbytes / 1e9 / (t_ms * 1e-3)
                table.append({"layer": l.name, "dir": d, "mb": round(nbytes / 1e6, 2), "ms": round(t_ms, 4), "gbs": round(gbs, 1),
                              "frac": round(gbs / hbm_peak, 3)})
                if nbytes >= 8e6:                                 # smaller launches are latency-bound, listed but not averaged in
                    tot_b += nbytes
                    tot_ms += t_ms
    arena = ctypes.c_long(0)
    ptr = ctypes.c_void_p()
    lib.sparknet_param_arena(st, ctypes.byref(ptr), ctypes.byref(arena))
    sgd_bytes = 28.0 * arena.value                               # 3 reads + 3 writes + the tf32 shadow write, padded arena
    if u.value > 0:
        gbs = sgd_bytes / 1e9 / (u.value * 1e-3)
        table.append({"layer": "sgd_update", "dir": "-", "mb": round(sgd_bytes / 1e6, 2), "ms": round(u.value, 4), "gbs": round(gbs, 1),
                      "frac": round(gbs / hbm_peak, 3)})
        if sgd_bytes >= 8e6:
            tot_b += sgd_bytes
            tot_ms += u.value
    roofline_hbm = {"bound": "hbm", "kernel": "pool / lrn / dropout / softmax / unfused relu + fused SGD update of one iteration: bytes the "
                                              "launched kernels move (work fused into tensor kernels is not counted here)",
                    "achieved": tot_b / 1e9 / (tot_ms * 1e-3) if tot_ms > 0 else 0.0, "peak": hbm_peak, "unit": "GB/s", "traffic": None,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650",
                    "ms": other_ms + u.value, "sgd_update_ms": u.value, "per_kernel": table,
                    "algorithmic_mb_per_iter_survey_8d": ELTWISE_MB_PER_ITER[name] + 24.0 * n_params / 1e6}
    roofline_hbm["frac"] = roofline_hbm["achieved"] / hbm_peak

    line = {"metric": f"{name}_train_images_per_sec", "value": value, "unit": "images/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "tf32" if precision == 1 else "f32", "data": "synthetic", "config": config,
            "pool_minibatches": pool_n, "resident_minibatches": resident_n,
            "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_ms / steps, "host_copy_threads": net.copy_threads, "input_path": e2e_stats,
                    "note": "solver_step through the reference's C ABI (set_train_data_callback, fp32 buffers); the Python data "
                            "callback widens the int8 pixels it holds into the library's pinned buffer on host threads "
                            "(sparknet_host_widen — what makeImageCallback / ByteImage.cropInto do in the JVM), H2D on the copy "
                            "stream, and every iteration's loss is copied back to a pinned host ring (4 bytes, asynchronous)"
                    if bytes_pool else "solver_step through the reference's C ABI; fp32 minibatches copied into the pinned buffer"},
            "e2e_u8": e2e_u8, "numa": numa.value.decode(),
            "gpu_launches": int(launches), "clocks": clk,
            "graph": {"replays": int(lib.sparknet_graph_replays(st)), "programmatic_edges": int(lib.sparknet_graph_programmatic_edges(st))},
            "averaging": {"rounds": n_rounds, "ms_per_round": avg_ms / n_rounds, "share_of_round": avg_ms / ms if ms else None,
                          "method": averager.method,
                          # ms_per_round is stream time from the end of this replica's last iteration to the end of the
                          # exchange: it contains the wait for the slowest replica.  The kernel's own clock splits it:
                          "wait_for_slowest_replica_ms": exchange_value.get("wait_ms"),
                          "exchange_ms": exchange_value.get("exchange_ms"),
                          "exchange_share_of_round": (exchange_value.get("exchange_ms", 0.0) * n_rounds / ms) if ms else None,
                          "e2e_ms_per_round": e2e_avg_ms / n_rounds},
            "roofline": roofline, "roofline_hbm": roofline_hbm, "layer_ms_fwd_bwd": layer_ms,
            "published_context": {"caffenet_k40_native_caffe_images_per_sec": 193,
                                  "source": "caffe/docs/performance_hardware.md:19 (other hardware, 1 GPU, no averaging)"}}
    if bit_exact:
        line["averaging"].update(bit_exact)
    if sweep:
        line["tau_sweep"] = sweep
    net.close()
    if rank == 0 and args.gpus == 1 and not args.no_parity:
        line["parity"] = parity_check(lib, name, batch, precision)
    if rank == 0 and not args.no_cpu_baseline and args.gpus == 1:
        r = cpu_reference_run(args.workload, 1, 0, budget_s=20.0)
        line["cpu_baseline"] = {"value": r["value"], "unit": "images/s", "cores": r["cores"], "kind": "port",
                                "sample": r["sample"]}
    elif rank == 0:
        line["cpu_baseline"] = None
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
