#!/usr/bin/env python
"""bench.py — the hot path of BASELINE.json measured on B200.

Metric: CaffeNet (bvlc_reference_caffenet, batch 256 per GPU) training images/sec with SparkNet's
parallel-SGD rounds (tau = 50 local solver iterations per worker, then weight averaging), on
synthetic data of the config's shape.  One "step" = one solver iteration over one batch; the timed
region runs exactly K steps as rounds of tau with the averaging inside it.

    python bench.py --gpus N --steps K --warmup W             # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W     # the reference's CPU path (oracle port)

Prints ONE JSON line (see the task contract): value = whole-job images/s with inputs resident in
HBM, e2e = the same through the reference-facing C ABI with host buffers (data callbacks + H2D in
the timed region), roofline for the dominant kernels, cpu_baseline on the box's host cores.
"""
from __future__ import annotations

import argparse
import ctypes
import glob
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
from sparknet_b200.net import CaffeNet, MinibatchSource  # noqa: E402

WORKLOADS = {
    # name: (zoo net, batch, tau)
    "caffenet": ("caffenet", 256, 50),
    "cifar10_quick": ("cifar10_quick", 100, 10),
    "cifar10_full": ("cifar10_full", 100, 50),
    "lenet": ("lenet", 64, 1),
}
# Algorithmic work per iteration, SURVEY.md §8(d) (FLOPs = 2*MACs of conv + inner-product fwd+bwd)
GEMM_GFLOP_PER_ITER = {"caffenet": 1058.7, "cifar10_quick": 6.921, "cifar10_full": 6.887, "lenet": 0.844}
ELTWISE_MB_PER_ITER = {"caffenet": 5915.6, "cifar10_quick": 92.6, "cifar10_full": 125.4, "lenet": 11.9}
PARAMS = {"caffenet": 60965224, "cifar10_quick": 145578, "cifar10_full": 89578, "lenet": 431080}


def synthetic_pool(name, batch, n, seed):
    c, h, w, classes, _, _ = zoo.GEOMETRY[name]
    rng = np.random.default_rng(seed)
    imgs, labs = [], []
    for _ in range(n):
        x = rng.random((batch, c, h, w), dtype=np.float32)
        if name != "lenet":
            x = x * np.float32(255.0) - np.float32(128.0)
        imgs.append(np.ascontiguousarray(x))
        labs.append(rng.integers(0, classes, size=(batch,)).astype(np.float32))
    return imgs, labs


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.gpu_index = gpu_index
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-i", str(gpu_index), "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [l.strip().split(", ") for l in open(self.f.name) if l.strip()]
        os.unlink(self.f.name)
        sm = [float(r[1]) for r in rows if len(r) >= 9]
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            for name, v in zip(names, r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        busy = [s for s in sm if s > 0]
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": float(rows[0][2]), "samples": len(sm),
                "power_w_max": max(float(r[3]) for r in rows), "reasons": sorted(reasons)}


def find_openblas():
    for pat in ("scipy.libs/libscipy_openblas*.so", "numpy.libs/libscipy_openblas*.so"):
        for sp in sys.path:
            hits = glob.glob(os.path.join(sp, pat))
            if hits:
                sym = "scipy_cblas_sgemm64_" if "openblas64_" in os.path.basename(hits[0]) else "scipy_cblas_sgemm"
                return hits[0], sym
    return None, None


def cpu_reference_run(workload, steps, warmup, budget_s):
    """The reference's CPU path for this workload (solver_step -> forward/backward/update through the
    raw C ABI with set_mode_cpu), here the oracle port because the reference cannot be compiled in
    this image.  Each step is one solver iteration on a bounded sample batch."""
    from tests.harness import load_oracle
    name, batch, tau = WORKLOADS[workload]
    oracle = load_oracle()
    cores = os.cpu_count() or 1
    oracle.oracle_set_threads(cores)
    blas_path, blas_sym = find_openblas()
    blas = "naive row-parallel sgemm (std::thread)"
    if blas_path and blas_sym == "scipy_cblas_sgemm":      # 32-bit-int interface only
        if oracle.oracle_use_blas(blas_path.encode(), blas_sym.encode()) == 0:
            blas = "OpenBLAS " + os.path.basename(blas_path)
    oracle.set_mode_cpu()

    def run(sample_batch, n_steps):
        sol = zoo.solver_with_net(name, train_bs=sample_batch, test_bs=1, random_seed=1701, display=0)
        net = CaffeNet.from_solver(sol, oracle)
        imgs, labs = synthetic_pool(name, sample_batch, 2, 1701)
        net.setTrainData(MinibatchSource(imgs, labs))
        t0 = time.perf_counter()
        oracle.solver_step(net.state, n_steps)
        dt = time.perf_counter() - t0
        net.close()
        return dt

    t_img = run(2, 1) / 2.0                                 # calibration: seconds per image
    total_steps = max(1, steps + warmup)
    sample = int(max(1, min(batch, budget_s / max(t_img * total_steps, 1e-9))))
    if warmup:
        run(sample, warmup)
    dt = run(sample, steps)
    return {"value": sample * steps / dt, "ms_per_step": dt / steps * 1e3, "cores": cores, "sample_batch": sample,
            "blas": blas, "sample": f"{steps} solver iteration(s) of {name} on a batch of {sample} (of {batch}) images, "
                                    f"CPU oracle port, {blas}, {cores} threads"}


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


# measured once per round with ncu (see profiles/): DRAM bytes moved by the conv / inner-product kernels of one iteration
NCU_GEMM_TRAFFIC_BYTES = {"caffenet": 3669.1e6}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="caffenet", choices=list(WORKLOADS))
    ap.add_argument("--tau", type=int, default=None)
    ap.add_argument("--precision", default=None, choices=[None, "ffma", "tf32"])
    ap.add_argument("--pool", type=int, default=8, help="distinct synthetic minibatches replayed cyclically")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    config = {"workload": f"{name} batch {batch}/GPU, tau={tau} local SGD iterations per averaging round "
                          f"(BASELINE.json configs[{ {'caffenet': 3, 'cifar10_quick': 1, 'cifar10_full': 2, 'lenet': 0}[name] }])",
              "net": name, "batch_per_gpu": batch, "tau": tau, "global_batch": batch * args.gpus,
              "parallelism": f"dp{args.gpus} (one replica per GPU, weights averaged every tau steps)",
              "l2_policy": f"inputs larger than L2 ({args.pool} distinct minibatches cycled; activations + parameters per step far exceed 126 MB)"
              if name == "caffenet" else f"{args.pool} distinct minibatches cycled"}

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_reference_run(args.workload, steps, warmup, budget_s=150.0)
        line = {"impl": "reference", "metric": f"{name}_train_images_per_sec", "value": r["value"], "unit": "images/s",
                "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config,
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
    imgs, labs = synthetic_pool(name, batch, args.pool, 1701 + rank)
    src = MinibatchSource(imgs, labs)
    net.copy_threads = int(os.environ.get("SPARKNET_COPY_THREADS", 0)) or max(1, min(16, (os.cpu_count() or 1) // max(1, world)))   # host threads of the data callback's copy
    net.setTrainData(src)
    averager = make_averager(lib, net, rank, world)
    averager.broadcast_from_rank0()                     # all replicas start from the same weights (CifarApp.scala:92-99)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        lib.sparknet_sync(st)

    def run_steps(n):
        """n solver iterations as rounds of tau, averaging after every round (CifarApp.scala:118-134)."""
        done, avg_ms = 0, 0.0
        while done < n:
            k = min(tau, n - done)
            lib.solver_step(st, k)
            avg_ms += averager.average()
            done += k
        return avg_ms

    exchange = {}                                       # device-side clock of the last timed region's exchanges

    def peer_clock():
        return averager.timing() if hasattr(averager, "timing") else (0, 0.0, 0.0)

    def timed(n):
        barrier()
        peer_clock()                                    # reset the kernel's own clock
        l0 = lib.sparknet_kernel_launches(st)
        lib.sparknet_timer_start(st)
        avg_ms = run_steps(n)
        ms = lib.sparknet_timer_stop_ms(st)
        barrier()
        _, wait_ms, xchg_ms = peer_clock()
        t = torch.tensor([ms, avg_ms, wait_ms, xchg_ms], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        exchange.update(wait_ms=float(t[2]), exchange_ms=float(t[3]))
        return float(t[0]), float(t[1]), lib.sparknet_kernel_launches(st) - l0

    # ---- (1) e2e: host buffers through the callback ABI, H2D inside the timed region ----------
    run_steps(warmup)
    e2e_ms, e2e_avg_ms, _ = timed(steps)
    h2d = int(imgs[0].nbytes + labs[0].nbytes)
    # ---- (2) value: the same steps with the minibatch pool resident in HBM --------------------
    lib.sparknet_set_resident_inputs(st, args.pool)
    run_steps(warmup)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    ms, avg_ms, launches = timed(steps)
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
                a = 0.0
                for _ in range(nr):
                    lib.solver_step(st, t)
                    a += averager.average()
                return a
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
    roofline = {"bound": "tensor", "kernel": "conv + inner-product implicit GEMMs (fwd, dgrad, wgrad) of one iteration",
                "achieved": achieved_tflops, "peak": tf32_peak, "unit": "TFLOP/s",
                "frac": achieved_tflops / tf32_peak if tf32_peak else None,
                # dram__bytes_read.sum + dram__bytes_write.sum summed over the 23 tensor-core launches of ONE iteration
                # (ncu --set full capture of this command, profiles/r1_ncu_gemm_iteration.md); null for other workloads
                "traffic": NCU_GEMM_TRAFFIC_BYTES.get(name), "traffic_unit": "bytes per iteration (all launches of the group)",
                "peak_source": "torch.matmul TF32 8192^3 best of 10, measured in this run (MEASURED_PEAKS.json has bf16 only)",
                "algorithmic_gflop_per_launch_group": GEMM_GFLOP_PER_ITER[name], "ms": gemm_ms,
                "share_of_step": gemm_ms / (gemm_ms + other_ms + u.value) if gemm_ms else None}
    bw_bytes = ELTWISE_MB_PER_ITER[name] * 1e6 + 24.0 * PARAMS[name]
    roofline_hbm = {"bound": "hbm", "kernel": "pool/relu/lrn/dropout/softmax/bias-grad + fused SGD update of one iteration",
                    "achieved": bw_bytes / 1e9 / ((other_ms + u.value) * 1e-3) if other_ms + u.value > 0 else 0.0,
                    "peak": hbm_peak, "unit": "GB/s", "traffic": None,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650",
                    "ms": other_ms + u.value, "sgd_update_ms": u.value,
                    "sgd_update_gbs": 24.0 * PARAMS[name] / 1e9 / (u.value * 1e-3) if u.value > 0 else None}
    roofline_hbm["frac"] = roofline_hbm["achieved"] / hbm_peak

    line = {"metric": f"{name}_train_images_per_sec", "value": value, "unit": "images/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "tf32" if precision == 1 else "f32", "data": "synthetic", "config": config,
            "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_ms / steps,
                    "host_copy_threads": net.copy_threads,
                    "note": "solver_step through the C ABI; Python data callbacks copy every minibatch from pageable host memory into "
                            "the library's pinned buffers (sparknet_host_copy), H2D on the copy stream, and every iteration's loss "
                            "is copied back to a pinned host ring (4 bytes, asynchronous)"},
            "gpu_launches": int(launches), "clocks": clk,
            "averaging": {"rounds": n_rounds, "ms_per_round": avg_ms / n_rounds, "share_of_round": avg_ms / ms if ms else None,
                          "method": averager.method,
                          # ms_per_round is stream time from the end of this replica's last iteration to the end of the
                          # exchange: it contains the wait for the slowest replica.  The kernel's own clock splits it:
                          "wait_for_slowest_replica_ms": exchange_value.get("wait_ms"),
                          "exchange_ms": exchange_value.get("exchange_ms"),
                          "exchange_share_of_round": (exchange_value.get("exchange_ms", 0.0) * n_rounds / ms) if ms else None},
            "roofline": roofline, "roofline_hbm": roofline_hbm, "layer_ms_fwd_bwd": layer_ms,
            "published_context": {"caffenet_k40_native_caffe_images_per_sec": 193,
                                  "source": "caffe/docs/performance_hardware.md:19 (other hardware, 1 GPU, no averaging)"}}
    if sweep:
        line["tau_sweep"] = sweep
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
    net.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
