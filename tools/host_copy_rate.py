"""GB/s of sparknet_host_copy (pageable -> pinned-sized destination) by thread count, 158 MB like a CaffeNet minibatch."""
import ctypes, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sparknet_b200.library import load_library
lib = load_library()
n = 256 * 3 * 227 * 227
src = [np.random.default_rng(i).random(n, dtype=np.float32) for i in range(3)]
dst = np.empty(n, dtype=np.float32)
print("cpus", os.cpu_count())
for th in (1, 4, 8, 16, 24, 32):
    best = 1e9
    for it in range(6):
        t0 = time.perf_counter()
        lib.sparknet_host_copy(dst.ctypes.data, src[it % 3].ctypes.data, dst.nbytes, th)
        best = min(best, time.perf_counter() - t0)
    print(f"threads {th:2d}: {best * 1e3:6.2f} ms  {dst.nbytes / best / 1e9:6.1f} GB/s", flush=True)
