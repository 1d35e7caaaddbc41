"""Clocks per TF32 tcgen05.mma issued back to back (csrc/conv_flat.cu mma_rate_kernel): the tensor-pipe ceiling."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sparknet_b200.library import load_library
lib = load_library()
lib.set_mode_gpu()
f = lib.sparknet_debug_mma_rate
f.restype = ctypes.c_float
f.argtypes = [ctypes.c_int] * 6 + [ctypes.POINTER(ctypes.c_float)]
print("pipeline-like pattern (per tap: nacc accumulators x 4 k-steps; A row start and B stage change per tap):")
for cg in (1, 2):
    for n, nacc in ((48, 4), (96, 4), (128, 4), (128, 2), (192, 2), (256, 2), (256, 1)):
        for mode, name in ((1000, "no commit"), (1001, "commit per tap")):
            io = ctypes.c_float(0)
            r = f(cg, n, 500, mode, nacc, 148, ctypes.byref(io))
            print(f"  cg {cg} N {n:3d} nacc {nacc} {name:15s}: {r:7.1f} clk/MMA (ideal {n/2:5.1f})", flush=True)
for ctas in (148,):
    for cg in (1, 2):
        for n in (48, 96, 128, 192, 256):
            for shift, nacc in ((0, 1), (3, 1), (0, 2)):
                if nacc * n > 512: continue
                io = ctypes.c_float(0)
                r = f(cg, n, 2000, shift, nacc, ctas, ctypes.byref(io))
                print(f"ctas {ctas:3d} cg {cg} N {n:3d} shift {shift} nacc {nacc}: {r:7.1f} clk/MMA (ideal {n/2:5.1f}), issue-only {io.value:6.1f}", flush=True)
