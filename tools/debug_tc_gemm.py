"""Exercise the tcgen05 GEMM core with dense operands in every storage / feed combination."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sparknet_b200.library import load_library
lib = load_library()
lib.set_mode_gpu()
fn = lib.sparknet_test_tc_gemm
fn.restype = ctypes.c_int
FP = ctypes.POINTER(ctypes.c_float)
fn.argtypes = [ctypes.c_int] * 6 + [FP, FP, FP]
def run(a_mn, b_mn, tma_a, M, N, R, A, B):
    As = np.ascontiguousarray(A.T if a_mn else A, dtype=np.float32)
    Bs = np.ascontiguousarray(B.T if b_mn else B, dtype=np.float32)
    D = np.zeros((M, N), dtype=np.float32)
    fn(a_mn, b_mn, tma_a, M, N, R, As.ctypes.data_as(FP), Bs.ctypes.data_as(FP), D.ctypes.data_as(FP))
    return D
verbose = len(sys.argv) > 1
rng = np.random.default_rng(0)
for (M, N, R) in [(128, 16, 8), (128, 16, 32), (128, 64, 64), (256, 128, 96), (200, 100, 70), (128, 256, 512), (384, 192, 40)]:
    A = rng.integers(-3, 4, size=(M, R)).astype(np.float32)
    B = rng.integers(-3, 4, size=(N, R)).astype(np.float32)
    want = A @ B.T
    for a_mn in (0, 1):
        for b_mn in (0, 1):
            for tma_a in (0, 1):
                if tma_a and ((a_mn == 0 and R % 4) or (a_mn == 1 and M % 4)):
                    continue
                D = run(a_mn, b_mn, tma_a, M, N, R, A, B)
                err = np.abs(D - want).max()
                print(f"M{M} N{N} R{R} a_mn={a_mn} b_mn={b_mn} tma_a={tma_a}: max err {err:g}  nan={np.isnan(D).sum()}")
                if verbose and err > 0 and (M, N, R) == (128, 16, 8):
                    # probe with identity-like operands to reveal where elements land
                    Ap = (np.arange(M)[:, None] * 8 + np.arange(R)[None, :]).astype(np.float32) % 1024
                    Bp = np.zeros((N, R), dtype=np.float32); Bp[np.arange(8), np.arange(8)] = 1
                    Dp = run(a_mn, b_mn, tma_a, M, N, R, Ap, Bp)
                    print("  A-probe rows 0..3:", Dp[:4, :8].tolist()); print("  A-probe rows 32..33:", Dp[32:34, :8].tolist())
                    Ap = np.zeros((M, R), dtype=np.float32); Ap[np.arange(8), np.arange(8)] = 1
                    Bp = (np.arange(N)[:, None] * 8 + np.arange(R)[None, :]).astype(np.float32)
                    Dp = run(a_mn, b_mn, tma_a, M, N, R, Ap, Bp)
                    print("  B-probe D[:8,:4].T (B rows 0..3):", Dp[:8, :4].T.tolist())
