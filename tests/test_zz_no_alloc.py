"""The allocation-free GEMM entry point, checked against the regular one.

`sparknet_tc_gemm_static_scratch` exists for callers that cannot allocate device memory while they
run: code inside a CUDA stream capture, latency-critical threads that cannot afford cudaMalloc's
device-wide synchronisation, and hosts that give the library a fixed memory budget.  Its split-K
slices live in a static 8 MiB device array instead of the on-demand workspace, so this test
  * runs two calls in a row (the second, smaller one sees the first one's leftovers in the shared
    array and must not be affected by them),
  * compares each against `sparknet_test_tc_gemm`, the existing entry point that allocates its own
    buffers and workspace (same kernels: the results must agree to the tolerance the TF32 tests use,
    1e-3 of the output scale — in practice they are identical),
  * checks that a problem needing more than the static array is refused with an error code.
The file name sorts last so that a failure here cannot hide any other test under `pytest -x`."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
F = np.float32
FP = ctypes.POINTER(ctypes.c_float)
TOL = 1e-3


def _bind(lib):
    new = lib.sparknet_tc_gemm_static_scratch
    new.restype = ctypes.c_int
    new.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    old = lib.sparknet_test_tc_gemm
    old.restype = ctypes.c_int
    old.argtypes = [ctypes.c_int] * 6 + [FP, FP, FP]
    return new, old


def test_static_scratch_gemm_matches_allocating_entry_point_twice_in_a_row(product):
    import torch
    new, old = _bind(product)
    rng = np.random.default_rng(1701)
    # (a_mn, b_mn, M, N, R): both split along R (1 output tile, long reduction); the second is smaller
    cases = [(0, 0, 128, 128, 8192), (1, 1, 100, 72, 4100)]
    for a_mn, b_mn, M, N, R in cases:
        A = rng.standard_normal((M, R)).astype(F)
        B = rng.standard_normal((N, R)).astype(F)
        As = np.ascontiguousarray(A.T if a_mn else A)
        Bs = np.ascontiguousarray(B.T if b_mn else B)
        want = np.zeros((M, N), dtype=F)
        assert old(a_mn, b_mn, 0, M, N, R, As.ctypes.data_as(FP), Bs.ctypes.data_as(FP), want.ctypes.data_as(FP)) == 0
        dA, dB = torch.from_numpy(As).cuda(), torch.from_numpy(Bs).cuda()     # the CALLER owns inputs and outputs
        dD = torch.full((M, N), float("nan"), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        rc = new(a_mn, b_mn, M, N, R, dA.data_ptr(), dB.data_ptr(), dD.data_ptr(), None)
        assert rc == 0
        torch.cuda.synchronize()
        got = dD.cpu().numpy()
        scale = np.abs(want).max()
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= TOL * scale
        exact = (A.astype(np.float64) @ B.astype(np.float64).T)
        assert np.abs(got - exact).max() <= TOL * np.abs(exact).max()          # and both are right


def test_static_scratch_gemm_refuses_what_does_not_fit(product):
    import torch
    new, _ = _bind(product)
    M, N, R = 4096, 256, 100000            # 32 tiles -> 10 slices of 4096 x 256 floats = 40 MiB > 8 MiB
    dA = torch.zeros((8,), dtype=torch.float32, device="cuda")                # never dereferenced: refused first
    assert new(0, 0, M, N, R, dA.data_ptr(), dA.data_ptr(), dA.data_ptr(), None) == -2
    assert new(0, 0, 0, N, R, dA.data_ptr(), dA.data_ptr(), dA.data_ptr(), None) == -1
