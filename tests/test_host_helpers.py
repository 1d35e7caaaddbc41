"""Host-side helpers of the library that need no GPU: the multi-threaded copy / widening that data callbacks use to fill
the pinned staging buffer (include/ccaffe.h sparknet_host_copy, sparknet_host_widen)."""
import ctypes

import numpy as np
import pytest

from sparknet_b200.library import load_library


@pytest.fixture(scope="module")
def lib():
    return load_library()


@pytest.mark.parametrize("nbytes,threads", [(1048580, 16), (1048577, 3), (4096 * 16 * 5 + 1, 16), (1 << 20, 16), ((1 << 22) + 4095, 7),
                                            (12345, 8), (1, 4), (158298112 // 64 + 3, 64)])
def test_host_copy_moves_every_byte(lib, nbytes, threads):
    # sizes whose per-thread share is already a multiple of the 4096-byte chunk granule, plus a remainder, used to lose the tail
    src = np.random.default_rng(nbytes).integers(0, 256, size=nbytes, dtype=np.uint8)
    dst = np.full(nbytes + 8, 0xEE, dtype=np.uint8)
    lib.sparknet_host_copy(dst.ctypes.data, src.ctypes.data, nbytes, threads)
    assert np.array_equal(dst[:nbytes], src)
    assert (dst[nbytes:] == 0xEE).all()                              # and nothing past the end


@pytest.mark.parametrize("n,threads,signed", [(1 << 20, 8, True), ((1 << 20) + 7, 16, False), (100, 4, True), (3 * 227 * 227 * 5, 5, True)])
def test_host_widen_is_the_java_byte_to_float_conversion(lib, n, threads, signed):
    src = np.random.default_rng(n).integers(-128 if signed else 0, 128 if signed else 256, size=n).astype(np.int8 if signed else np.uint8)
    dst = np.full(n + 2, -7.0, dtype=np.float32)
    lib.sparknet_host_widen(dst.ctypes.data, src.ctypes.data, n, 1 if signed else 0, threads)
    assert np.array_equal(dst[:n], src.astype(np.float32))
    assert (dst[n:] == -7.0).all()
