"""CPU-side checks of the drop-in boundary: the built library exports every symbol that
include/ccaffe.h declares (no compute calls, no GPU needed)."""
import ctypes
import os
import re

from sparknet_b200.library import ABI_SYMBOLS, DEFAULT_LIB

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ccaffe.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = "\n".join(l for l in src.splitlines() if not l.lstrip().startswith(("#", "typedef")))
    names = re.findall(r"\b([a-z_][a-z0-9_]*)\s*\(", src)
    return [n for n in names if n not in ("defined", "void", "int", "float", "char", "long") and not n.endswith("_t")]


def test_library_exports_every_declared_symbol():
    assert os.path.exists(DEFAULT_LIB), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(DEFAULT_LIB)
    names = declared_symbols()
    assert len(names) >= 43 + 10
    for n in names:
        assert hasattr(lib, n), f"libccaffe.so does not export {n}"


def test_reference_abi_is_fully_declared():
    # the 37 functions CaffeLibrary.java binds + the 6 it does not (db writers, error callback): 43
    names = set(declared_symbols())
    for n in ABI_SYMBOLS:
        assert n in names
    for n in ("create_db", "write_to_db", "commit_db_txn", "close_db", "save_mean_image", "set_global_error_callback"):
        assert n in names


def test_scalar_queries_work_without_a_gpu():
    lib = ctypes.CDLL(DEFAULT_LIB)
    lib.get_int_size.restype = ctypes.c_int
    lib.get_dtype_size.restype = ctypes.c_int
    assert lib.get_int_size() == 4 and lib.get_dtype_size() == 4
    lib.sparknet_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.sparknet_version()
