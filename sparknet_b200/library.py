"""ctypes binding of the libccaffe C ABI — a 1:1 mirror of the JNA interface in
src/main/java/libs/CaffeLibrary.java:8-76 (same names, argument order and types), plus the
`sparknet_*` extension symbols declared in include/ccaffe.h.

`load_library()` loads the product (sparknet_b200/csrc/libccaffe.so).  It fails loudly if the
shared object is missing; there is no CPU fallback.  The CPU oracle (oracle/liboracle_ccaffe.so)
exports the same ABI and can be loaded with the same function by tests."""
from __future__ import annotations

import ctypes
import os
from ctypes import CFUNCTYPE, POINTER, c_bool, c_char_p, c_float, c_int, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.path.join(HERE, "csrc", "libccaffe.so")

# CaffeLibrary.java:12-14  interface java_callback_t { void invoke(Pointer data, int batch_size, int num_dims, Pointer shape); }
java_callback_t = CFUNCTYPE(None, c_void_p, c_int, c_int, POINTER(c_int))

_SIGNATURES = {
    # name: (restype, argtypes)                                       CaffeLibrary.java line
    "init_logging": (None, [c_char_p, c_int]),                       # :16
    "set_basepath": (None, [c_char_p]),                              # :17
    "get_int_size": (c_int, []),                                     # :19
    "get_dtype_size": (c_int, []),                                   # :20
    "create_state": (c_void_p, []),                                  # :30
    "destroy_state": (None, [c_void_p]),                             # :31
    "load_net_from_protobuf": (None, [c_void_p, c_char_p, c_int]),   # :33
    "load_solver_from_protobuf": (None, [c_void_p, c_char_p, c_int]),  # :34
    "set_train_data_callback": (c_int, [c_void_p, c_int, java_callback_t]),  # :36
    "set_test_data_callback": (c_int, [c_void_p, c_int, java_callback_t]),   # :37
    "forward": (None, [c_void_p]),                                   # :39
    "backward": (None, [c_void_p]),                                  # :40
    "solver_step": (c_int, [c_void_p, c_int]),                       # :42
    "solver_test": (None, [c_void_p, c_int]),                        # :43
    "num_test_scores": (c_int, [c_void_p]),                          # :45
    "get_test_score": (c_float, [c_void_p, c_int]),                  # :46
    "num_layers": (c_int, [c_void_p]),                               # :48
    "layer_name": (c_char_p, [c_void_p, c_int]),                     # :49
    "num_layer_weights": (c_int, [c_void_p, c_int]),                 # :50
    "num_data_blobs": (c_int, [c_void_p]),                           # :52
    "get_data_blob": (c_void_p, [c_void_p, c_int]),                  # :53
    "get_data_blob_name": (c_char_p, [c_void_p, c_int]),             # :54
    "get_weight_blob": (c_void_p, [c_void_p, c_int, c_int]),         # :55
    "get_data": (POINTER(c_float), [c_void_p]),                      # :57
    "get_diff": (POINTER(c_float), [c_void_p]),                      # :58
    "get_num_axes": (c_int, [c_void_p]),                             # :60
    "get_axis_shape": (c_int, [c_void_p, c_int]),                    # :61
    "set_mode_cpu": (None, []),                                      # :63
    "set_mode_gpu": (None, []),                                      # :64
    "set_device": (None, [c_int]),                                   # :65
    "load_weights_from_file": (None, [c_void_p, c_char_p]),          # :67
    "save_weights_to_file": (None, [c_void_p, c_char_p]),            # :68
    "restore_solver_from_file": (None, [c_void_p, c_char_p]),        # :69
    "parse_net_prototxt": (c_bool, [c_void_p, c_char_p]),            # :71
    "parse_solver_prototxt": (c_bool, [c_void_p, c_char_p]),         # :72
    "get_prototxt_len": (c_int, [c_void_p]),                         # :74
    "get_prototxt_data": (c_void_p, [c_void_p]),                     # :75
}

# ABI superset (include/ccaffe.h, "extensions"); optional per library
_EXTENSIONS = {
    "sparknet_get_iter": (c_int, [c_void_p]),
    "sparknet_last_loss": (c_float, [c_void_p]),
    "sparknet_smoothed_loss": (c_float, [c_void_p]),
    "sparknet_inject_dropout_mask": (c_int, [c_void_p, c_int, POINTER(ctypes.c_uint), c_int]),
    "sparknet_set_precision": (c_int, [c_void_p, c_int]),
    "sparknet_version": (c_char_p, []),
    "sparknet_get_precision": (c_int, [c_void_p]),
    "sparknet_tc_available": (c_int, []),
    "sparknet_compute_stream": (c_void_p, [c_void_p]),
    "sparknet_param_arena": (c_int, [c_void_p, POINTER(c_void_p), POINTER(ctypes.c_long)]),
    "sparknet_sync": (None, [c_void_p]),
    "sparknet_num_params": (ctypes.c_long, [c_void_p]),
    "sparknet_get_weights_flat": (c_int, [c_void_p, POINTER(c_float), ctypes.c_long]),
    "sparknet_set_weights_flat": (c_int, [c_void_p, POINTER(c_float), ctypes.c_long]),
    "sparknet_set_resident_inputs": (c_int, [c_void_p, c_int]),
    "sparknet_kernel_launches": (ctypes.c_long, [c_void_p]),
    "sparknet_comm_init": (c_int, [c_void_p, c_int, c_int, c_char_p, c_int]),
    "sparknet_comm_unique_id": (c_int, [c_char_p, c_int]),
    "sparknet_average_weights": (c_int, [c_void_p]),
    "sparknet_last_average_ms": (c_float, [c_void_p]),
    "sparknet_average_replicas_device": (c_int, [POINTER(c_void_p), c_int, ctypes.c_long, c_void_p]),
    "sparknet_average_replicas_async": (c_int, [POINTER(c_void_p), c_int, ctypes.c_long, c_void_p]),
    "sparknet_set_graph_mode": (c_int, [c_void_p, c_int]),
    "sparknet_graph_replays": (ctypes.c_long, [c_void_p]),
    "sparknet_graph_programmatic_edges": (c_int, [c_void_p]),
    "sparknet_peer_handle_bytes": (c_int, []),
    "sparknet_peer_export": (c_int, [c_void_p, c_void_p, c_void_p]),
    "sparknet_peer_attach": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "sparknet_average_peers": (c_int, [c_void_p]),
    "sparknet_peer_status": (c_int, [c_void_p]),
    "sparknet_peer_timing": (c_int, [c_void_p, POINTER(c_float), POINTER(c_float)]),
    "sparknet_test_average_peers": (c_int, [c_int, ctypes.c_long, POINTER(c_float), POINTER(c_float), c_int]),
    "sparknet_timer_start": (None, [c_void_p]),
    "sparknet_timer_stop_ms": (c_float, [c_void_p]),
    "sparknet_profile_step": (c_int, [c_void_p, POINTER(c_float), POINTER(c_float), POINTER(c_float), c_int]),
    "sparknet_get_loss_history": (c_int, [c_void_p, POINTER(c_float), c_int]),
    "sparknet_profile_launches": (c_int, [c_void_p, POINTER(c_int), POINTER(c_int), c_int]),
    "sparknet_host_copy": (None, [c_void_p, c_void_p, ctypes.c_long, c_int]),
    "sparknet_snapshot": (c_int, [c_void_p, c_char_p, c_char_p]),
    "sparknet_get_history": (c_int, [c_void_p, c_int, POINTER(c_float), ctypes.c_long]),
    "sparknet_get_pool_mask": (c_int, [c_void_p, c_int, POINTER(c_int), ctypes.c_long]),
    "sparknet_set_train_data_callback_bytes": (c_int, [c_void_p, c_int, java_callback_t, c_int, POINTER(c_float), c_int, c_float]),
    "sparknet_input_stats": (c_int, [c_void_p, POINTER(ctypes.c_double)]),
    "sparknet_forward_range": (c_int, [c_void_p, c_int, c_int]),
    "sparknet_backward_range": (c_int, [c_void_p, c_int, c_int]),
    "sparknet_reserve_workspace": (ctypes.c_long, [ctypes.c_long]),
    "sparknet_param_copy_counts": (None, [c_void_p, POINTER(ctypes.c_long), POINTER(ctypes.c_long)]),
    "sparknet_numa_info": (c_int, [c_char_p, c_int]),
    "sparknet_host_widen": (None, [c_void_p, c_void_p, ctypes.c_long, c_int, c_int]),
    # oracle-only helpers
    "oracle_pool_mask": (POINTER(c_int), [c_void_p, c_int]),
    "oracle_history": (POINTER(c_float), [c_void_p, c_int]),
    "oracle_num_learnable": (c_int, [c_void_p]),
    "oracle_average": (None, [POINTER(POINTER(c_float)), c_int, ctypes.c_long, POINTER(c_float)]),
    "oracle_gemm": (None, [c_int, c_int, c_int, c_int, c_int, c_float, POINTER(c_float), POINTER(c_float), c_float, POINTER(c_float)]),
    "oracle_gemv": (None, [c_int, c_int, c_int, c_float, POINTER(c_float), POINTER(c_float), c_float, POINTER(c_float)]),
    "oracle_seed": (None, [ctypes.c_uint]),
    "oracle_set_threads": (None, [c_int]),
    "oracle_use_blas": (c_int, [c_char_p, c_char_p]),
}

ABI_SYMBOLS = tuple(_SIGNATURES)


def load_library(path: str | None = None, strict: bool = True) -> ctypes.CDLL:
    path = path or os.environ.get("SPARKNET_LIBCCAFFE", DEFAULT_LIB)
    if not os.path.exists(path):
        raise FileNotFoundError(
            f"{path} not found: build it first (python -c 'import __graft_entry__ as g; g.build()' "
            "or make -C sparknet_b200/csrc). There is no CPU fallback.")
    lib = ctypes.CDLL(path, mode=ctypes.RTLD_LOCAL)
    for name, (res, args) in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)      # AttributeError here == missing ABI symbol
        except AttributeError:
            if strict:
                raise
            continue                     # the CPU oracle implements the hot-path subset only
        fn.restype, fn.argtypes = res, args
    for name, (res, args) in _EXTENSIONS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.restype, fn.argtypes = res, args
    lib._path = path
    return lib
