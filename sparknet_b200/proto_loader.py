"""Host-side mirror of src/main/scala/libs/ProtoLoader.scala.

`load_*_prototxt` use this package's own text-format parser when no library handle is given,
or the library's `parse_*_prototxt` + `get_prototxt_{len,data}` entry points exactly as the
Scala code does (ProtoLoader.scala:9-29)."""
from __future__ import annotations

import ctypes
from typing import Optional

from .caffe_proto import Msg, decode, parse_text
from .layers import TEST, TRAIN, RDDLayer


def _via_library(lib, fn_name: str, filename: str) -> bytes:
    state = lib.create_state()
    try:
        ok = getattr(lib, fn_name)(state, filename.encode())
        if not ok:
            raise IOError(f"{fn_name}({filename}) failed")
        n = lib.get_prototxt_len(state)
        return ctypes.string_at(lib.get_prototxt_data(state), n)
    finally:
        lib.destroy_state(state)


def load_solver_prototxt(filename: str, lib=None) -> Msg:          # ProtoLoader.scala:9-18
    if lib is not None:
        return decode("SolverParameter", _via_library(lib, "parse_solver_prototxt", filename))
    return parse_text("SolverParameter", open(filename).read())


def load_net_prototxt(filename: str, lib=None) -> Msg:             # ProtoLoader.scala:20-29
    if lib is not None:
        return decode("NetParameter", _via_library(lib, "parse_net_prototxt", filename))
    return parse_text("NetParameter", open(filename).read())


def load_solver_prototxt_with_net(solver: Msg, net: Msg, snapshot_path: Optional[str] = None) -> Msg:
    """ProtoLoader.scala:31-43 (takes an already-parsed SolverParameter)."""
    sol = solver.copy()
    if snapshot_path is None:
        sol.clear("snapshot")
        sol.clear("snapshot_prefix")
    else:
        sol.snapshot_prefix = snapshot_path
    sol.clear("net")
    sol.net_param = net.copy()
    return sol


def replace_data_layers(net: Msg, train_batch: int, test_batch: int, channels: int, height: int, width: int) -> Msg:
    """ProtoLoader.scala:50-57: layers 0,1 become TRAIN JavaData layers; a TEST pair is prepended."""
    out = net.copy()
    layers = list(out.layer)
    layers[0] = RDDLayer("data", [train_batch, channels, height, width], TRAIN)
    layers[1] = RDDLayer("label", [train_batch, 1], TRAIN)
    layers = [RDDLayer("data", [test_batch, channels, height, width], TEST),
              RDDLayer("label", [test_batch, 1], TEST)] + layers
    out.layer = layers
    return out
