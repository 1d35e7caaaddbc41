"""Dependency-free codec for the Caffe protobuf messages that cross the libccaffe boundary.

SparkNet hands `load_solver_from_protobuf` a serialized `SolverParameter` that embeds the
`NetParameter` (reference: src/main/scala/libs/Net.scala:253-262,
libccaffe/ccaffe.cpp:128-134).  The JVM side builds those bytes with protobuf-java; this module
is the host-side mirror used by the Python driver, the tests and bench.py: a hand-written
schema table (field numbers from caffe/src/caffe/proto/caffe.proto, cross-checked in
tests/test_proto.py against the schema embedded in the reference's generated Caffe.java), a
prototxt (text-format) parser, and a proto2 wire encoder/decoder whose output is byte-identical
to protobuf's own serializer for these messages (checked against tests/golden/*.binpb).

Nothing here touches the GPU; the C++ library has its own decoder (csrc/caffe_proto.hpp).
"""
from __future__ import annotations

import re
import struct
from typing import Any, Dict, Iterator, List, Tuple

# field kinds
OPT, REP, PACKED = 0, 1, 2

# ---------------------------------------------------------------------------------------------
# Schema: message -> field name -> (number, type, kind).  `type` is a scalar type name, an
# "enum:<Name>" or a message name.  Source: caffe.proto (line numbers in the reference tree).
# ---------------------------------------------------------------------------------------------
ENUMS: Dict[str, Dict[str, int]] = {
    "Phase": {"TRAIN": 0, "TEST": 1},                                   # caffe.proto:257
    "SolverMode": {"CPU": 0, "GPU": 1},                                 # :190
    "SnapshotFormat": {"HDF5": 0, "BINARYPROTO": 1},                    # :183
    "SolverType": {"SGD": 0, "NESTEROV": 1, "ADAGRAD": 2, "RMSPROP": 3, "ADADELTA": 4, "ADAM": 5},
    "VarianceNorm": {"FAN_IN": 0, "FAN_OUT": 1, "AVERAGE": 2},          # :56
    "DimCheckMode": {"STRICT": 0, "PERMISSIVE": 1},                     # :287
    "Engine": {"DEFAULT": 0, "CAFFE": 1, "CUDNN": 2},
    "PoolMethod": {"MAX": 0, "AVE": 1, "STOCHASTIC": 2},                # :775
    "NormRegion": {"ACROSS_CHANNELS": 0, "WITHIN_CHANNEL": 1},          # :742
    "EltwiseOp": {"PROD": 0, "SUM": 1, "MAX": 2},                       # :603
    "DB": {"LEVELDB": 0, "LMDB": 1},                                    # DataParameter.DB
}

SCHEMA: Dict[str, Dict[str, Tuple[int, str, int]]] = {
    "BlobShape": {"dim": (1, "int64", PACKED)},
    "BlobProto": {
        "num": (1, "int32", OPT), "channels": (2, "int32", OPT), "height": (3, "int32", OPT),
        "width": (4, "int32", OPT), "data": (5, "float", PACKED), "diff": (6, "float", PACKED),
        "shape": (7, "BlobShape", OPT), "double_data": (8, "double", PACKED),
        "double_diff": (9, "double", PACKED),
    },
    "FillerParameter": {
        "type": (1, "string", OPT), "value": (2, "float", OPT), "min": (3, "float", OPT),
        "max": (4, "float", OPT), "mean": (5, "float", OPT), "std": (6, "float", OPT),
        "sparse": (7, "int32", OPT), "variance_norm": (8, "enum:VarianceNorm", OPT),
    },
    "NetState": {"phase": (1, "enum:Phase", OPT), "level": (2, "int32", OPT), "stage": (3, "string", REP)},
    "NetStateRule": {
        "phase": (1, "enum:Phase", OPT), "min_level": (2, "int32", OPT), "max_level": (3, "int32", OPT),
        "stage": (4, "string", REP), "not_stage": (5, "string", REP),
    },
    "ParamSpec": {
        "name": (1, "string", OPT), "share_mode": (2, "enum:DimCheckMode", OPT),
        "lr_mult": (3, "float", OPT), "decay_mult": (4, "float", OPT),
    },
    "NetParameter": {
        "name": (1, "string", OPT), "input": (3, "string", REP), "input_dim": (4, "int32", REP),
        "force_backward": (5, "bool", OPT), "state": (6, "NetState", OPT), "debug_info": (7, "bool", OPT),
        "input_shape": (8, "BlobShape", REP), "layer": (100, "LayerParameter", REP),
    },
    "SolverParameter": {
        "train_net": (1, "string", OPT), "test_net": (2, "string", REP), "test_iter": (3, "int32", REP),
        "test_interval": (4, "int32", OPT), "base_lr": (5, "float", OPT), "display": (6, "int32", OPT),
        "max_iter": (7, "int32", OPT), "lr_policy": (8, "string", OPT), "gamma": (9, "float", OPT),
        "power": (10, "float", OPT), "momentum": (11, "float", OPT), "weight_decay": (12, "float", OPT),
        "stepsize": (13, "int32", OPT), "snapshot": (14, "int32", OPT), "snapshot_prefix": (15, "string", OPT),
        "snapshot_diff": (16, "bool", OPT), "solver_mode": (17, "enum:SolverMode", OPT),
        "device_id": (18, "int32", OPT), "test_compute_loss": (19, "bool", OPT),
        "random_seed": (20, "int64", OPT), "train_net_param": (21, "NetParameter", OPT),
        "test_net_param": (22, "NetParameter", REP), "debug_info": (23, "bool", OPT),
        "net": (24, "string", OPT), "net_param": (25, "NetParameter", OPT),
        "train_state": (26, "NetState", OPT), "test_state": (27, "NetState", REP),
        "average_loss": (33, "int32", OPT), "stepvalue": (34, "int32", REP),
        "regularization_type": (29, "string", OPT), "solver_type": (30, "enum:SolverType", OPT),
        "delta": (31, "float", OPT), "test_initialization": (32, "bool", OPT),
        "clip_gradients": (35, "float", OPT), "iter_size": (36, "int32", OPT),
        "snapshot_format": (37, "enum:SnapshotFormat", OPT), "rms_decay": (38, "float", OPT),
        "momentum2": (39, "float", OPT), "type": (40, "string", OPT),
        "snapshot_after_train": (28, "bool", OPT),
    },
    "SolverState": {
        "iter": (1, "int32", OPT), "learned_net": (2, "string", OPT), "history": (3, "BlobProto", REP),
        "current_step": (4, "int32", OPT),
    },
    "LayerParameter": {
        "name": (1, "string", OPT), "type": (2, "string", OPT), "bottom": (3, "string", REP),
        "top": (4, "string", REP), "loss_weight": (5, "float", REP), "param": (6, "ParamSpec", REP),
        "blobs": (7, "BlobProto", REP), "include": (8, "NetStateRule", REP),
        "exclude": (9, "NetStateRule", REP), "phase": (10, "enum:Phase", OPT),
        "propagate_down": (11, "bool", REP),
        "transform_param": (100, "TransformationParameter", OPT),
        "loss_param": (101, "LossParameter", OPT), "accuracy_param": (102, "AccuracyParameter", OPT),
        "convolution_param": (106, "ConvolutionParameter", OPT), "data_param": (107, "DataParameter", OPT),
        "dropout_param": (108, "DropoutParameter", OPT), "eltwise_param": (110, "EltwiseParameter", OPT),
        "inner_product_param": (117, "InnerProductParameter", OPT), "lrn_param": (118, "LRNParameter", OPT),
        "pooling_param": (121, "PoolingParameter", OPT), "power_param": (122, "PowerParameter", OPT),
        "relu_param": (123, "ReLUParameter", OPT), "softmax_param": (125, "SoftmaxParameter", OPT),
        "java_data_param": (149, "JavaDataParameter", OPT),
    },
    "TransformationParameter": {
        "scale": (1, "float", OPT), "mirror": (2, "bool", OPT), "crop_size": (3, "uint32", OPT),
        "mean_file": (4, "string", OPT), "mean_value": (5, "float", REP),
        "force_color": (6, "bool", OPT), "force_gray": (7, "bool", OPT),
    },
    "DataParameter": {
        "source": (1, "string", OPT), "batch_size": (4, "uint32", OPT), "rand_skip": (7, "uint32", OPT),
        "backend": (8, "enum:DB", OPT), "scale": (2, "float", OPT), "mean_file": (3, "string", OPT),
        "crop_size": (5, "uint32", OPT), "mirror": (6, "bool", OPT),
        "force_encoded_color": (9, "bool", OPT), "prefetch": (10, "uint32", OPT),
    },
    "LossParameter": {"ignore_label": (1, "int32", OPT), "normalize": (2, "bool", OPT)},
    "AccuracyParameter": {"top_k": (1, "uint32", OPT), "axis": (2, "int32", OPT), "ignore_label": (3, "int32", OPT)},
    "ConvolutionParameter": {
        "num_output": (1, "uint32", OPT), "bias_term": (2, "bool", OPT), "pad": (3, "uint32", REP),
        "kernel_size": (4, "uint32", REP), "group": (5, "uint32", OPT), "stride": (6, "uint32", REP),
        "weight_filler": (7, "FillerParameter", OPT), "bias_filler": (8, "FillerParameter", OPT),
        "pad_h": (9, "uint32", OPT), "pad_w": (10, "uint32", OPT), "kernel_h": (11, "uint32", OPT),
        "kernel_w": (12, "uint32", OPT), "stride_h": (13, "uint32", OPT), "stride_w": (14, "uint32", OPT),
        "engine": (15, "enum:Engine", OPT), "axis": (16, "int32", OPT), "force_nd_im2col": (17, "bool", OPT),
    },
    "DropoutParameter": {"dropout_ratio": (1, "float", OPT)},
    "EltwiseParameter": {
        "operation": (1, "enum:EltwiseOp", OPT), "coeff": (2, "float", REP), "stable_prod_grad": (3, "bool", OPT),
    },
    "InnerProductParameter": {
        "num_output": (1, "uint32", OPT), "bias_term": (2, "bool", OPT),
        "weight_filler": (3, "FillerParameter", OPT), "bias_filler": (4, "FillerParameter", OPT),
        "axis": (5, "int32", OPT),
    },
    "LRNParameter": {
        "local_size": (1, "uint32", OPT), "alpha": (2, "float", OPT), "beta": (3, "float", OPT),
        "norm_region": (4, "enum:NormRegion", OPT), "k": (5, "float", OPT), "engine": (6, "enum:Engine", OPT),
    },
    "PoolingParameter": {
        "pool": (1, "enum:PoolMethod", OPT), "pad": (4, "uint32", OPT), "pad_h": (9, "uint32", OPT),
        "pad_w": (10, "uint32", OPT), "kernel_size": (2, "uint32", OPT), "kernel_h": (5, "uint32", OPT),
        "kernel_w": (6, "uint32", OPT), "stride": (3, "uint32", OPT), "stride_h": (7, "uint32", OPT),
        "stride_w": (8, "uint32", OPT), "engine": (11, "enum:Engine", OPT), "global_pooling": (12, "bool", OPT),
    },
    "PowerParameter": {"power": (1, "float", OPT), "scale": (2, "float", OPT), "shift": (3, "float", OPT)},
    "ReLUParameter": {"negative_slope": (1, "float", OPT), "engine": (2, "enum:Engine", OPT)},
    "SoftmaxParameter": {"engine": (1, "enum:Engine", OPT), "axis": (2, "int32", OPT)},
    "JavaDataParameter": {"shape": (1, "BlobShape", OPT)},             # SparkNet addition, caffe.proto:991-993
}

_VARINT_TYPES = {"int32", "int64", "uint32", "uint64", "bool"}


class Msg:
    """A protobuf message instance: ordered dict of explicitly-set fields."""

    __slots__ = ("_type", "_f")

    def __init__(self, type_name: str, **kw: Any):
        if type_name not in SCHEMA:
            raise KeyError(f"unknown message type {type_name}")
        object.__setattr__(self, "_type", type_name)
        object.__setattr__(self, "_f", {})
        for k, v in kw.items():
            setattr(self, k, v)

    # -- attribute access ---------------------------------------------------------------------
    def _spec(self, name: str) -> Tuple[int, str, int]:
        try:
            return SCHEMA[self._type][name]
        except KeyError:
            raise AttributeError(f"{self._type} has no field {name!r}") from None

    def __getattr__(self, name: str) -> Any:
        num, typ, kind = self._spec(name)
        if name in self._f:
            return self._f[name]
        if kind != OPT:
            lst: List[Any] = []
            self._f[name] = lst
            return lst
        if typ in SCHEMA:                       # auto-vivify sub-message like protobuf's mutable access
            sub = Msg(typ)
            self._f[name] = sub
            return sub
        return None

    def __setattr__(self, name: str, value: Any) -> None:
        num, typ, kind = self._spec(name)
        if kind != OPT:
            value = list(value)
        self._f[name] = value

    def has(self, name: str) -> bool:
        self._spec(name)
        v = self._f.get(name)
        return v is not None and (not isinstance(v, list) or len(v) > 0)

    def clear(self, name: str) -> None:
        self._spec(name)
        self._f.pop(name, None)

    def get(self, name: str, default: Any = None) -> Any:
        return self._f[name] if self.has(name) else default

    @property
    def type_name(self) -> str:
        return self._type

    def fields(self) -> Iterator[Tuple[str, Any]]:
        sch = SCHEMA[self._type]
        for name in sorted((n for n in self._f if self.has(n)), key=lambda n: sch[n][0]):
            yield name, self._f[name]

    def copy(self) -> "Msg":
        return decode(self._type, encode(self))

    def __eq__(self, other: object) -> bool:
        return isinstance(other, Msg) and other._type == self._type and encode(self) == encode(other)

    def __repr__(self) -> str:
        return f"<{self._type} {to_text(self, indent=None)}>"


# ---------------------------------------------------------------------------------------------
# wire encoding (proto2, fields in number order, like protobuf's serializers)
# ---------------------------------------------------------------------------------------------
def _varint(v: int) -> bytes:
    if v < 0:
        v += 1 << 64
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _scalar_payload(typ: str, v: Any) -> Tuple[int, bytes]:
    if typ in _VARINT_TYPES:
        return 0, _varint(int(v))
    if typ.startswith("enum:"):
        if isinstance(v, str):
            v = ENUMS[typ[5:]][v]
        return 0, _varint(int(v))
    if typ == "float":
        return 5, struct.pack("<f", float(v))
    if typ == "double":
        return 1, struct.pack("<d", float(v))
    if typ == "string":
        b = v.encode("utf-8") if isinstance(v, str) else bytes(v)
        return 2, _varint(len(b)) + b
    raise TypeError(typ)


def encode(msg: Msg) -> bytes:
    out = bytearray()
    sch = SCHEMA[msg.type_name]
    for name, val in msg.fields():
        num, typ, kind = sch[name]
        if typ in SCHEMA:
            for sub in (val if kind != OPT else [val]):
                b = encode(sub)
                out += _varint((num << 3) | 2) + _varint(len(b)) + b
        elif kind == PACKED:
            body = b"".join(_scalar_payload(typ, x)[1] for x in val)
            out += _varint((num << 3) | 2) + _varint(len(body)) + body
        else:
            for x in (val if kind == REP else [val]):
                wt, payload = _scalar_payload(typ, x)
                out += _varint((num << 3) | wt) + payload
    return bytes(out)


def _read_varint(b: bytes, i: int) -> Tuple[int, int]:
    shift = 0
    v = 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7F) << shift
        if not c & 0x80:
            return v, i
        shift += 7


def _from_varint(typ: str, v: int) -> Any:
    if typ == "bool":
        return bool(v)
    if typ in ("int32", "int64") or typ.startswith("enum:"):
        if v >= 1 << 63:
            v -= 1 << 64
        return v
    return v


def decode(type_name: str, data: bytes) -> Msg:
    msg = Msg(type_name)
    by_num = {spec[0]: (name, spec) for name, spec in SCHEMA[type_name].items()}
    i, n = 0, len(data)
    while i < n:
        key, i = _read_varint(data, i)
        num, wt = key >> 3, key & 7
        ent = by_num.get(num)
        if wt == 0:
            raw, i = _read_varint(data, i)
            payload: Any = raw
        elif wt == 1:
            payload, i = data[i:i + 8], i + 8
        elif wt == 5:
            payload, i = data[i:i + 4], i + 4
        elif wt == 2:
            ln, i = _read_varint(data, i)
            payload, i = data[i:i + ln], i + ln
        else:
            raise ValueError(f"unsupported wire type {wt}")
        if ent is None:
            continue                               # unknown field: skipped (the C++ decoder does the same)
        name, (_, typ, kind) = ent
        vals: List[Any]
        if typ in SCHEMA:
            vals = [decode(typ, payload)]
        elif typ == "string":
            vals = [payload.decode("utf-8")]
        elif wt == 2:                              # packed scalars
            vals, j = [], 0
            while j < len(payload):
                if typ == "float":
                    vals.append(struct.unpack_from("<f", payload, j)[0]); j += 4
                elif typ == "double":
                    vals.append(struct.unpack_from("<d", payload, j)[0]); j += 8
                else:
                    raw, j = _read_varint(payload, j)
                    vals.append(_from_varint(typ, raw))
        elif typ == "float":
            vals = [struct.unpack("<f", payload)[0]]
        elif typ == "double":
            vals = [struct.unpack("<d", payload)[0]]
        else:
            vals = [_from_varint(typ, payload)]
        if kind == OPT:
            msg._f[name] = vals[-1]
        else:
            msg._f.setdefault(name, []).extend(vals)
    return msg


# ---------------------------------------------------------------------------------------------
# text format (prototxt)
# ---------------------------------------------------------------------------------------------
_TOKEN = re.compile(r"""\s*(?:(\#[^\n]*)|([{}:<>\[\],;])|("(?:[^"\\]|\\.)*"|'(?:[^'\\]|\\.)*')|([^\s{}:<>\[\],;"'#]+))""")


def _tokenize(text: str) -> List[str]:
    toks, i = [], 0
    while i < len(text):
        m = _TOKEN.match(text, i)
        if not m:
            if text[i:].strip() == "":
                break
            raise ValueError(f"prototxt: cannot tokenize at offset {i}: {text[i:i+30]!r}")
        i = m.end()
        if m.group(1):
            continue
        toks.append(m.group(2) or m.group(3) or m.group(4))
    return toks


def _unquote(tok: str) -> str:
    body = tok[1:-1]
    return bytes(body, "utf-8").decode("unicode_escape") if "\\" in body else body


def _parse_scalar(typ: str, tok: str) -> Any:
    if typ == "string":
        return _unquote(tok)
    if typ == "bool":
        if tok in ("true", "True", "1", "t"):
            return True
        if tok in ("false", "False", "0", "f"):
            return False
        raise ValueError(f"bad bool {tok}")
    if typ.startswith("enum:"):
        table = ENUMS[typ[5:]]
        return table[tok] if tok in table else int(tok)
    if typ in ("float", "double"):
        return float(tok.rstrip("fF")) if tok[-1:] in "fF" and tok[:-1].replace(".", "").lstrip("+-").isdigit() else float(tok)
    return int(tok, 0)


def parse_text(type_name: str, text: str) -> Msg:
    toks = _tokenize(text)
    pos = 0

    def parse_msg(tname: str, closer: str | None) -> Msg:
        nonlocal pos
        msg = Msg(tname)
        sch = SCHEMA[tname]
        while pos < len(toks):
            t = toks[pos]
            if closer is not None and t == closer:
                pos += 1
                return msg
            if t in (",", ";"):
                pos += 1
                continue
            name = t
            pos += 1
            if name not in sch:
                raise ValueError(f"prototxt: {tname} has no field {name!r} (not on the supported path)")
            num, typ, kind = sch[name]
            if toks[pos] == ":":
                pos += 1
            if typ in SCHEMA:
                opener = toks[pos]
                if opener not in ("{", "<"):
                    raise ValueError(f"prototxt: expected '{{' after {name}")
                pos += 1
                val: Any = parse_msg(typ, "}" if opener == "{" else ">")
                vals = [val]
            elif toks[pos] == "[":
                pos += 1
                vals = []
                while toks[pos] != "]":
                    if toks[pos] != ",":
                        vals.append(_parse_scalar(typ, toks[pos]))
                    pos += 1
                pos += 1
            else:
                vals = [_parse_scalar(typ, toks[pos])]
                pos += 1
            if kind == OPT:
                msg._f[name] = vals[-1]
            else:
                msg._f.setdefault(name, []).extend(vals)
        if closer is not None:
            raise ValueError("prototxt: unexpected end of input")
        return msg

    return parse_msg(type_name, None)


def _fmt_scalar(typ: str, v: Any) -> str:
    if typ == "string":
        return '"' + str(v).replace("\\", "\\\\").replace('"', '\\"') + '"'
    if typ == "bool":
        return "true" if v else "false"
    if typ.startswith("enum:"):
        for k, n in ENUMS[typ[5:]].items():
            if n == v or k == v:
                return k
        return str(v)
    if typ in ("float", "double"):
        return repr(float(v))
    return str(int(v))


def to_text(msg: Msg, indent: int | None = 0) -> str:
    lines: List[str] = []
    pad = "" if indent is None else "  " * indent
    sch = SCHEMA[msg.type_name]
    for name, val in msg.fields():
        num, typ, kind = sch[name]
        for x in (val if kind != OPT else [val]):
            if typ in SCHEMA:
                if indent is None:
                    lines.append(f"{name} {{ {to_text(x, None)} }}")
                else:
                    lines.append(f"{pad}{name} {{\n{to_text(x, indent + 1)}{pad}}}")
            else:
                lines.append(f"{pad}{name}: {_fmt_scalar(typ, x)}")
    if indent is None:
        return " ".join(lines)
    return "".join(l + "\n" for l in lines)
