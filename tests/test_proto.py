"""Host logic: the protobuf codec and the net zoo against the reference-generated fixtures."""
import json
import os

import pytest

from sparknet_b200 import zoo
from sparknet_b200.caffe_proto import ENUMS, SCHEMA, Msg, decode, encode, parse_text, to_text
from sparknet_b200.proto_loader import load_solver_prototxt_with_net, replace_data_layers

GOLD = os.path.join(os.path.dirname(__file__), "golden")
NETS = ["lenet", "cifar10_quick", "cifar10_full", "caffenet"]
TYPE_CODES = {"double": 1, "float": 2, "int64": 3, "uint64": 4, "int32": 5, "bool": 8, "string": 9,
              "uint32": 13}


def gold(name):
    return open(os.path.join(GOLD, name), "rb").read()


@pytest.mark.parametrize("net", NETS)
def test_decode_encode_roundtrip_is_byte_identical(net):
    raw = gold(f"{net}_solver.binpb")
    assert encode(decode("SolverParameter", raw)) == raw


@pytest.mark.parametrize("net", NETS)
def test_text_parser_matches_protobuf_text_format(net):
    txt = gold(f"{net}_solver.prototxt.txt").decode()
    assert encode(parse_text("SolverParameter", txt)) == gold(f"{net}_solver.binpb")
    again = parse_text("SolverParameter", to_text(parse_text("SolverParameter", txt)))
    assert encode(again) == gold(f"{net}_solver.binpb")


@pytest.mark.parametrize("net", NETS)
def test_zoo_reproduces_what_sparknet_sends(net):
    # fixture = reference prototxt -> replaceDataLayers -> loadSolverPrototxtWithNet, via the reference's own schema
    assert encode(zoo.solver_with_net(net)) == gold(f"{net}_solver.binpb")


def test_replace_data_layers_and_solver_embedding():
    # ProtoLoader.scala:50-57 and :31-43 on a net that still has its original two data layers
    net = decode("NetParameter", gold("cifar10_quick_net.binpb"))
    layers = list(net.layer)[2:]                      # drop the TEST pair -> "original" layout
    orig = Msg("NetParameter", name=net.name, layer=layers)
    out = replace_data_layers(orig, 100, 100, 3, 32, 32)
    assert encode(out) == gold("cifar10_quick_net.binpb")
    sol = zoo.SOLVERS["cifar10_quick"]()
    sol.snapshot = 4000
    sol.snapshot_prefix = "examples/cifar10/cifar10_quick"
    sol.net = "examples/cifar10/cifar10_quick_train_test.prototxt"
    assert encode(load_solver_prototxt_with_net(sol, out)) == gold("cifar10_quick_solver.binpb")


def test_schema_matches_reference_descriptor():
    ref = json.load(open(os.path.join(GOLD, "caffe_schema_ref.json")))
    for mname, fields in SCHEMA.items():
        rf = ref["messages"][mname]
        for fname, (num, typ, kind) in fields.items():
            assert fname in rf, (mname, fname)
            rnum, rtype, rlabel, rtname, _, rpacked = rf[fname]
            assert rnum == num, (mname, fname)
            assert (rlabel == 3) == (kind != 0), (mname, fname)
            assert rpacked == (kind == 2), (mname, fname)
            if typ in SCHEMA:
                assert rtype == 11 and rtname.endswith("." + typ), (mname, fname)
            elif typ.startswith("enum:"):
                assert rtype == 14, (mname, fname)
                en = typ[5:]
                cands = [v for k, v in ref["enums"].items() if k.endswith(en)]
                assert ENUMS[en] in cands, (mname, fname)
            else:
                assert TYPE_CODES[typ] == rtype, (mname, fname)


def test_negative_and_packed_scalars():
    s = Msg("SolverParameter", random_seed=-1, test_iter=[1, 2, 300])
    d = decode("SolverParameter", encode(s))
    assert d.random_seed == -1 and d.test_iter == [1, 2, 300]
    b = Msg("BlobProto", data=[1.5, -2.0], shape=Msg("BlobShape", dim=[2]))
    d = decode("BlobProto", encode(b))
    assert d.data == [1.5, -2.0] and d.shape.dim == [2]
