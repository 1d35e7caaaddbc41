"""parse_net_prototxt / parse_solver_prototxt + get_prototxt_len / get_prototxt_data (reference
libccaffe/ccaffe.cpp:275-304, src/main/scala/libs/ProtoLoader.scala:9-29): the library's C++ text-format parser must
produce the message the Python codec of the reference schema produces.  No GPU involved: these entry points only parse."""
import glob
import os

import pytest

from sparknet_b200 import zoo
from sparknet_b200.caffe_proto import Msg, decode, parse_text, to_text
from sparknet_b200.library import load_library
from sparknet_b200.proto_loader import load_net_prototxt, load_solver_prototxt

NETS = {"lenet": zoo.lenet_net, "cifar10_quick": zoo.cifar10_quick_net, "cifar10_full": zoo.cifar10_full_net,
        "caffenet": zoo.caffenet_net}


@pytest.fixture(scope="module")
def lib():
    return load_library()


@pytest.mark.parametrize("name", list(NETS))
def test_net_prototxt_round_trip_through_the_library(lib, tmp_path, name):
    net = NETS[name]()
    path = tmp_path / f"{name}.prototxt"
    path.write_text("# " + name + " written by the test\n" + to_text(net))
    got = load_net_prototxt(str(path), lib)              # parse_net_prototxt + get_prototxt_len/data, as ProtoLoader.scala does
    assert got == net
    assert got == load_net_prototxt(str(path))           # and the same message as the Python text parser


@pytest.mark.parametrize("name", list(NETS))
def test_solver_prototxt_with_embedded_net(lib, tmp_path, name):
    sol = zoo.solver_with_net(name, random_seed=1701, display=0)
    path = tmp_path / f"{name}_solver.prototxt"
    path.write_text(to_text(sol))
    assert load_solver_prototxt(str(path), lib) == sol


HAND_WRITTEN = r'''
name: "hand"   # trailing comment
state < phase: TEST level: 0x2 stage: 'a' stage: "b\n\"q\"" >
force_backward: true
layer {
  name: "conv"  type: "Convolution"
  bottom: "data", top: "conv";
  param { lr_mult: 1 decay_mult: 1. }
  param: { lr_mult: 2e0 decay_mult: 0 }
  include: { phase: 0 }
  convolution_param {
    num_output: 96  kernel_size: [11]  stride: 4
    weight_filler { type: "gaussian" std: 1e-2 }
    bias_filler { type: "constant" value: .5f }
  }
}
layer { name: "norm" type: "LRN" bottom: "conv" top: "norm"
  lrn_param { local_size: 5 alpha: 0.0001 beta: 0.75 norm_region: WITHIN_CHANNEL } }
layer { name: "pool" type: "Pooling" bottom: "norm" top: "pool" pooling_param { pool: 1 kernel_size: 3 stride: 2 } }
'''


def test_text_format_details(lib, tmp_path):
    path = tmp_path / "hand.prototxt"
    path.write_text(HAND_WRITTEN)
    got = load_net_prototxt(str(path), lib)
    assert got == parse_text("NetParameter", HAND_WRITTEN)
    assert got.name == "hand" and got.force_backward is True
    assert got.state.level == 2 and list(got.state.stage) == ["a", 'b\n"q"']
    conv, norm, pool = got.layer
    assert list(conv.convolution_param.kernel_size) == [11] and conv.convolution_param.stride == [4]
    assert [p.lr_mult for p in conv.param] == [1.0, 2.0]
    assert abs(conv.convolution_param.weight_filler.std - 0.01) < 1e-9 and conv.convolution_param.bias_filler.value == 0.5
    assert norm.lrn_param.norm_region == 1 and pool.pooling_param.pool == 1


def test_errors_return_false(lib, tmp_path, capfd):
    state = lib.create_state()
    try:
        assert not lib.parse_net_prototxt(state, str(tmp_path / "missing.prototxt").encode())
        bad = tmp_path / "bad.prototxt"
        bad.write_text('name: "x"\nlayer { name: "l" no_such_field: 3 }\n')
        assert not lib.parse_net_prototxt(state, str(bad).encode())
        assert lib.get_prototxt_len(state) == 0
        assert "no_such_field" in capfd.readouterr().err and True
        unterminated = tmp_path / "open.prototxt"
        unterminated.write_text('layer { name: "l" ')
        assert not lib.parse_net_prototxt(state, str(unterminated).encode())
    finally:
        lib.destroy_state(state)


REFERENCE = "/root/reference/caffe"
SHIPPED = sorted(glob.glob(REFERENCE + "/examples/mnist/lenet_solver.prototxt") +
                 glob.glob(REFERENCE + "/examples/cifar10/cifar10_*_solver.prototxt") +
                 glob.glob(REFERENCE + "/models/bvlc_reference_caffenet/solver.prototxt"))


SHIPPED_NETS = sorted(glob.glob(REFERENCE + "/examples/mnist/lenet_train_test.prototxt") +
                      glob.glob(REFERENCE + "/examples/cifar10/cifar10_*_train_test.prototxt") +
                      glob.glob(REFERENCE + "/models/bvlc_reference_caffenet/train_val.prototxt"))


@pytest.mark.skipif(not SHIPPED, reason="the reference tree is only mounted in the build container")
@pytest.mark.parametrize("path", SHIPPED)
def test_the_references_own_solver_prototxts(lib, path):
    assert load_solver_prototxt(path, lib) == load_solver_prototxt(path)


@pytest.mark.skipif(not SHIPPED_NETS, reason="the reference tree is only mounted in the build container")
@pytest.mark.parametrize("path", SHIPPED_NETS)
def test_the_references_own_net_prototxts(lib, path):
    net = load_net_prototxt(path, lib)
    assert net == load_net_prototxt(path) and len(net.layer) >= 9
