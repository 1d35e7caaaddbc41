#!/usr/bin/env python
"""Generate tests/golden/*.binpb from the REFERENCE (run in the build container only).

The reference cannot be compiled here (no protoc/boost/glog), but its
generated Java class embeds the serialized FileDescriptorProto of caffe.proto
(/root/reference/src/main/java/caffe/Caffe.java, `descriptorData`).  We un-escape that Java
string, hand it to google.protobuf, and obtain the reference's *own* message classes.  With
those we parse the reference's prototxts, apply the two SparkNet transformations
(ProtoLoader.scala:31-43 loadSolverPrototxtWithNet, :50-57 replaceDataLayers) and write the
exact SolverParameter bytes SparkNet would hand to load_solver_from_protobuf.

Outputs (small, committed):
  tests/golden/<net>_solver.binpb   SolverParameter with embedded net_param
  tests/golden/<net>_net.binpb      NetParameter after replaceDataLayers
  tests/golden/<net>_net.prototxt.txt  text_format dump of the same (for the text parser test)
  tests/golden/manifest.json        shapes / batch sizes used
"""
import json, os, re, sys

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def java_unescape(s: str) -> bytes:
    out = bytearray()
    i = 0
    while i < len(s):
        ch = s[i]
        if ch != "\\":
            out += ch.encode("latin-1") if ord(ch) < 256 else ch.encode("utf-8")
            i += 1
            continue
        i += 1
        ch = s[i]
        if ch in "01234567":
            # Java octal escape: first digit 0-3 => up to 3 digits, else up to 2
            maxd = 3 if ch in "0123" else 2
            j = i
            while j < len(s) and j - i < maxd and s[j] in "01234567":
                j += 1
            out.append(int(s[i:j], 8))
            i = j
        else:
            out += {"n": b"\n", "t": b"\t", "r": b"\r", '"': b'"', "'": b"'", "\\": b"\\",
                    "b": b"\b", "f": b"\f"}[ch]
            i += 1
    return bytes(out)


def load_caffe_pb():
    src = open(os.path.join(REF, "src/main/java/caffe/Caffe.java"), encoding="latin-1").read()
    start = src.index("java.lang.String[] descriptorData = {")
    end = src.index("};", start)
    body = src[start:end]
    pieces = re.findall(r'"((?:[^"\\]|\\.)*)"', body)
    raw = b"".join(java_unescape(p) for p in pieces)
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.ParseFromString(raw)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName("caffe." + n))
    return get, fdp


def rdd_layer(LayerParameter, name, shape, phase):
    # Layers.scala:19-41 RDDLayer
    l = LayerParameter()
    l.type = "JavaData"
    l.name = name
    l.top.append(name)
    l.java_data_param.shape.dim.extend(shape)
    r = l.include.add()
    r.phase = phase
    return l


def replace_data_layers(net, LayerParameter, train_bs, test_bs, c, h, w):
    # ProtoLoader.scala:50-57
    net.layer[0].CopyFrom(rdd_layer(LayerParameter, "data", [train_bs, c, h, w], 0))
    net.layer[1].CopyFrom(rdd_layer(LayerParameter, "label", [train_bs, 1], 0))
    rest = list(net.layer)
    del net.layer[:]
    net.layer.add().CopyFrom(rdd_layer(LayerParameter, "data", [test_bs, c, h, w], 1))
    net.layer.add().CopyFrom(rdd_layer(LayerParameter, "label", [test_bs, 1], 1))
    for l in rest:
        net.layer.add().CopyFrom(l)
    return net


NETS = {
    # name: (net prototxt, solver prototxt, train_bs, test_bs, C, H, W)
    "lenet": ("caffe/examples/mnist/lenet_train_test.prototxt",
              "caffe/examples/mnist/lenet_solver.prototxt", 64, 100, 1, 28, 28),
    "cifar10_quick": ("caffe/examples/cifar10/cifar10_quick_train_test.prototxt",
                      "caffe/examples/cifar10/cifar10_quick_solver.prototxt", 100, 100, 3, 32, 32),
    "cifar10_full": ("caffe/examples/cifar10/cifar10_full_train_test.prototxt",
                     "caffe/examples/cifar10/cifar10_full_solver.prototxt", 100, 100, 3, 32, 32),
    "caffenet": ("caffe/models/bvlc_reference_caffenet/train_val.prototxt",
                 "caffe/models/bvlc_reference_caffenet/solver.prototxt", 256, 50, 3, 227, 227),
}


def main():
    from google.protobuf import text_format
    get, fdp = load_caffe_pb()
    NetParameter, SolverParameter, LayerParameter = get("NetParameter"), get("SolverParameter"), get("LayerParameter")
    os.makedirs(OUT, exist_ok=True)
    manifest = {}
    for name, (netf, solf, trb, teb, c, h, w) in NETS.items():
        net = NetParameter()
        text_format.Parse(open(os.path.join(REF, netf)).read(), net)
        net = replace_data_layers(net, LayerParameter, trb, teb, c, h, w)
        sol = SolverParameter()
        text_format.Parse(open(os.path.join(REF, solf)).read(), sol)
        # ProtoLoader.scala:31-43 (snapshotPath == None)
        sol.ClearField("snapshot")
        sol.ClearField("snapshot_prefix")
        sol.ClearField("net")
        sol.net_param.CopyFrom(net)
        open(os.path.join(OUT, name + "_solver.binpb"), "wb").write(sol.SerializeToString())
        open(os.path.join(OUT, name + "_net.binpb"), "wb").write(net.SerializeToString())
        open(os.path.join(OUT, name + "_net.prototxt.txt"), "w").write(text_format.MessageToString(net))
        open(os.path.join(OUT, name + "_solver.prototxt.txt"), "w").write(text_format.MessageToString(sol))
        manifest[name] = dict(train_batch=trb, test_batch=teb, channels=c, height=h, width=w,
                              n_layers_total=len(net.layer), solver_bytes=len(sol.SerializeToString()))
        print(name, manifest[name])
    json.dump(manifest, open(os.path.join(OUT, "manifest.json"), "w"), indent=1, sort_keys=True)
    # field-number table of every message, used to cross-check sparknet_b200/caffe_proto.py
    table = {}
    def walk(msgs, prefix=""):
        for m in msgs:
            table[prefix + m.name] = {f.name: [f.number, f.type, f.label, f.type_name, f.default_value,
                                                bool(f.options.packed)] for f in m.field}
            walk(m.nested_type, prefix + m.name + ".")
    walk(fdp.message_type)
    enums = {}
    def walk_e(msgs, prefix=""):
        for m in msgs:
            for e in m.enum_type:
                enums[prefix + m.name + "." + e.name] = {v.name: v.number for v in e.value}
            walk_e(m.nested_type, prefix + m.name + ".")
    walk_e(fdp.message_type)
    for e in fdp.enum_type:
        enums[e.name] = {v.name: v.number for v in e.value}
    json.dump({"messages": table, "enums": enums}, open(os.path.join(OUT, "caffe_schema_ref.json"), "w"),
              indent=0, sort_keys=True)
    print("messages:", len(table))


if __name__ == "__main__":
    main()
