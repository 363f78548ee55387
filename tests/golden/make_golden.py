#!/usr/bin/env python
"""Generate tests/golden/upb_vectors.json with an INDEPENDENT protobuf implementation (upb, the C
backend of python `protobuf`, with tensorboard's compiled tensorflow.TensorProto).

Why upb: the reference's own arithmetic lives in prost 0.6.0, which is neither vendored under
/root/reference nor buildable here (no rustc/cargo/protoc), and the reference has no tests or golden
vectors for this path.  upb agrees with prost on every canonical encoding; the
known divergences (default map key/value, version_label ordering, map order)
are avoided here and covered by hand-written vectors in tests/test_oracle.py instead.

PredictRequest / PredictResponse / ModelSpec are not shipped compiled anywhere in this image, so their
descriptors are rebuilt at run time from the field numbers of the reference's
protos/tensorflow_serving/apis/predict.proto:12-40 and model.proto:9-33.

Run (CPU container only; needs tensorboard + protobuf):  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory, wrappers_pb2  # noqa: F401
from tensorboard.compat.proto import tensor_pb2, tensor_shape_pb2, types_pb2  # noqa: F401

HERE = os.path.dirname(os.path.abspath(__file__))

FIELD_NP = {
    "tensor_content": "u1", "float_val": "<f4", "double_val": "<f8", "int_val": "<i4", "scomplex_val": "<f4",
    "int64_val": "<i8", "bool_val": "u1", "dcomplex_val": "<f8", "half_val": "<i4", "uint32_val": "<u4",
    "uint64_val": "<u8",
}


def build_predict_classes():
    pool = descriptor_pool.Default()
    tensor_file = tensor_pb2.DESCRIPTOR.name
    f = descriptor_pb2.FileDescriptorProto()
    f.name = "golden/tfs_model.proto"
    f.package = "tensorflow.serving"
    f.syntax = "proto3"
    f.dependency.append("google/protobuf/wrappers.proto")
    m = f.message_type.add()
    m.name = "ModelSpec"
    o = m.oneof_decl.add()
    o.name = "version_choice"
    for name, num, typ, tname, oneof in (("name", 1, 9, None, None), ("version", 2, 11, ".google.protobuf.Int64Value", 0),
                                         ("signature_name", 3, 9, None, None), ("version_label", 4, 9, None, 0)):
        fd = m.field.add()
        fd.name, fd.number, fd.type, fd.label = name, num, typ, 1
        if tname:
            fd.type_name = tname
        if oneof is not None:
            fd.oneof_index = oneof
    pool.Add(f)

    g = descriptor_pb2.FileDescriptorProto()
    g.name = "golden/tfs_predict.proto"
    g.package = "tensorflow.serving"
    g.syntax = "proto3"
    g.dependency.extend([tensor_file, "golden/tfs_model.proto"])

    def add_map_msg(parent, entry_name):
        e = parent.nested_type.add()
        e.name = entry_name
        e.options.map_entry = True
        k = e.field.add()
        k.name, k.number, k.type, k.label = "key", 1, 9, 1
        v = e.field.add()
        v.name, v.number, v.type, v.label, v.type_name = "value", 2, 11, 1, "." + tensor_pb2.TensorProto.DESCRIPTOR.full_name

    req = g.message_type.add()
    req.name = "PredictRequest"
    add_map_msg(req, "InputsEntry")
    fd = req.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "model_spec", 1, 11, 1, ".tensorflow.serving.ModelSpec"
    fd = req.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "inputs", 2, 11, 3, ".tensorflow.serving.PredictRequest.InputsEntry"
    fd = req.field.add()
    fd.name, fd.number, fd.type, fd.label = "output_filter", 3, 9, 3
    resp = g.message_type.add()
    resp.name = "PredictResponse"
    add_map_msg(resp, "OutputsEntry")
    fd = resp.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "outputs", 1, 11, 3, ".tensorflow.serving.PredictResponse.OutputsEntry"
    fd = resp.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "model_spec", 2, 11, 1, ".tensorflow.serving.ModelSpec"
    pool.Add(g)
    get = message_factory.GetMessageClass
    return (get(pool.FindMessageTypeByName("tensorflow.serving.PredictRequest")),
            get(pool.FindMessageTypeByName("tensorflow.serving.PredictResponse")))


PredictRequestPB, PredictResponsePB = build_predict_classes()


def build_metadata_classes():
    """GetModelMetadataRequest / GetModelMetadataResponse (protos/tensorflow_serving/apis/get_model_metadata.proto:16-35 of the
    reference), rebuilt at run time: model_spec = 1; repeated string metadata_field = 2 / map<string, google.protobuf.Any>
    metadata = 2."""
    from google.protobuf import any_pb2  # noqa: F401  (registers google/protobuf/any.proto in the default pool)
    pool = descriptor_pool.Default()
    g = descriptor_pb2.FileDescriptorProto()
    g.name = "golden/tfs_get_model_metadata.proto"
    g.package = "tensorflow.serving"
    g.syntax = "proto3"
    g.dependency.extend(["google/protobuf/any.proto", "golden/tfs_model.proto"])
    req = g.message_type.add()
    req.name = "GetModelMetadataRequest"
    fd = req.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "model_spec", 1, 11, 1, ".tensorflow.serving.ModelSpec"
    fd = req.field.add()
    fd.name, fd.number, fd.type, fd.label = "metadata_field", 2, 9, 3
    resp = g.message_type.add()
    resp.name = "GetModelMetadataResponse"
    e = resp.nested_type.add()
    e.name = "MetadataEntry"
    e.options.map_entry = True
    k = e.field.add()
    k.name, k.number, k.type, k.label = "key", 1, 9, 1
    v = e.field.add()
    v.name, v.number, v.type, v.label, v.type_name = "value", 2, 11, 1, ".google.protobuf.Any"
    fd = resp.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "model_spec", 1, 11, 1, ".tensorflow.serving.ModelSpec"
    fd = resp.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "metadata", 2, 11, 3, ".tensorflow.serving.GetModelMetadataResponse.MetadataEntry"
    pool.Add(g)
    get = message_factory.GetMessageClass
    return (get(pool.FindMessageTypeByName("tensorflow.serving.GetModelMetadataRequest")),
            get(pool.FindMessageTypeByName("tensorflow.serving.GetModelMetadataResponse")))


MetadataRequestPB, MetadataResponsePB = build_metadata_classes()


# ---- plain-dict message descriptions <-> upb messages ---------------------------------------------
def build_example_classes():
    """Feature / Features / Example (feature.proto:65-87, example.proto:88-90), ExampleList / Input (input.proto:15-82) and
    ClassificationRequest (classification.proto:33-40), rebuilt at run time like the predict messages above."""
    pool = descriptor_pool.Default()
    f = descriptor_pb2.FileDescriptorProto()
    f.name, f.package, f.syntax = "golden/tf_example.proto", "tensorflow", "proto3"

    def scalar_list(name, typ, packed):
        m = f.message_type.add()
        m.name = name
        fd = m.field.add()
        fd.name, fd.number, fd.type, fd.label = "value", 1, typ, 3
        if packed:
            fd.options.packed = True
    scalar_list("BytesList", 12, False)
    scalar_list("FloatList", 2, True)
    scalar_list("Int64List", 3, True)
    ft = f.message_type.add()
    ft.name = "Feature"
    ft.oneof_decl.add().name = "kind"
    for name, num, tname in (("bytes_list", 1, "BytesList"), ("float_list", 2, "FloatList"), ("int64_list", 3, "Int64List")):
        fd = ft.field.add()
        fd.name, fd.number, fd.type, fd.label, fd.type_name, fd.oneof_index = name, num, 11, 1, ".tensorflow." + tname, 0
    fs = f.message_type.add()
    fs.name = "Features"
    e = fs.nested_type.add()
    e.name = "FeatureEntry"
    e.options.map_entry = True
    k = e.field.add()
    k.name, k.number, k.type, k.label = "key", 1, 9, 1
    v = e.field.add()
    v.name, v.number, v.type, v.label, v.type_name = "value", 2, 11, 1, ".tensorflow.Feature"
    fd = fs.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "feature", 1, 11, 3, ".tensorflow.Features.FeatureEntry"
    ex = f.message_type.add()
    ex.name = "Example"
    fd = ex.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "features", 1, 11, 1, ".tensorflow.Features"
    pool.Add(f)

    g = descriptor_pb2.FileDescriptorProto()
    g.name, g.package, g.syntax = "golden/tfs_input.proto", "tensorflow.serving", "proto3"
    g.dependency.extend(["golden/tf_example.proto", "golden/tfs_model.proto"])
    el = g.message_type.add()
    el.name = "ExampleList"
    fd = el.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "examples", 1, 11, 3, ".tensorflow.Example"
    inp = g.message_type.add()
    inp.name = "Input"
    inp.oneof_decl.add().name = "kind"
    fd = inp.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name, fd.oneof_index = "example_list", 1, 11, 1, ".tensorflow.serving.ExampleList", 0
    cr = g.message_type.add()
    cr.name = "ClassificationRequest"
    fd = cr.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "model_spec", 1, 11, 1, ".tensorflow.serving.ModelSpec"
    fd = cr.field.add()
    fd.name, fd.number, fd.type, fd.label, fd.type_name = "input", 2, 11, 1, ".tensorflow.serving.Input"
    pool.Add(g)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName(n))
    return get("tensorflow.serving.Input"), get("tensorflow.serving.ClassificationRequest")


def fill_input(pb, examples):
    pb.example_list.SetInParent()
    for ex in examples:
        e = pb.example_list.examples.add()
        e.features.SetInParent()
        for key, ft in ex:
            dst = e.features.feature[key]
            if "bytes_list" in ft:
                dst.bytes_list.SetInParent()
                dst.bytes_list.value.extend([bytes.fromhex(x) for x in ft["bytes_list"]])
            elif "float_list" in ft:
                # through the wire form so that NaN payloads survive
                a = np.frombuffer(bytes.fromhex(ft["float_list"]), dtype="<f4")
                dst.float_list.SetInParent()
                if a.size:
                    dst.float_list.MergeFromString(bytes([0x0a]) + varint(a.nbytes) + a.tobytes())
            else:
                dst.int64_list.SetInParent()
                dst.int64_list.value.extend([int(x) for x in np.frombuffer(bytes.fromhex(ft["int64_list"]), dtype="<i8")])


def input_cases():
    """Encode-only: what build_input produces (src/lib.rs:413-431).  Feature keys are listed in sorted order (upb's
    deterministic map order) and never empty, every Feature has a kind: the default-key / default-value cases where
    upb and prost differ are hand-written in tests/test_oracle.py."""
    rng = np.random.default_rng(777)
    i64 = lambda v: np.asarray(v, dtype="<i8").tobytes().hex()
    f4 = lambda v: np.asarray(v, dtype="<f4").tobytes().hex()
    cases = []
    one = [[["age", {"int64_list": i64([29])}], ["name", {"bytes_list": [b"bob".hex(), b"".hex()]}], ["w", {"float_list": f4([1.5, 2.5])}]]]
    cases.append({"name": "input_one_example", "kind": "input", "msg": {"model_spec": None, "examples": one}})
    cases.append({"name": "input_no_examples", "kind": "input", "msg": {"model_spec": None, "examples": []}})
    cases.append({"name": "input_empty_examples_and_lists", "kind": "input", "msg": {"model_spec": None, "examples": [
        [], [["a", {"float_list": f4([])}], ["b", {"int64_list": i64([])}], ["c", {"bytes_list": []}]], []]}})
    many = []
    for e in range(40):
        feats = []
        for j in range(int(rng.integers(0, 6))):
            which = int(rng.integers(0, 3))
            n = int(rng.integers(0, 50))
            if which == 0:
                ft = {"int64_list": i64(rng.integers(-(1 << 40), 1 << 40, n))}
            elif which == 1:
                ft = {"float_list": f4(rng.standard_normal(n))}
            else:
                ft = {"bytes_list": [rng.integers(0, 256, int(rng.integers(0, 30)), dtype=np.uint8).tobytes().hex() for _ in range(n % 7)]}
            feats.append(["f%02d" % j, ft])
        many.append(feats)
    cases.append({"name": "input_40_examples", "kind": "input", "msg": {"model_spec": None, "examples": many}})
    cases.append({"name": "classify_request", "kind": "classify", "msg": {
        "model_spec": {"name": "census", "signature_name": "classify", "version": 3, "version_label": None}, "examples": one + many[:5]}})
    cases.append({"name": "classify_big_lists", "kind": "classify", "msg": {
        "model_spec": {"name": "wide", "signature_name": "", "version": None, "version_label": None},
        "examples": [[["ids", {"int64_list": i64(rng.integers(0, 30522, 2100))}], ["img", {"bytes_list": [rng.integers(0, 256, 5000, dtype=np.uint8).tobytes().hex()]}],
                      ["x", {"float_list": f4(rng.standard_normal(1000))}]]]}})
    for c in cases:
        if c["kind"] == "input":
            pb = InputPB()
            fill_input(pb, c["msg"]["examples"])
        else:
            pb = ClassificationRequestPB()
            if c["msg"]["model_spec"] is not None:
                fill_model_spec(pb.model_spec, c["msg"]["model_spec"])
            pb.input.SetInParent()
            fill_input(pb.input, c["msg"]["examples"])
        c["wire"] = pb.SerializeToString(deterministic=True).hex()
    return cases


InputPB, ClassificationRequestPB = build_example_classes()


def arr_hex(a):
    return np.ascontiguousarray(a).tobytes().hex()


def tensor_desc(dtype=0, shape=None, unknown_rank=False, version_number=0, **fields):
    d = {"dtype": int(dtype), "shape": None if shape is None else [int(x) for x in shape],
         "unknown_rank": bool(unknown_rank), "version_number": int(version_number), "fields": {}}
    for k, v in fields.items():
        if k == "string_val":  # repeated bytes: a list of hex strings
            d["fields"][k] = [bytes(x).hex() for x in v]
            continue
        d["fields"][k] = arr_hex(np.asarray(v).astype(FIELD_NP[k]))
    return d


def fill_tensor(pb, d):
    pb.dtype = d["dtype"]
    pb.version_number = d["version_number"]
    if d["shape"] is not None:
        pb.tensor_shape.SetInParent()
        for s in d["shape"]:
            pb.tensor_shape.dim.add().size = s
        pb.tensor_shape.unknown_rank = d["unknown_rank"]
    for k, hx in d["fields"].items():
        if k == "string_val":
            pb.string_val.extend([bytes.fromhex(x) for x in hx])
            continue
        a = np.frombuffer(bytes.fromhex(hx), dtype=FIELD_NP[k])
        if k == "tensor_content":
            pb.tensor_content = a.tobytes()
        elif k == "bool_val":
            pb.bool_val.extend([bool(x) for x in a])
        elif k in ("float_val", "scomplex_val"):
            # go through the packed wire form so NaN payloads survive (python floats quiet sNaNs)
            tmp = tensor_pb2.TensorProto()
            num = 5 if k == "float_val" else 9
            tmp.ParseFromString(bytes([(num << 3) | 2]) + varint(a.nbytes) + a.tobytes())
            pb.MergeFrom(tmp)
        elif k in ("double_val", "dcomplex_val"):
            tmp = tensor_pb2.TensorProto()
            num = 6 if k == "double_val" else 12
            tmp.ParseFromString(bytes([(num << 3) | 2]) + varint(a.nbytes) + a.tobytes())
            pb.MergeFrom(tmp)
        else:
            getattr(pb, k).extend([int(x) for x in a])


def varint(v):
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def fill_model_spec(pb, d):
    pb.SetInParent()
    pb.name = d["name"]
    pb.signature_name = d["signature_name"]
    if d.get("version") is not None:
        pb.version.value = d["version"]
        pb.version.SetInParent()


def upb_serialize(case):
    kind, m = case["kind"], case["msg"]
    if kind == "tensor":
        pb = tensor_pb2.TensorProto()
        fill_tensor(pb, m)
    else:
        pb = PredictRequestPB() if kind == "request" else PredictResponsePB()
        if m["model_spec"] is not None:
            fill_model_spec(pb.model_spec, m["model_spec"])
        mp = pb.inputs if kind == "request" else pb.outputs
        for key, td in m["entries"]:
            fill_tensor(mp[key], td)
        if kind == "request":
            pb.output_filter.extend(m["output_filter"])
    return pb.SerializeToString(deterministic=True)


def tensor_from_pb(pb):
    d = {"dtype": int(pb.dtype), "shape": None, "unknown_rank": False, "version_number": int(pb.version_number), "fields": {}}
    if pb.HasField("tensor_shape"):
        d["shape"] = [int(x.size) for x in pb.tensor_shape.dim]
        d["unknown_rank"] = bool(pb.tensor_shape.unknown_rank)
    if len(pb.string_val):
        d["fields"]["string_val"] = [bytes(x).hex() for x in pb.string_val]
    # re-serialise each repeated field alone to get exact bits (python float access quiets sNaNs)
    for k, npdt in FIELD_NP.items():
        if k == "tensor_content":
            if pb.tensor_content:
                d["fields"][k] = pb.tensor_content.hex()
            continue
        vals = getattr(pb, k)
        if len(vals) == 0:
            continue
        if npdt in ("<f4", "<f8"):
            one = tensor_pb2.TensorProto()
            getattr(one, k).extend(vals)  # may quiet sNaN: decode cases avoid sNaN
            raw = one.SerializeToString()
            # strip key + len
            i = 1
            while raw[i] & 0x80:
                i += 1
            d["fields"][k] = raw[i + 1:].hex()
        else:
            d["fields"][k] = arr_hex(np.array([int(x) for x in vals], dtype=np.int64 if npdt != "<u8" else np.uint64).astype(npdt))
    return d


def upb_parse(kind, wire):
    if kind == "tensor":
        pb = tensor_pb2.TensorProto()
        pb.ParseFromString(wire)
        return tensor_from_pb(pb)
    pb = PredictRequestPB() if kind == "request" else PredictResponsePB()
    pb.ParseFromString(wire)
    ms = None
    if pb.HasField("model_spec"):
        s = pb.model_spec
        ms = {"name": s.name, "signature_name": s.signature_name,
              "version": int(s.version.value) if s.WhichOneof("version_choice") == "version" else None,
              "version_label": s.version_label if s.WhichOneof("version_choice") == "version_label" else None}
    mp = pb.inputs if kind == "request" else pb.outputs
    out = {"model_spec": ms, "entries": [[k, tensor_from_pb(mp[k])] for k in sorted(mp.keys())]}
    if kind == "request":
        out["output_filter"] = list(pb.output_filter)
    return out


def f32(bits):
    return np.array(bits, dtype=np.uint32).view(np.float32)


def f64(bits):
    return np.array(bits, dtype=np.uint64).view(np.float64)


def encode_cases():
    rng = np.random.default_rng(12345)
    cases = []

    def add(name, kind, msg):
        cases.append({"name": name, "kind": kind, "msg": msg})

    i64_edges = [0, 1, 127, 128, 16383, 16384, (1 << 21) - 1, 1 << 21, (1 << 28) - 1, 1 << 28, (1 << 35) - 1, 1 << 35,
                 (1 << 42) - 1, 1 << 42, (1 << 49) - 1, 1 << 49, (1 << 56) - 1, 1 << 56, (1 << 63) - 1, -1, -2, -128,
                 -(1 << 31), -(1 << 63)]
    i32_edges = [0, 1, 127, 128, 16383, 16384, (1 << 21) - 1, 1 << 21, (1 << 28) - 1, 1 << 28, (1 << 31) - 1, -1, -127,
                 -128, -(1 << 31)]
    u32_edges = [0, 1, 127, 128, 16383, 16384, (1 << 28) - 1, 1 << 28, (1 << 32) - 1]
    u64_edges = [0, 1, 127, 128, (1 << 35), (1 << 63) - 1, 1 << 63, (1 << 64) - 1]
    f32_special = f32([0x00000000, 0x80000000, 0x3F800000, 0x7F800000, 0xFF800000, 0x7FC00000, 0x7FA00001, 0xFFC12345,
                       0x00000001, 0x807FFFFF, 0x7F7FFFFF])
    f64_special = f64([0, 1 << 63, 0x3FF0000000000000, 0x7FF0000000000000, 0xFFF0000000000000, 0x7FF8000000000000,
                       0x7FF4000000000001, 1, 0x7FEFFFFFFFFFFFFF])

    add("b3_float_2_3_4", "tensor", tensor_desc(1, [3], float_val=[2, 3, 4]))
    add("b3_int64", "tensor", tensor_desc(9, None, int64_val=[-1, 0, 1, 300]))
    add("b3_int_minus1", "tensor", tensor_desc(0, None, int_val=[-1]))
    add("b3_tag_order", "tensor", tensor_desc(0, None, half_val=[1], dcomplex_val=[0.0], uint32_val=[2]))
    add("b3_empty_shape", "tensor", tensor_desc(1, []))
    add("b3_float_bits", "tensor", tensor_desc(0, None, float_val=f32([0x80000000, 0x7FA00001])))
    add("int64_edges", "tensor", tensor_desc(9, [len(i64_edges)], int64_val=np.array(i64_edges, dtype=np.int64)))
    add("int32_edges", "tensor", tensor_desc(3, [len(i32_edges)], int_val=np.array(i32_edges, dtype=np.int32)))
    add("uint32_edges", "tensor", tensor_desc(22, [len(u32_edges)], uint32_val=np.array(u32_edges, dtype=np.uint32)))
    add("uint64_edges", "tensor", tensor_desc(23, [len(u64_edges)], uint64_val=np.array(u64_edges, dtype=np.uint64)))
    add("bool_vals", "tensor", tensor_desc(10, [5], bool_val=[1, 0, 0, 1, 1]))
    add("half_patterns", "tensor", tensor_desc(19, [8], half_val=[0, 1, 0x7F, 0x80, 0x3C00, 0x3FFF, 0x4000, 0xFFFF]))
    add("float_special", "tensor", tensor_desc(1, [len(f32_special)], float_val=f32_special))
    add("double_special", "tensor", tensor_desc(2, [len(f64_special)], double_val=f64_special))
    add("scomplex", "tensor", tensor_desc(8, [2], scomplex_val=[1.5, -2.5, 0.0, 3.25]))
    add("dcomplex", "tensor", tensor_desc(18, [1], dcomplex_val=[1e300, -1e-300]))
    add("content_raw", "tensor", tensor_desc(4, [7], tensor_content=np.arange(7, dtype=np.uint8)))
    add("content_f32", "tensor", tensor_desc(1, [2, 2], tensor_content=np.array([1, 2, 3, 4], np.float32).view(np.uint8)))
    add("version_number", "tensor", tensor_desc(1, [1], version_number=7, float_val=[9.0]))
    add("neg_version_number", "tensor", tensor_desc(0, None, version_number=-1))
    add("unknown_rank", "tensor", tensor_desc(1, [], unknown_rank=True))
    add("dims_zero_neg_big", "tensor", tensor_desc(1, [0, -1, 1 << 40, 127, 128]))
    add("rank8", "tensor", tensor_desc(3, [1, 2, 3, 4, 5, 6, 7, 8], int_val=[5]))
    add("all_fields", "tensor", tensor_desc(
        1, [2], version_number=3, tensor_content=np.array([1, 2, 3], np.uint8), half_val=[15360, 1], float_val=[1.0, -1.0],
        double_val=[2.0], int_val=[-5, 5], scomplex_val=[0.5, 0.25], int64_val=[1 << 40, -7], bool_val=[1, 0],
        dcomplex_val=[4.0, 8.0], uint32_val=[4000000000], uint64_val=[(1 << 64) - 1]))
    add("empty_tensor", "tensor", tensor_desc())
    add("dtype_only_negative", "tensor", tensor_desc(-3))
    # seeded random payloads, a few hundred elements each, lengths that cross the 1/2-byte varint(len) edge
    for n in (1, 32, 33, 127, 128, 129, 1000):
        add(f"rand_int64_{n}", "tensor", tensor_desc(9, [n], int64_val=rng.integers(-(1 << 62), 1 << 62, n) >> rng.integers(0, 62, n)))
        add(f"rand_int32_{n}", "tensor", tensor_desc(3, [n], int_val=(rng.integers(-(1 << 31), 1 << 31, n) >> rng.integers(0, 31, n)).astype(np.int32)))
        add(f"rand_half_{n}", "tensor", tensor_desc(19, [n], half_val=rng.standard_normal(n).astype(np.float16).view(np.uint16).astype(np.int32)))
        add(f"rand_float_{n}", "tensor", tensor_desc(1, [n], float_val=rng.standard_normal(n).astype(np.float32)))
        add(f"rand_double_{n}", "tensor", tensor_desc(2, [n], double_val=rng.standard_normal(n)))
        add(f"rand_u64_{n}", "tensor", tensor_desc(23, [n], uint64_val=rng.integers(0, 1 << 63, n, dtype=np.uint64) >> rng.integers(0, 63, n).astype(np.uint64)))
        add(f"rand_content_{n}", "tensor", tensor_desc(4, [n], tensor_content=rng.integers(0, 256, n, dtype=np.uint8)))
    add("bert_ids_row", "tensor", tensor_desc(9, [1, 512], int64_val=np.random.default_rng(3001).integers(0, 30522, 512)))

    spec = {"name": "half_plus_two", "signature_name": "serving_default", "version": None, "version_label": None}
    add("c1_request", "request", {"model_spec": spec, "entries": [["x", tensor_desc(1, [3], float_val=[1.0, 2.0, 5.0])]],
                                  "output_filter": []})
    add("c1_response", "response", {"model_spec": dict(spec, version=123),
                                    "entries": [["y", tensor_desc(1, [3], float_val=[2.5, 3.0, 4.5])]]})
    add("request_no_spec", "request", {"model_spec": None, "entries": [["a", tensor_desc(3, [1], int_val=[7])]], "output_filter": []})
    add("request_empty_spec", "request", {"model_spec": {"name": "", "signature_name": "", "version": None, "version_label": None},
                                          "entries": [], "output_filter": []})
    add("request_version0", "request", {"model_spec": {"name": "m", "signature_name": "", "version": 0, "version_label": None},
                                        "entries": [], "output_filter": ["out_a", "", "out_b"]})
    add("request_neg_version", "request", {"model_spec": {"name": "m", "signature_name": "s", "version": -5, "version_label": None},
                                           "entries": [["k", tensor_desc(9, [2], int64_val=[1, -1])]], "output_filter": ["y"]})
    add("bert_request", "request", {
        "model_spec": {"name": "bert", "signature_name": "serving_default", "version": None, "version_label": None},
        "entries": [["attention_mask", tensor_desc(9, [2, 16], int64_val=[1] * 20 + [0] * 12)],
                    ["input_ids", tensor_desc(9, [2, 16], int64_val=np.random.default_rng(7).integers(0, 30522, 32))]],
        "output_filter": []})
    add("mixed_response", "response", {
        "model_spec": {"name": "mixed", "signature_name": "sig", "version": 1 << 40, "version_label": None},
        "entries": [["out0", tensor_desc(1, [4], float_val=[1, 2, 3, 4])],
                    ["out1", tensor_desc(19, [3], half_val=[15360, 16384, 0])],
                    ["out2", tensor_desc(10, [2, 2], bool_val=[1, 0, 1, 1])],
                    ["out3", tensor_desc(2, [1], tensor_content=np.array([3.5], np.float64).view(np.uint8))]]})
    add("utf8_names", "response", {"model_spec": {"name": "mödel-✓", "signature_name": "签名", "version": None, "version_label": None},
                                   "entries": [["ключ", tensor_desc(1, [1], float_val=[1.0])]]})
    # string_val (tensor.proto:57): repeated bytes, one record per element, never packed
    add("string_val_basic", "tensor", tensor_desc(7, [3], string_val=[b"hello", b"", b"world!!"]))
    add("string_val_between_int_and_scomplex", "tensor", tensor_desc(7, [2], int_val=[1, -2], string_val=[b"a", b"\x00\xff\x80"], scomplex_val=[1.5, 2.5]))
    add("string_val_len_127_128", "tensor", tensor_desc(7, [2], string_val=[bytes(range(127)), bytes(i % 251 for i in range(128))]))
    srng = np.random.default_rng(808)
    add("string_val_many_small", "tensor", tensor_desc(7, [300], string_val=[srng.integers(0, 256, int(srng.integers(0, 20)), dtype=np.uint8).tobytes() for _ in range(300)]))
    add("string_val_one_16k", "tensor", tensor_desc(7, [1], string_val=[srng.integers(0, 256, 16384 + 3, dtype=np.uint8).tobytes()]))
    add("string_request", "request", {
        "model_spec": {"name": "text", "signature_name": "serving_default", "version": None, "version_label": None},
        "entries": [["examples", tensor_desc(7, [2], string_val=["serialized example one".encode(), "zwei \u2713".encode()])],
                    ["weights", tensor_desc(1, [2], float_val=[0.5, 0.25])]], "output_filter": ["scores"]})
    big = np.random.default_rng(2001).standard_normal(3 * 32 * 32).astype(np.float32)
    add("resnet_shape_small_image_float_val", "request", {
        "model_spec": {"name": "resnet", "signature_name": "serving_default", "version": None, "version_label": None},
        "entries": [["input", tensor_desc(1, [1, 32, 32, 3], float_val=big)]], "output_filter": []})
    return cases


def key(num, wt):
    return varint((num << 3) | wt)


def ld(num, payload):
    return key(num, 2) + varint(len(payload)) + payload


def vi(num, v):
    return key(num, 0) + varint(v)


def decode_cases():
    """Non-canonical but valid wires: what upb parses out of them is the expectation."""
    import struct
    out = []

    def add(name, kind, wire):
        out.append({"name": name, "kind": kind, "wire": wire.hex(), "expect": upb_parse(kind, wire)})

    f32le = lambda *v: struct.pack("<%df" % len(v), *v)
    f64le = lambda *v: struct.pack("<%dd" % len(v), *v)
    add("unpacked_and_chunks_int_val", "tensor", bytes.fromhex("38013a0202033a0104"))
    add("unpacked_floats", "tensor", key(5, 5) + f32le(1.0) + key(5, 5) + f32le(2.0) + ld(5, f32le(3.0, 4.0)))
    add("unpacked_doubles", "tensor", key(6, 1) + f64le(1.0) + ld(6, f64le(2.0)) + key(6, 1) + f64le(3.0))
    add("string_val_interleaved", "tensor", ld(8, b"first") + vi(1, 7) + ld(8, b"") + ld(7, b"\x01\x02") + ld(8, b"third one") + vi(7, 3))
    add("string_val_not_utf8", "tensor", ld(8, b"\xff\xfe\x80") + ld(8, b"\xc3\x28"))
    add("out_of_order_fields", "tensor", ld(5, f32le(1.0)) + vi(1, 1) + ld(2, ld(2, vi(1, 3))))
    add("scalar_last_wins", "tensor", vi(1, 1) + vi(1, 9) + vi(3, 7) + vi(3, 5))
    add("nonminimal_varints", "tensor", key(1, 0) + bytes.fromhex("8100") + ld(7, bytes.fromhex("818000ff8000")))
    add("int32_truncation", "tensor", ld(7, bytes.fromhex("ffffffff1f")))
    add("int32_10byte_wrap", "tensor", ld(7, bytes.fromhex("ffffffffffffffffff7f")))
    add("int64_10byte_wrap", "tensor", ld(10, bytes.fromhex("ffffffffffffffffff7f") + bytes.fromhex("80808080808080808002")))
    add("bool_multibyte", "tensor", ld(11, bytes.fromhex("00" "02" "808001" "8000")))
    add("unknown_fields_skipped", "tensor",
        vi(1, 1) + vi(100, 1) + ld(101, b"abc") + key(102, 1) + bytes(range(1, 9)) + key(103, 5) + bytes([1, 2, 3, 4]) + ld(5, f32le(1.0)))
    add("unknown_group_skipped", "tensor", vi(1, 3) + key(100, 3) + vi(1, 1) + key(101, 3) + key(101, 4) + key(100, 4) + ld(7, b"\x05"))
    add("shape_merge_two_chunks", "tensor", ld(2, ld(2, vi(1, 3))) + ld(2, ld(2, vi(1, 5))) + vi(3, 1))
    add("shape_unknown_rank_then_dims", "tensor", ld(2, vi(3, 1)) + ld(2, ld(2, vi(1, 2)) + vi(3, 0)))
    add("content_twice_last_wins", "tensor", ld(4, b"\x01\x02\x03") + ld(4, b"\x0a\x0b"))
    add("empty_packed_chunks", "tensor", ld(5, b"") + ld(7, b"") + ld(10, b"") + ld(7, b"\x01"))
    add("half_upper_bits", "tensor", vi(1, 19) + ld(13, bytes.fromhex("80807c" "ffff0f" "01")))
    add("uint32_truncation", "tensor", ld(16, bytes.fromhex("ffffffff1f" "05")))
    add("uint64_max", "tensor", ld(17, bytes.fromhex("ffffffffffffffffff01")))
    add("dim_size_twice_and_empty_dim", "tensor", ld(2, ld(2, vi(1, 3) + vi(1, 4)) + ld(2, b"")))
    t1 = vi(1, 1) + ld(7, b"\x01\x02")
    t2 = vi(1, 3) + ld(7, b"\x09")
    add("dup_map_key_last_wins", "response", ld(1, ld(1, b"k") + ld(2, t1)) + ld(1, ld(1, b"k") + ld(2, t2)))
    add("entry_fields_reversed", "response", ld(1, ld(2, vi(1, 1) + ld(7, b"\x07")) + ld(1, b"k")))
    add("entry_value_twice_merges", "response", ld(1, ld(1, b"k") + ld(2, ld(7, b"\x01")) + ld(2, vi(1, 3) + ld(7, b"\x02"))))
    add("entry_key_twice_last_wins", "response", ld(1, ld(1, b"a") + ld(1, b"b") + ld(2, vi(1, 1))))
    add("model_spec_twice_merges", "request", ld(1, ld(1, b"m")) + ld(1, ld(3, b"s")))
    add("version_then_label", "request", ld(1, ld(2, vi(1, 5)) + ld(4, b"abc")))
    add("label_then_version", "request", ld(1, ld(4, b"abc") + ld(2, vi(1, 5))))
    add("version_twice_merges", "request", ld(1, ld(2, vi(1, 5)) + ld(2, b"")))
    add("unknown_top_level", "response", ld(3, b"abc") + ld(1, ld(1, b"k") + ld(2, vi(1, 2))) + vi(4, 1))
    add("filter_and_entries_interleaved", "request", ld(3, b"a") + ld(2, ld(1, b"k") + ld(2, vi(1, 1))) + ld(3, b"b"))
    add("two_entries_multi_kind", "response",
        ld(1, ld(1, b"a") + ld(2, vi(1, 9) + ld(10, bytes.fromhex("ac02" "ffffffffffffffffff01")))) +
        ld(1, ld(1, b"b") + ld(2, vi(1, 1) + ld(5, f32le(1.5, -2.5)) + ld(4, b"xyz"))))
    return out


def error_cases():
    """Malformed wires -> prost 0.6.0 DecodeError class.  Written by hand from prost's
    rules; `upb_rejects` only records that an independent parser refuses them too."""
    E = {"varint": 1, "underflow": 2, "exceeded": 3, "wiretype": 4, "wiretype_value": 5, "key": 6, "tag": 7,
         "recursion": 8, "endgroup": 9, "utf8": 10}
    rows = [
        ("truncated_key_varint", "tensor", "80", E["varint"]),
        ("scalar_unterminated", "tensor", "08ffff", E["varint"]),
        ("varint_11_bytes", "tensor", "08" + "ff" * 10 + "01", E["varint"]),
        ("packed_len_past_end", "tensor", "3a0501", E["underflow"]),
        ("packed_varint_crosses_limit", "tensor", "3a0180" "01" "0801", E["exceeded"]),
        ("packed_varint_unterminated_at_end", "tensor", "3a0280" "80", E["varint"]),
        ("packed_varint_run_of_11", "tensor", "3a0c" + "80" * 11 + "01", E["varint"]),
        ("packed_float_len_not_multiple_inside", "tensor", "2a06" "0000803f" "0000" "08010801", E["exceeded"]),
        ("packed_float_len_not_multiple_at_end", "tensor", "2a06" "0000803f" "0000", E["underflow"]),
        ("unpacked_float_short", "tensor", "2d0000", E["underflow"]),
        ("content_len_past_end", "tensor", "2205010203", E["underflow"]),
        ("dtype_wrong_wire_type", "tensor", "0a0101", E["wiretype"]),
        ("float_val_wrong_wire_type", "tensor", "2801", E["wiretype"]),
        ("string_val_wrong_wire_type", "tensor", "4005", E["wiretype"]),
        ("string_val_len_past_end", "tensor", "4205616263", E["underflow"]),
        ("string_val_len_unterminated", "tensor", "42ff", E["varint"]),
        ("int_val_fixed32_wire_type", "tensor", "3d01000000", E["wiretype"]),
        ("content_varint_wire_type", "tensor", "2001", E["wiretype"]),
        ("wire_type_6", "tensor", "0e", E["wiretype_value"]),
        ("wire_type_7", "tensor", "0f", E["wiretype_value"]),
        ("key_over_u32", "tensor", "808080808001" + "01", E["key"]),
        ("tag_zero", "tensor", "0001", E["tag"]),
        ("end_group_alone", "tensor", "a406", E["endgroup"]),
        ("end_group_mismatch", "tensor", "a306" "ac06", E["endgroup"]),
        ("group_unterminated", "tensor", "a306" "0801", E["varint"]),
        ("shape_len_past_end", "tensor", "120a1202", E["underflow"]),
        ("dim_crosses_shape_limit", "tensor", "1203" "1202" "0803" "00", E["exceeded"]),
        ("shape_wrong_wire_type", "tensor", "1001", E["wiretype"]),
        ("unknown_len_past_end", "tensor", "aa0605" "6162", E["underflow"]),
        ("unknown_fixed64_short", "tensor", "b106" "01020304", E["underflow"]),
        ("map_key_bad_utf8", "response", "0a05" "0a01ff" "1200", E["utf8"]),
        ("model_name_bad_utf8", "response", "1204" "0a02c328", E["utf8"]),
        ("filter_bad_utf8", "request", "1a02" "ed a0".replace(" ", ""), E["utf8"]),
        ("overlong_utf8", "request", "1a02c080", E["utf8"]),
        ("entry_len_past_end", "response", "0a10" "0a016b", E["underflow"]),
        ("entry_value_crosses_limit", "response", "0a06" "0a016b" "1205" "0801" "0801", E["underflow"]),
        ("entry_inner_exceeds", "response", "0a07" "0a016b" "1203" "0801" "0801" "00", E["exceeded"]),
        ("value_wrong_wire_type", "response", "0a05" "0a016b" "1001", E["wiretype"]),
        ("model_spec_wrong_wire_type", "response", "1001", E["wiretype"]),
        ("version_wrong_inner_type", "request", "0a04" "1202" "0a00", E["wiretype"]),
        ("nested_groups_101", "tensor", "a306" * 101 + "a406" * 101, E["recursion"]),
    ]
    out = []
    for name, kind, hx, code in rows:
        wire = bytes.fromhex(hx)
        try:
            upb_parse(kind, wire)
            rejects = False
        except Exception:
            rejects = True
        out.append({"name": name, "kind": kind, "wire": wire.hex(), "code": code, "upb_rejects": rejects})
    return out


def main():
    enc = encode_cases()
    for c in enc:
        c["wire"] = upb_serialize(c).hex()
        # upb must also read its own bytes back to the same description (sanity of this generator)
        back = upb_parse(c["kind"], bytes.fromhex(c["wire"]))
        if c["kind"] == "tensor":
            ok = back == c["msg"] or any("7fa00001" in v or "0100a07f" in v or "f47f" in v for v in c["msg"]["fields"].values() if isinstance(v, str))
            assert ok, (c["name"], back, c["msg"])
    doc = {"generator": "tests/golden/make_golden.py", "protobuf": __import__("google.protobuf").protobuf.__version__,
           "encode": enc, "decode": decode_cases(), "errors": error_cases(), "input": input_cases()}
    path = os.path.join(HERE, "upb_vectors.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=0, ensure_ascii=True)
    print(f"wrote {path}: {len(enc)} encode, {len(doc['decode'])} decode, {len(doc['errors'])} error cases, "
          f"{os.path.getsize(path)} bytes")
    bad = [e["name"] for e in doc["errors"] if not e["upb_rejects"]]
    print("error cases upb accepts (prost-specific strictness):", bad)


if __name__ == "__main__":
    sys.exit(main())
