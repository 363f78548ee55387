"""resource_handle_val (14) and variant_val (15): the two message-typed TensorProto fields (tensor.proto:72-76, 86-94;
resource_handle.proto:13-30), SURVEY.md section 8(f) row 4.

On the ABI an element is the ENCODED BODY of its message.  CPU tests pin the oracle against upb (tensorboard ships the
compiled TensorProto with both fields) and against hand-built vectors for prost's recursion limit; GPU tests compare the
library with the oracle: bytes on encode, element bodies / statuses on decode, through every entry point."""
import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import abi

import helpers
import oracle_bind as ob

try:
    from tensorboard.compat.proto import tensor_pb2
except Exception:  # pragma: no cover
    tensor_pb2 = None


def oracle_tensor_bytes(t):
    return ob.encode_messages([t], abi.MSG_TENSOR_PROTO)[0]


def sample_tensor():
    """A DT_VARIANT / DT_RESOURCE tensor with two handles and two variants, one of which holds a variant tensor itself."""
    inner = tfs.TensorProto(dtype=9, tensor_shape=tfs.TensorShapeProto([3]), int64_val=np.array([-1, 300, 1 << 40], np.int64))
    leaf = tfs.VariantTensorDataProto("leaf", b"\x00\x01meta", [tfs.TensorProto(dtype=1, float_val=np.arange(5, dtype=np.float32))])
    nested = tfs.TensorProto(dtype=21, variant_val=[tfs.variant_body(leaf, oracle_tensor_bytes)])
    v0 = tfs.VariantTensorDataProto("tensorflow::TensorList", b"", [inner, nested])
    v1 = tfs.VariantTensorDataProto("", b"only metadata", [])
    h0 = tfs.ResourceHandleProto("/job:localhost/device:CPU:0", "localhost", "table_7", 0xDEADBEEFCAFEF00D, "N10tensorflow6lookup15LookupInterfaceE")
    h1 = tfs.ResourceHandleProto(name="x")
    t = tfs.TensorProto(dtype=21, tensor_shape=tfs.TensorShapeProto([2]),
                        resource_handle_val=[tfs.resource_handle_body(h) for h in (h0, h1)] + [b""],
                        variant_val=[tfs.variant_body(v, oracle_tensor_bytes) for v in (v0, v1)],
                        uint32_val=np.array([7], np.uint32), half_val=np.array([1], np.int32))
    return t, (h0, h1), (v0, v1, inner, leaf)


@pytest.mark.skipif(tensor_pb2 is None, reason="tensorboard protos not importable")
def test_oracle_matches_upb_on_message_typed_fields():
    t, (h0, h1), (v0, v1, inner, leaf) = sample_tensor()
    pb = tensor_pb2.TensorProto(dtype=21)
    pb.tensor_shape.dim.add().size = 2
    for h in (h0, h1):
        e = pb.resource_handle_val.add()
        e.device, e.container, e.name, e.hash_code, e.maybe_type_name = h.device, h.container, h.name, h.hash_code, h.maybe_type_name
    pb.resource_handle_val.add()  # an all-default element: `72 00`
    e = pb.variant_val.add()
    e.type_name = v0.type_name
    ti = e.tensors.add()
    ti.dtype = 9
    ti.tensor_shape.dim.add().size = 3
    ti.int64_val.extend([-1, 300, 1 << 40])
    tn = e.tensors.add()
    tn.dtype = 21
    lf = tn.variant_val.add()
    lf.type_name, lf.metadata = "leaf", b"\x00\x01meta"
    lt = lf.tensors.add()
    lt.dtype = 1
    lt.float_val.extend([0, 1, 2, 3, 4])
    e = pb.variant_val.add()
    e.metadata = b"only metadata"
    pb.uint32_val.append(7)
    pb.half_val.append(1)
    want = pb.SerializeToString()
    # encode: ascending field numbers put 14 / 15 between half_val (13) and uint32_val (16)
    got = ob.encode_messages([t], abi.MSG_TENSOR_PROTO)[0]
    assert got == want, (got.hex(), want.hex())
    # the element bodies are what upb writes for the elements alone
    assert [bytes(x) for x in t.resource_handle_val] == [x.SerializeToString() for x in pb.resource_handle_val]
    assert [bytes(x) for x in t.variant_val] == [x.SerializeToString() for x in pb.variant_val]
    # decode: upb's bytes come back as the same element bodies
    st, msgs = ob.decode_bytes([want], abi.MSG_TENSOR_PROTO)
    assert st == 0
    m = msgs[0].message
    assert [bytes(x) for x in m.resource_handle_val] == [bytes(x) for x in t.resource_handle_val]
    assert [bytes(x) for x in m.variant_val] == [bytes(x) for x in t.variant_val]
    assert np.array_equal(m.uint32_val, [7]) and np.array_equal(m.half_val, [1])
    # and the host-side helpers open them again
    assert tfs.parse_resource_handle(m.resource_handle_val[0]) == h0 and tfs.parse_resource_handle(m.resource_handle_val[2]) == tfs.ResourceHandleProto()
    name, meta, tensors = tfs.parse_variant(m.variant_val[0])
    assert name == "tensorflow::TensorList" and meta == b"" and len(tensors) == 2
    st, inner_dec = ob.decode_bytes(tensors, abi.MSG_TENSOR_PROTO)
    assert st == 0 and np.array_equal(inner_dec[0].message.int64_val, [-1, 300, 1 << 40])
    assert tfs.parse_variant(inner_dec[1].message.variant_val[0])[0] == "leaf"


def nested_variants(pairs: int, innermost: bytes = b"\x08\x01") -> bytes:
    """TensorProto { variant_val: [ Variant { tensors: [ TensorProto { variant_val: [ ... ] } ] } ] }, `pairs` levels deep."""
    body = innermost  # a TensorProto body
    for _ in range(pairs):
        variant = tfs.messages._len_field(3, body)       # Variant { tensors: [body] }
        body = tfs.messages._len_field(15, variant)      # TensorProto { variant_val: [variant] }
    return body


def malformed_cases():
    ok = tfs.variant_body(tfs.VariantTensorDataProto("t", b"m", [tfs.TensorProto(dtype=9, int64_val=np.array([1, 2, 3], np.int64))]), oracle_tensor_bytes)
    L = tfs.messages._len_field
    return [
        ("depth 50 is accepted", nested_variants(50), abi.TFS_OK),
        ("depth 51 hits prost's recursion limit", nested_variants(51), abi.TFS_ERR_RECURSION),
        ("type_name is not UTF-8", L(15, L(1, b"\xff\xfe")), abi.TFS_ERR_UTF8),
        ("device is not UTF-8", L(14, L(1, b"\xc0\x80")), abi.TFS_ERR_UTF8),
        ("variant element with the varint wire type", bytes([15 << 3 | 0, 1]), abi.TFS_ERR_WIRE_TYPE),
        ("hash_code with a length-delimited wire type", L(14, L(4, b"\x01")), abi.TFS_ERR_WIRE_TYPE),
        ("element longer than the buffer", bytes([15 << 3 | 2, 50]) + ok[:10], abi.TFS_ERR_BUFFER_UNDERFLOW),
        ("nested tensor longer than its variant", L(15, bytes([3 << 3 | 2, 40]) + b"\x08\x01") + b"\x08\x01" * 30, abi.TFS_ERR_LENGTH_EXCEEDED),
        ("over-long varint inside a nested packed int64_val", L(15, L(3, L(10, b"\x01" + b"\xff" * 10 + b"\x01"))), abi.TFS_ERR_INVALID_VARINT),
        ("nested packed float_val of 5 bytes at the end of the buffer", L(15, L(3, L(5, b"\x00" * 5))), abi.TFS_ERR_BUFFER_UNDERFLOW),
        ("nested packed float_val of 5 bytes with more behind it", L(15, L(3, L(5, b"\x00" * 5))) + b"\x08\x01\x18\x02", abi.TFS_ERR_LENGTH_EXCEEDED),
        ("unknown fields and groups inside the elements are skipped", L(15, b"\x9b\x01\x08\x01\x9c\x01" + L(1, b"t") + b"\x38\x07") + L(14, b"\x38\x07" + L(3, b"n")), abi.TFS_OK),
        ("a stray end group (of an unknown field) inside a handle", L(14, b"\x4c"), abi.TFS_ERR_END_GROUP),
        ("an end group under a KNOWN field number is a wire type error first", L(14, b"\x0c"), abi.TFS_ERR_WIRE_TYPE),
        ("dim name inside a NESTED tensor is none of our business", L(15, L(3, L(2, L(2, b"\x08\x03" + L(2, b"rows"))))), abi.TFS_OK),
        ("valid element followed by a truncated one", L(15, ok) + bytes([14 << 3 | 2, 9, 0x0a]), abi.TFS_ERR_BUFFER_UNDERFLOW),
    ]


@pytest.mark.parametrize("name,wire,code", malformed_cases(), ids=[c[0] for c in malformed_cases()])
def test_oracle_validates_elements_like_prost(name, wire, code):
    st, msgs = ob.decode_bytes([wire], abi.MSG_TENSOR_PROTO)
    assert msgs[0].status == code, (name, abi.STATUS_NAMES.get(msgs[0].status))
    if tensor_pb2 is not None and "depth" not in name and "dim name" not in name:  # accept / reject agrees with upb, except wire types (see test_oracle_upb_messages)
        pb = tensor_pb2.TensorProto()
        try:
            pb.ParseFromString(wire)
            upb_ok = True
        except Exception:
            upb_ok = False
        if code != abi.TFS_ERR_WIRE_TYPE:
            assert upb_ok == (code == abi.TFS_OK), name


# ---- GPU ---------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def codec():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    c = tfs.GpuCodec(tfs.GpuContext(0))
    yield c
    c.ctx.close()


@pytest.mark.gpu
def test_gpu_message_typed_fields_roundtrip(codec):
    import test_gpu_parity as P
    t, _, _ = sample_tensor()
    big = tfs.VariantTensorDataProto("big", b"", [tfs.TensorProto(dtype=1, float_val=np.random.default_rng(2).standard_normal(70_000).astype(np.float32))])
    t2 = tfs.TensorProto(dtype=21, variant_val=[tfs.variant_body(big, oracle_tensor_bytes)] * 3 + [b""] * 40)
    msgs = [tfs.PredictResponse([("handles", t), ("plain", tfs.TensorProto(dtype=1, float_val=np.ones(9, np.float32))), ("v", t2)],
                                tfs.ModelSpec("m", version=2)), tfs.PredictResponse([("only", t2)])]
    kind = abi.MSG_PREDICT_RESPONSE
    want = ob.encode_messages(msgs, kind)
    got = P.gpu_encode_both(codec, msgs, kind)
    P.assert_bytes_equal(got, want, "message-typed fields encode")
    st, dec = P.gpu_decode_both(codec, got, kind)
    st_o, dec_o = ob.decode_bytes(want, kind)
    assert st == st_o == 0
    helpers.assert_same_decoded(dec_o, dec, "response", "message-typed fields decode")
    # the tensors inside a decoded variant go through the library again
    name, meta, inner = tfs.parse_variant(dec[0].message.outputs[2][1].variant_val[0])
    st, inner_dec = P.gpu_decode_both(codec, inner, abi.MSG_TENSOR_PROTO)
    assert st == 0 and name == "big" and np.array_equal(inner_dec[0].message.float_val, big.tensors[0].float_val)
    # and the codec encodes the tensors of a variant itself
    enc = lambda x: codec.encoder().encode_host([x], abi.MSG_TENSOR_PROTO).all_bytes()[0]
    assert tfs.variant_body(big, enc) == tfs.variant_body(big, oracle_tensor_bytes)


@pytest.mark.gpu
def test_gpu_validates_elements_like_the_oracle(codec):
    import test_gpu_parity as P
    cases = malformed_cases()
    for kind, wrap in ((abi.MSG_TENSOR_PROTO, lambda b: b),
                       (abi.MSG_PREDICT_RESPONSE, lambda b: tfs.messages._len_field(1, tfs.messages._len_field(1, b"k") + tfs.messages._len_field(2, b)))):
        blobs = [wrap(c[1]) for c in cases]
        st_o, want = ob.decode_bytes(blobs, kind)
        st, got = P.gpu_decode_both(codec, blobs, kind)
        for c, w, g in zip(cases, want, got):
            assert w.status == g.status, (c[0], abi.STATUS_NAMES.get(w.status), abi.STATUS_NAMES.get(g.status))
        helpers.assert_same_decoded(want, got, P.helpers_kind(kind), "message-typed fields, malformed table")
    # inside a response the budget is two lower (map entry, TensorProto): depth 49 still fits, 50 does not
    for pairs, code in ((49, abi.TFS_OK), (50, abi.TFS_ERR_RECURSION)):
        b = wrap(nested_variants(pairs))
        st_o, want = ob.decode_bytes([b], abi.MSG_PREDICT_RESPONSE)
        st, got = P.gpu_decode_both(codec, [b], abi.MSG_PREDICT_RESPONSE)
        assert want[0].status == got[0].status == code, (pairs, want[0].status, got[0].status)
