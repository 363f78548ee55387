"""GetModelMetadataRequest / GetModelMetadataResponse (the reference's model_metadata call,
src/lib.rs:300-312; IDL get_model_metadata.proto:16-35 + google/protobuf/any.proto).

CPU part (no GPU): the oracle against upb on messages rebuilt from the IDL at run time, and against hand-built wire for
what only prost does (merge of repeated occurrences, error classes).  GPU part: the library against the oracle - requests
encoded byte for byte, responses (valid, merged, malformed, mutated) decoded to the same values and the same errors.
"""
import os
import sys

import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import abi

import oracle_bind as ob

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
try:
    import make_golden as mg
except Exception:  # pragma: no cover
    mg = None

SIG_URL = "type.googleapis.com/tensorflow.serving.SignatureDefMap"


def ld(num, payload):
    out = bytearray([num << 3 | 2])
    n = len(payload)
    while n >= 0x80:
        out.append((n & 0x7F) | 0x80)
        n >>= 7
    out.append(n)
    return bytes(out) + bytes(payload)


def requests():
    rng = np.random.default_rng(5)
    out = [tfs.GetModelMetadataRequest(None, []), tfs.GetModelMetadataRequest(tfs.ModelSpec("half_plus_two"), ["signature_def"]),
           tfs.GetModelMetadataRequest(tfs.ModelSpec("", version=0), [""]), tfs.GetModelMetadataRequest(None, ["a", "", "signature_def" * 20])]
    for i in range(40):
        spec = None if i % 7 == 0 else tfs.ModelSpec("m%d" % i * (i % 3), version=int(rng.integers(-5, 1 << 40)) if i % 2 else None,
                                                     version_label="canary" if i % 4 == 2 else None, signature_name="s" * (i % 5))
        out.append(tfs.GetModelMetadataRequest(spec, ["f%d" % k * k for k in range(i % 4)]))
    return out


def responses():
    """(wire bytes, what they are) - canonical ones as upb writes them plus hand-built ones"""
    rng = np.random.default_rng(6)
    blobs = []
    if mg is not None:
        for i in range(30):
            pb = mg.MetadataResponsePB()
            if i % 5:
                pb.model_spec.name = "model-%d" % i
                if i % 2:
                    pb.model_spec.version.value = int(rng.integers(0, 1 << 50))
            for k in range(i % 4):
                a = pb.metadata["key%d" % k * (1 + k % 2)]
                a.type_url = SIG_URL if k % 2 == 0 else "tést/%d" % k
                a.value = bytes(rng.integers(0, 256, int(rng.integers(0, 3000)), dtype=np.uint8))
            blobs.append(pb.SerializeToString())
    any1 = ld(1, SIG_URL.encode()) + ld(2, b"\x0a\x02hi")
    blobs += [
        b"",
        ld(2, b""),                                                        # an entry of defaults: key "", Any::default()
        ld(2, ld(1, b"k")),                                                # value left out
        ld(2, ld(2, any1)),                                                # key left out
        ld(2, ld(1, b"k") + ld(2, any1) + ld(2, ld(1, b"second/url"))),     # the Any twice: fields merge, the later type_url wins
        ld(2, ld(1, b"k") + ld(2, ld(1, b"a") + ld(1, b"b") + ld(2, b"x") + ld(2, b"yz"))),  # repeated scalars: last wins
        ld(2, ld(1, b"k1") + ld(1, b"k2") + ld(2, any1)),                   # repeated key: last wins
        ld(2, ld(2, any1) + ld(1, b"late-key")),                           # value before key
        ld(2, ld(1, b"k") + ld(2, any1) + bytes([0x18, 0x05])) + ld(1, ld(1, b"m")),  # unknown varint field in the entry; spec after the map
        ld(2, ld(1, b"k") + ld(2, any1 + bytes([0x1D, 1, 2, 3, 4]))),       # unknown fixed32 inside the Any
        ld(2, ld(1, b"k") + ld(2, any1)) * 3,                               # the same key three times: three records, in order
        # malformed
        ld(2, ld(1, b"k") + ld(2, ld(1, b"\xff\xfe"))),                     # type_url is a string: invalid UTF-8
        ld(2, ld(1, b"\xc0\x80") + ld(2, any1)),                           # overlong encoding in the key
        ld(2, ld(1, b"k") + bytes([0x10, 0x01])),                           # value with wire type varint
        ld(2, ld(1, b"k") + ld(2, bytes([0x08, 0x01]))),                    # type_url with wire type varint
        ld(2, ld(1, b"k") + ld(2, bytes([0x15, 1, 2, 3, 4]))),              # value bytes with wire type fixed32
        ld(2, ld(1, b"k") + ld(2, any1))[:-3],                              # truncated
        ld(2, ld(1, b"k") + ld(2, any1[:-1]) + b"\x00"),                    # inner length runs over its parent
        bytes([0x12, 0x05]) + ld(1, b"k"),                                  # entry shorter than its length prefix says
        ld(2, ld(1, b"k") + ld(2, any1) + bytes([0x0C])),                   # stray end-group
        ld(1, bytes([0x0A, 0x03, 0x61])),                                   # model_spec: name runs out of bytes
    ]
    return blobs


def same_messages(a, b, ctx):
    assert len(a) == len(b), ctx
    for i, (x, y) in enumerate(zip(a, b)):
        assert x.status == y.status, f"{ctx}: message {i}: status {x.status} != {y.status} (offsets {x.err_offset}, {y.err_offset})"
        if x.status == 0:
            assert x.message == y.message, f"{ctx}: message {i}: {x.message} != {y.message}"


# ---- CPU: oracle against upb -----------------------------------------------------------------------------------------
@pytest.mark.skipif(mg is None, reason="protobuf (upb) not importable")
def test_oracle_request_bytes_match_upb():
    for r in requests():
        if r.model_spec is not None and r.model_spec.version_label is not None:
            continue  # upb writes the oneof members in field order, prost the label last
        pb = mg.MetadataRequestPB()
        if r.model_spec is not None:
            pb.model_spec.SetInParent()
            pb.model_spec.name, pb.model_spec.signature_name = r.model_spec.name, r.model_spec.signature_name
            if r.model_spec.version is not None:
                pb.model_spec.version.value = r.model_spec.version
                pb.model_spec.version.SetInParent()
        pb.metadata_field.extend(r.metadata_field)
        want = pb.SerializeToString()
        assert ob.encode_messages([r], abi.MSG_METADATA_REQUEST)[0] == want, r
        st, back = ob.decode_bytes([want], abi.MSG_METADATA_REQUEST)
        assert st == 0 and back[0].message == r
        assert ob.encode_size(tfs.build_batch([r], abi.MSG_METADATA_REQUEST)) == [len(want)]


@pytest.mark.skipif(mg is None, reason="protobuf (upb) not importable")
def test_oracle_response_values_match_upb():
    n_ok = 0
    for wire in responses():
        st, got = ob.decode_bytes([wire], abi.MSG_METADATA_RESPONSE)
        pb = mg.MetadataResponsePB()
        try:
            pb.ParseFromString(wire)
            upb_ok = True
        except Exception:
            upb_ok = False
        if got[0].status != 0:
            # prost checks the wire type of known fields (upb files them under unknown fields) and, like upb, UTF-8 and lengths
            assert not upb_ok or got[0].status == abi.TFS_ERR_WIRE_TYPE, (wire.hex(), got[0].status)
            continue
        assert upb_ok, wire.hex()
        n_ok += 1
        m = got[0].message
        # a HashMap: the last entry of a key wins, as in upb's map
        last = {k: a for k, a in m.metadata}
        if bytes([0x18, 0x05]) in wire and not pb.metadata:
            continue  # an unknown field INSIDE a map entry: upb keeps the whole entry as an unknown field, prost skips the field
        assert set(last) == set(pb.metadata.keys())
        for k, a in last.items():
            assert (a.type_url, a.value) == (pb.metadata[k].type_url, pb.metadata[k].value), (wire.hex(), k)
        assert (m.model_spec is not None) == pb.HasField("model_spec")
        if m.model_spec is not None:
            assert m.model_spec.name == pb.model_spec.name
            assert (m.model_spec.version or 0) == pb.model_spec.version.value
    assert n_ok >= 35


def test_oracle_response_prost_rules():
    """What the hand-built cases pin beyond upb: merge order and error classes, as prost's derived code decides them."""
    dec = lambda w: ob.decode_bytes([w], abi.MSG_METADATA_RESPONSE)[1][0]
    any1 = ld(1, SIG_URL.encode()) + ld(2, b"\x0a\x02hi")
    m = dec(ld(2, ld(1, b"k") + ld(2, any1) + ld(2, ld(1, b"second/url")))).message
    assert m.metadata == [("k", tfs.AnyProto("second/url", b"\x0a\x02hi"))]
    m = dec(ld(2, ld(1, b"k") + ld(2, ld(1, b"a") + ld(1, b"b") + ld(2, b"x") + ld(2, b"yz")))).message
    assert m.metadata == [("k", tfs.AnyProto("b", b"yz"))]
    assert dec(ld(2, b"")).message.metadata == [("", tfs.AnyProto("", b""))]
    assert dec(ld(2, ld(1, b"k") + ld(2, ld(1, b"\xff\xfe")))).status == abi.TFS_ERR_UTF8
    assert dec(ld(2, ld(1, b"k") + bytes([0x10, 0x01]))).status == abi.TFS_ERR_WIRE_TYPE
    assert dec(ld(2, ld(1, b"k") + ld(2, bytes([0x15, 1, 2, 3, 4])))).status == abi.TFS_ERR_WIRE_TYPE
    assert dec(ld(2, ld(1, b"k") + ld(2, any1))[:-3]).status == abi.TFS_ERR_BUFFER_UNDERFLOW
    assert dec(ld(2, ld(1, b"k") + ld(2, any1) + bytes([0x0C]))).status in (abi.TFS_ERR_END_GROUP, abi.TFS_ERR_WIRE_TYPE)


# ---- GPU: the library against the oracle -----------------------------------------------------------------------------
@pytest.fixture(scope="module")
def codec():
    from tfs_b200.codec import GpuCodec
    c = GpuCodec()
    yield c
    c.ctx.close()


@pytest.mark.gpu
def test_gpu_metadata_requests_encode_like_the_oracle(codec):
    import torch
    reqs = requests()
    want = ob.encode_messages(reqs, abi.MSG_METADATA_REQUEST)
    enc, dec = codec.encoder(), codec.decoder()
    for flags in (0, abi.F_GRPC_PREFIX):
        want_f = ob.encode_messages(reqs, abi.MSG_METADATA_REQUEST, flags)
        assert enc.encode_host(reqs, abi.MSG_METADATA_REQUEST, flags).all_bytes() == want_f
        assert enc.encode(reqs, abi.MSG_METADATA_REQUEST, flags).all_bytes() == want_f
    assert enc.encoded_len(reqs, abi.MSG_METADATA_REQUEST) == [len(w) for w in want]
    # and back
    eb = enc.encode_host(reqs, abi.MSG_METADATA_REQUEST)
    d = dec.decode_host(eb.wire[: eb.total], eb.offsets, eb.lengths, abi.MSG_METADATA_REQUEST)
    same_messages(ob.decode_bytes(want, abi.MSG_METADATA_REQUEST)[1], d.messages(), "request decode")
    d.close()
    built = tfs.build_batch([tfs.GetModelMetadataRequest()], abi.MSG_METADATA_REQUEST)
    built.msgs[0].kind = abi.MSG_METADATA_RESPONSE  # the response is decode only: refused by the library itself
    with pytest.raises(tfs.TfsError) as e:
        enc.encode_host(built, abi.MSG_METADATA_RESPONSE)
    assert e.value.status == abi.TFS_ERR_INVALID_ARG
    del torch


def gpu_decode(codec, blobs, kind):
    import torch
    offs, lens, cur = [], [], 0
    for b in blobs:
        offs.append(cur)
        lens.append(len(b))
        cur += len(b)
    wire = np.frombuffer(b"".join(blobs), dtype=np.uint8) if cur else np.zeros(0, np.uint8)
    dec = codec.decoder()
    host = dec.decode_host(wire, offs, lens, kind)
    host_msgs = host.messages()
    dwire = torch.zeros(cur + 64, dtype=torch.uint8, device=f"cuda:{codec.ctx.device}")
    if cur:
        dwire[5: 5 + cur] = torch.from_numpy(wire.copy()).to(dwire.device)
    d = dec.decode(dwire[5: 5 + cur], offs, lens, kind)
    codec.ctx.sync()
    same_messages(host_msgs, d.messages(), "host vs device decode")
    return host_msgs


@pytest.mark.gpu
def test_gpu_metadata_responses_decode_like_the_oracle(codec):
    blobs = responses()
    want = ob.decode_bytes(blobs, abi.MSG_METADATA_RESPONSE)[1]
    assert sum(m.status == 0 for m in want) >= 10 and sum(m.status != 0 for m in want) >= 8
    same_messages(want, gpu_decode(codec, blobs, abi.MSG_METADATA_RESPONSE), "responses, one batch")
    for i, b in enumerate(blobs):  # and one by one (small-batch plan kernel, per-message quotas)
        same_messages([want[i]], gpu_decode(codec, [b], abi.MSG_METADATA_RESPONSE), f"response {i}")


@pytest.mark.gpu
def test_gpu_metadata_response_mutations(codec):
    rng = np.random.default_rng(17)
    base = [b for b in responses() if ob.decode_bytes([b], abi.MSG_METADATA_RESPONSE)[1][0].status == 0 and len(b) > 8]
    blobs = []
    for i in range(600):
        b = bytearray(base[i % len(base)])
        op = rng.integers(0, 4)
        if op == 0:
            b = b[: rng.integers(0, len(b))]
        elif op == 1:
            for _ in range(int(rng.integers(1, 4))):
                b[rng.integers(0, len(b))] = rng.integers(0, 256)
        elif op == 2:
            j = rng.integers(0, len(b))
            b[j:j] = bytes(rng.integers(0, 256, rng.integers(1, 5), dtype=np.uint8))
        else:
            j = rng.integers(0, min(len(b), 40))
            b[j] ^= 1 << int(rng.integers(0, 8))
        blobs.append(bytes(b))
    want = ob.decode_bytes(blobs, abi.MSG_METADATA_RESPONSE)[1]
    assert sum(m.status != 0 for m in want) > 100
    same_messages(want, gpu_decode(codec, blobs, abi.MSG_METADATA_RESPONSE), "mutated responses")


@pytest.mark.gpu
def test_gpu_metadata_large_batch_takes_the_two_walk_path(codec):
    """More than 1 024 messages: the count / fill walks instead of the single-walk plan kernel."""
    base = responses()
    blobs = [base[i % len(base)] for i in range(3000)]
    want = ob.decode_bytes(blobs, abi.MSG_METADATA_RESPONSE)[1]
    same_messages(want, gpu_decode(codec, blobs, abi.MSG_METADATA_RESPONSE), "3 000 responses")
    reqs = requests() * 40
    wire = ob.encode_messages(reqs, abi.MSG_METADATA_REQUEST)
    assert codec.encoder().encode_host(reqs, abi.MSG_METADATA_REQUEST).all_bytes() == wire
    same_messages(ob.decode_bytes(wire, abi.MSG_METADATA_REQUEST)[1], gpu_decode(codec, wire, abi.MSG_METADATA_REQUEST), "1 760 requests")
