"""GPU parity tests (run on the B200 box with `-m gpu`): the CUDA path, called through the C ABI
(ctypes -> csrc/libtfs_gpu.so), against the CPU oracle and the committed golden vectors.

Bar: bit-exact — wire bytes identical on encode, values identical on decode, same DecodeError class on
malformed input.  Nothing here reads /root/reference.
"""
import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import abi, workloads

import helpers
import oracle_bind as ob

pytestmark = pytest.mark.gpu

G = helpers.golden()


@pytest.fixture(scope="module")
def codec():
    from tfs_b200.codec import GpuCodec
    c = GpuCodec()
    yield c
    c.ctx.close()


def gpu_encode_both(codec, msgs, kind, flags=0):
    """Encode through the host entry point and through the device entry point; both must agree."""
    import torch
    enc = codec.encoder()
    host = enc.encode_host(msgs, kind, flags).all_bytes()
    dev_msgs = workloads.to_device(msgs, kind, torch.device(f"cuda:{codec.ctx.device}"))
    dev = enc.encode(dev_msgs, kind, flags).all_bytes()
    assert host == dev, "host-staged and device-resident encode differ"
    # pinned payloads and a pinned destination: the kernels read and write host memory in place
    pmsgs, keep = pinned_like(msgs, kind)
    built = tfs.build_batch(pmsgs, kind)
    pdst = torch.empty(enc.bound(built, flags) + 16, dtype=torch.uint8).pin_memory()
    pinned = enc.encode_host(built, kind, flags, dst=pdst.numpy()).all_bytes()
    assert host == pinned, "pageable and pinned (in place) host encode differ"
    return host


def pinned_like(msgs, kind):
    """The same messages with every payload array moved to pinned host memory (numpy views of torch tensors)."""
    import copy
    import torch

    def mv(v):
        if v is None:
            return None
        if isinstance(v, tfs.Payload):
            return tfs.Payload(mv(v.data), v.src_type if v.src_type != abi.T_NONE else tfs.messages._array_info(v.data)[2], v.dst_type, v.op, v.op_param)
        if isinstance(v, tuple):
            return tuple(mv(x) for x in v)
        if isinstance(v, list):  # string_val
            blk = tfs.ragged_block(v)
            return tfs.Ragged(mv(blk), len(v), blk.size - 8 * (len(v) + 1))
        a = np.ascontiguousarray(v)
        if a.size == 0:
            return a
        t = torch.empty(a.nbytes + 16, dtype=torch.uint8).pin_memory()
        view = t.numpy()[: a.nbytes].view(a.dtype if a.dtype != np.bool_ else np.uint8)
        view[:] = a.reshape(-1) if a.dtype != np.bool_ else a.reshape(-1).view(np.uint8)
        keep.append(t)
        return view.view(np.bool_) if a.dtype == np.bool_ else view

    keep = []

    def mv_tensor(t):
        t2 = copy.copy(t)
        for k in abi.SLOT_KINDS:
            name = abi.KIND_FIELD_NAMES[k]
            setattr(t2, name, mv(getattr(t, name)))
        return t2

    def mv_input(inp):
        return tfs.Input([[(k, tfs.Feature(mv(f.bytes_list), mv(f.float_list), mv(f.int64_list))) for k, f in ex] for ex in inp.examples])

    out = []
    for m in msgs:
        if kind == abi.MSG_TENSOR_PROTO:
            out.append(mv_tensor(m))
        elif kind == abi.MSG_PREDICT_REQUEST:
            out.append(tfs.PredictRequest(m.model_spec, [(k, mv_tensor(t)) for k, t in m.inputs], list(m.output_filter)))
        elif kind == abi.MSG_INPUT:
            out.append(mv_input(m))
        elif kind == abi.MSG_CLASSIFY_REQUEST:
            out.append(tfs.ClassificationRequest(m.model_spec, mv_input(m.input)))
        else:
            out.append(tfs.PredictResponse([(k, mv_tensor(t)) for k, t in m.outputs], m.model_spec))
    return out, keep


def first_diff(a: bytes, b: bytes):
    n = min(len(a), len(b))
    i = next((i for i in range(n) if a[i] != b[i]), n)
    return f"len {len(a)} vs {len(b)}, first diff at byte {i}: {a[max(0, i - 8): i + 8].hex()} vs {b[max(0, i - 8): i + 8].hex()}"


def assert_bytes_equal(got, want, ctx=""):
    assert len(got) == len(want), f"{ctx}: message count"
    for i, (g, w) in enumerate(zip(got, want)):
        assert g == w, f"{ctx}: message {i}: {first_diff(g, w)}"


# ---- encode: golden vectors and the oracle ---------------------------------------------------------
@pytest.mark.parametrize("kind_name", ["tensor", "request", "response"])
def test_encode_golden_vectors(codec, kind_name):
    cases = [c for c in G["encode"] if c["kind"] == kind_name]
    msgs = [helpers.message_from_desc(kind_name, c["msg"]) for c in cases]
    got = gpu_encode_both(codec, msgs, helpers.KIND_OF[kind_name])
    for c, g in zip(cases, got):
        assert g.hex() == c["wire"], f"{c['name']}: {first_diff(g, bytes.fromhex(c['wire']))}"


def test_encode_golden_one_by_one(codec):
    # single-message batches exercise the n_msgs == 1 layout path
    for c in G["encode"][:40]:
        msg = helpers.message_from_desc(c["kind"], c["msg"])
        got = codec.encoder().encode_host([msg], helpers.KIND_OF[c["kind"]]).all_bytes()[0]
        assert got.hex() == c["wire"], c["name"]


def test_encode_prost_specific_cases(codec):
    cases = [
        (tfs.PredictResponse([("k", tfs.TensorProto())]), abi.MSG_PREDICT_RESPONSE, "0a030a016b"),
        (tfs.PredictResponse([("", tfs.TensorProto(dtype=1))]), abi.MSG_PREDICT_RESPONSE, "0a0412020801"),
        (tfs.PredictResponse([("", tfs.TensorProto())]), abi.MSG_PREDICT_RESPONSE, "0a00"),
        (tfs.PredictRequest(tfs.ModelSpec("m", version_label="stable", signature_name="s")), abi.MSG_PREDICT_REQUEST,
         "0a0e0a016d2206737461626c651a0173"),
        (tfs.PredictRequest(tfs.ModelSpec("", version_label="", signature_name="")), abi.MSG_PREDICT_REQUEST, "0a022200"),
        (tfs.PredictResponse([("b", tfs.TensorProto(dtype=1)), ("a", tfs.TensorProto(dtype=2))]), abi.MSG_PREDICT_RESPONSE,
         "0a070a0162120208010a070a016112020802"),
    ]
    for msg, kind, want in cases:
        assert codec.encoder().encode_host([msg], kind).all_bytes()[0].hex() == want


def test_encode_size_matches_oracle(codec):
    msgs = workloads.c5_responses(64, seed=11, max_log2=12)
    built = tfs.build_batch(msgs, abi.MSG_PREDICT_RESPONSE)
    want = ob.encode_size(built)
    import torch
    dev = workloads.to_device(msgs, abi.MSG_PREDICT_RESPONSE, torch.device(f"cuda:{codec.ctx.device}"))
    assert codec.encoder().encoded_len(dev, abi.MSG_PREDICT_RESPONSE) == want


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 255, 256, 257, 2047, 2048, 2049, 4096, 6145, 100_003])
@pytest.mark.parametrize("field", ["int64_val", "int_val", "half_val", "uint32_val", "uint64_val", "bool_val", "float_val",
                                   "double_val", "tensor_content"])
def test_encode_sizes_around_tile_edges(codec, field, n):
    rng = np.random.default_rng(n * 131 + len(field))
    shift = rng.integers(0, 62, n)
    if field == "int64_val":
        v = (rng.integers(-(1 << 62), 1 << 62, n) >> shift).astype(np.int64)
    elif field == "int_val":
        v = (rng.integers(-(1 << 31), 1 << 31, n) >> (shift % 31)).astype(np.int32)
    elif field == "half_val":
        v = rng.integers(0, 65536, n).astype(np.int32)
    elif field == "uint32_val":
        v = (rng.integers(0, 1 << 32, n, dtype=np.uint64) >> (shift % 32).astype(np.uint64)).astype(np.uint32)
    elif field == "uint64_val":
        v = rng.integers(0, 1 << 63, n, dtype=np.uint64) * 2 >> shift.astype(np.uint64)
    elif field == "bool_val":
        v = rng.random(n) < 0.5
    elif field == "float_val":
        v = rng.standard_normal(n).astype(np.float32)
    elif field == "double_val":
        v = rng.standard_normal(n)
    else:
        v = rng.integers(0, 256, n, dtype=np.uint8)
    # three map keys of different length put the payload at different byte alignments
    msgs = [tfs.PredictResponse([("k" * j, tfs.TensorProto(dtype=1, **{field: v}))]) for j in (1, 2, 7)]
    want = ob.encode_messages(msgs, abi.MSG_PREDICT_RESPONSE)
    assert_bytes_equal(gpu_encode_both(codec, msgs, abi.MSG_PREDICT_RESPONSE), want, f"{field} n={n}")


def test_encode_every_payload_alignment(codec):
    # key lengths 0..31 move the payload start through all 16-byte phases; sources are offset too
    rng = np.random.default_rng(7)
    base_f = rng.standard_normal(5000).astype(np.float32)
    base_i = rng.integers(-1000, 1 << 40, 5000).astype(np.int64)
    msgs = []
    for j in range(32):
        msgs.append(tfs.PredictRequest(tfs.ModelSpec("m" * (j % 5)), [
            ("a" * j, tfs.TensorProto(dtype=1, float_val=base_f[j % 7: 4000 + j])),
            ("b" * (31 - j), tfs.TensorProto(dtype=9, int64_val=base_i[j % 3: 3000 + 2 * j])),
            ("c", tfs.TensorProto(dtype=4, tensor_content=np.ascontiguousarray(base_f.view(np.uint8)[j: 9000 + j]))),
        ], ["f" * j]))
    want = ob.encode_messages(msgs, abi.MSG_PREDICT_REQUEST)
    for flags in (0, abi.F_GRPC_PREFIX, abi.F_ALIGN_PAYLOAD, abi.F_GRPC_PREFIX | abi.F_ALIGN_PAYLOAD):
        got = gpu_encode_both(codec, msgs, abi.MSG_PREDICT_REQUEST, flags)
        ref = ob.encode_messages(msgs, abi.MSG_PREDICT_REQUEST, flags & abi.F_GRPC_PREFIX)
        assert_bytes_equal(got, ref, f"flags={flags}")
    assert want == ob.encode_messages(msgs, abi.MSG_PREDICT_REQUEST)


def test_align_payload_flag_aligns_the_bulk_payload(codec):
    import torch
    x = np.random.default_rng(3).standard_normal(100_000).astype(np.float32)
    msgs = [tfs.PredictRequest(tfs.ModelSpec("resnet" + "x" * j), [("input", tfs.TensorProto(dtype=1, float_val=x))]) for j in range(5)]
    dev = workloads.to_device(msgs, abi.MSG_PREDICT_REQUEST, torch.device(f"cuda:{codec.ctx.device}"))
    eb = codec.encoder().encode(dev, abi.MSG_PREDICT_REQUEST, abi.F_ALIGN_PAYLOAD)
    base = eb.wire.data_ptr()
    for o, l in zip(eb.offsets, eb.lengths):
        payload_addr = base + o + l - 4 * x.size  # float_val is the last field of the message
        assert payload_addr % 16 == 0
    assert_bytes_equal(eb.all_bytes(), ob.encode_messages(msgs, abi.MSG_PREDICT_REQUEST))


def test_fused_casts(codec):
    x = np.concatenate([workloads.c4_payload(50_000), np.array([np.inf, -np.inf, np.nan, 0.0, -0.0, 65504.0, 65520.0, 1e-8, 6e-8],
                                                                 np.float32),
                        np.array([0x7FA00001, 0xFFC00001, 0x00000001, 0x33000000, 0x33000001], np.uint32).view(np.float32)])
    px = np.random.default_rng(1001).integers(0, 256, 224 * 224 * 3, dtype=np.uint8)
    msgs = [
        tfs.TensorProto(dtype=19, half_val=tfs.Payload(x, abi.T_F32, abi.T_F16)),
        tfs.TensorProto(dtype=14, half_val=tfs.Payload(x, abi.T_F32, abi.T_BF16)),
        tfs.TensorProto(dtype=19, tensor_content=tfs.Payload(x, abi.T_F32, abi.T_F16)),
        tfs.TensorProto(dtype=14, tensor_content=tfs.Payload(x, abi.T_F32, abi.T_BF16)),
        tfs.TensorProto(dtype=1, float_val=tfs.Payload(px, abi.T_U8, abi.T_F32, abi.OP_DIV, 255.0)),
        tfs.TensorProto(dtype=1, float_val=tfs.Payload(px, abi.T_U8, abi.T_F32, abi.OP_MUL, 1.0 / 255.0)),
        tfs.TensorProto(dtype=1, float_val=tfs.Payload(px, abi.T_U8)),
        tfs.TensorProto(dtype=3, int_val=np.arange(-128, 128).astype(np.int8)),
        tfs.TensorProto(dtype=3, int_val=np.arange(0, 70000, 7).astype(np.uint16)),
        tfs.TensorProto(dtype=3, int_val=np.arange(-32768, 32768, 5).astype(np.int16)),
        tfs.TensorProto(dtype=19, half_val=x.astype(np.float16)),
    ]
    want = ob.encode_messages(msgs, abi.MSG_TENSOR_PROTO)
    assert_bytes_equal(gpu_encode_both(codec, msgs, abi.MSG_TENSOR_PROTO), want, "fused casts")


@pytest.mark.parametrize("flags", [0, abi.F_ALIGN_PAYLOAD])
def test_staged_casts_alignment_and_tile_edges(codec, flags):
    """The staged cast kernel (u8 -> f32 x1 / div / mul, f32 -> f16 / bf16, and on decode f16 / bf16 -> f32): element
    counts around its 32 KiB output tiles, sources that start at every element phase inside a 16-byte word, destinations at
    whatever byte phase the framing in front of them gives (names of different lengths), with and without payload
    alignment; bytes against the oracle, decoded values (widened with TFS_F_CONTENT_HALF_TO_F32) too."""
    rng = np.random.default_rng(777)
    base_f = np.concatenate([workloads.c4_payload(70_000), np.array([np.inf, -np.inf, np.nan, -0.0, 65520.0, 1e-8], np.float32),
                             np.array([0x7FA00001, 0xFFC00001, 0x00000001, 0x33000001], np.uint32).view(np.float32)])
    base_u = rng.integers(0, 256, 80_000, dtype=np.uint8)
    msgs = []
    sizes = [1024, 1025, 4095, 8191, 8192, 8193, 16383, 16384, 16385, 16384 * 2 + 3, 50_001]
    for i, n in enumerate(sizes):
        fo, uo = i % 4, i % 16            # element phase of the source inside a 16-byte word
        xf, xu = base_f[fo: fo + n], base_u[uo: uo + n]
        msgs.append(tfs.PredictRequest(tfs.ModelSpec("m", signature_name="s"), [
            ("a" * (i + 1), tfs.TensorProto(dtype=14, tensor_content=tfs.Payload(xf, abi.T_F32, abi.T_BF16))),
            ("b" * (i % 5), tfs.TensorProto(dtype=19, tensor_content=tfs.Payload(xf, abi.T_F32, abi.T_F16))),
            ("c", tfs.TensorProto(dtype=1, float_val=tfs.Payload(xu, abi.T_U8, abi.T_F32, abi.OP_DIV, 255.0))),
            ("d" * (i % 3 + 1), tfs.TensorProto(dtype=1, float_val=tfs.Payload(xu, abi.T_U8, abi.T_F32, abi.OP_MUL, 0.5))),
            ("e", tfs.TensorProto(dtype=1, float_val=tfs.Payload(xu, abi.T_U8)))]))
    # one field per message too: with TFS_F_ALIGN_PAYLOAD the destination is then 16-byte aligned (the bulk-store flush)
    for n in (8192, 16384 + 5, 100_000):
        msgs.append(tfs.PredictRequest(None, [("x", tfs.TensorProto(dtype=14, tensor_content=tfs.Payload(np.resize(base_f, n), abi.T_F32, abi.T_BF16)))]))
        msgs.append(tfs.PredictRequest(None, [("y", tfs.TensorProto(dtype=1, float_val=tfs.Payload(np.resize(base_u, n), abi.T_U8, abi.T_F32, abi.OP_DIV, 255.0)))]))
    kind = abi.MSG_PREDICT_REQUEST
    want = ob.encode_messages(msgs, kind)
    got = gpu_encode_both(codec, msgs, kind, flags)
    assert_bytes_equal(got, want, "staged casts encode")
    dflags = abi.F_CONTENT_HALF_TO_F32
    st, dec = gpu_decode_both(codec, got, kind, dflags)
    st_o, dec_o = ob.decode_bytes(want, kind, dflags)
    assert st == st_o == 0
    helpers.assert_same_decoded(dec_o, dec, "request", "staged casts decode (widened)")


def test_u8_scale_by_any_divisor_is_the_ieee_operation(codec):
    """u8 -> f32 with `/ p` is computed with a reciprocal and two fused multiply-adds only when the kernel has checked that
    this gives the bits of the IEEE division for all 256 inputs, and looked up in a table of true quotients otherwise; `* p`
    is the multiplication itself.  Divisors and factors of every kind (including ones where the shortcut is NOT exact, zero,
    denormals, infinities and NaN) against the oracle's plain `x / p`, `x * p`."""
    all_bytes = np.tile(np.arange(256, dtype=np.uint8), 9)[:2100]  # every input value, more than the staged kernel's minimum
    params = [255.0, 3.0, 7.1, 127.5, 1.0, 0.1, 1e-30, 1e-40, 3.4e38, 0.0, -255.0, float("inf"), float("nan"), 2.0 ** -126, 65535.0, 0.00392156862745098]
    msgs = [tfs.PredictRequest(None, [
        ("d", tfs.TensorProto(dtype=1, float_val=tfs.Payload(all_bytes, abi.T_U8, abi.T_F32, abi.OP_DIV, p))),
        ("m", tfs.TensorProto(dtype=1, float_val=tfs.Payload(all_bytes[3:], abi.T_U8, abi.T_F32, abi.OP_MUL, p)))]) for p in params]
    kind = abi.MSG_PREDICT_REQUEST
    assert_bytes_equal(gpu_encode_both(codec, msgs, kind), ob.encode_messages(msgs, kind), "u8 scale")


def test_invalid_descriptors_are_rejected(codec):
    from tfs_b200._lib import TfsError
    enc = codec.encoder()
    bad = [
        tfs.TensorProto(int64_val=np.zeros(3, np.float32)),
        tfs.TensorProto(float_val=tfs.Payload(np.zeros(3, np.float32), abi.T_F32, abi.T_F32, abi.OP_DIV, 2.0)),
        tfs.TensorProto(double_val=np.zeros(3, np.float32)),
    ]
    for t in bad:
        with pytest.raises(TfsError) as e:
            enc.encode_host([t], abi.MSG_TENSOR_PROTO)
        assert e.value.status == abi.TFS_ERR_INVALID_ARG
    # a destination that is too small is refused and nothing is written past it
    t = tfs.TensorProto(dtype=1, float_val=np.ones(10_000, np.float32))
    dst = np.full(1000, 0xAB, np.uint8)
    with pytest.raises(TfsError) as e:
        enc.encode_host([t], abi.MSG_TENSOR_PROTO, dst=dst)
    assert e.value.status == abi.TFS_ERR_NOMEM and (dst == 0xAB).all()
    # and the context keeps working afterwards
    assert enc.encode_host([t], abi.MSG_TENSOR_PROTO).all_bytes() == ob.encode_messages([t], abi.MSG_TENSOR_PROTO)


# ---- decode ----------------------------------------------------------------------------------------
def gpu_decode_both(codec, blobs, kind, flags=0):
    import torch
    offs, lens, cur = [], [], 0
    for b in blobs:
        offs.append(cur)
        lens.append(len(b))
        cur += len(b)
    wire = np.frombuffer(b"".join(blobs), dtype=np.uint8) if cur else np.zeros(0, np.uint8)
    dec = codec.decoder()
    host = dec.decode_host(wire, offs, lens, kind, flags)
    host_msgs = host.messages()
    # device path, with the wire at an odd device address
    dwire = torch.zeros(cur + 64, dtype=torch.uint8, device=f"cuda:{codec.ctx.device}")
    if cur:
        dwire[3: 3 + cur] = torch.from_numpy(wire.copy()).to(dwire.device)
    d = dec.decode(dwire[3: 3 + cur] if cur else dwire[:0], offs, lens, kind, flags)
    codec.ctx.sync()
    helpers.assert_same_decoded(host_msgs, d.messages(), helpers_kind(kind), "host vs device decode")
    # one-call forms: device pool of exactly the needed size, and pinned host buffers used in place
    pool = torch.empty(max(16, d.pool_bytes), dtype=torch.uint8, device=dwire.device)
    d2 = dec.decode_into(dwire[3: 3 + cur] if cur else dwire[:0], offs, lens, kind, pool, flags)
    codec.ctx.sync()  # the device-pool forms are asynchronous
    helpers.assert_same_decoded(host_msgs, d2.messages(), helpers_kind(kind), "two-call vs one-call decode")
    pwire = torch.empty(max(1, cur), dtype=torch.uint8).pin_memory()
    pwire.numpy()[:cur] = wire
    ppool = torch.zeros(max(16, d.pool_bytes), dtype=torch.uint8).pin_memory()
    d3 = dec.decode_host_into(pwire.numpy()[:cur], offs, lens, kind, ppool.numpy(), flags)
    helpers.assert_same_decoded(host_msgs, d3.messages(), helpers_kind(kind), "pageable vs pinned (in place) host decode")
    assert d2.status == d3.status == host.status
    return host.status, host_msgs


def helpers_kind(kind):
    return {abi.MSG_TENSOR_PROTO: "tensor", abi.MSG_PREDICT_REQUEST: "request", abi.MSG_PREDICT_RESPONSE: "response"}[kind]


@pytest.mark.parametrize("kind_name", ["tensor", "request", "response"])
def test_decode_golden_canonical(codec, kind_name):
    cases = [c for c in G["encode"] if c["kind"] == kind_name]
    blobs = [bytes.fromhex(c["wire"]) for c in cases]
    kind = helpers.KIND_OF[kind_name]
    st, got = gpu_decode_both(codec, blobs, kind)
    st_o, want = ob.decode_bytes(blobs, kind)
    assert st == st_o == 0
    helpers.assert_same_decoded(want, got, kind_name, "golden canonical")


@pytest.mark.parametrize("kind_name", ["tensor", "request", "response"])
def test_decode_golden_noncanonical(codec, kind_name):
    cases = [c for c in G["decode"] if c["kind"] == kind_name]
    blobs = [bytes.fromhex(c["wire"]) for c in cases]
    kind = helpers.KIND_OF[kind_name]
    st, got = gpu_decode_both(codec, blobs, kind)
    assert st == 0
    for c, m in zip(cases, got):
        assert m.status == 0, c["name"]
        assert helpers.message_to_desc(kind_name, m.message) == c["expect"], c["name"]


def test_decode_error_table(codec):
    for kind_name in ("tensor", "request", "response"):
        cases = [c for c in G["errors"] if c["kind"] == kind_name]
        blobs = [bytes.fromhex(c["wire"]) for c in cases]
        st, got = gpu_decode_both(codec, blobs, helpers.KIND_OF[kind_name])
        assert st == abi.TFS_ERR_BATCH
        for c, m in zip(cases, got):
            assert m.status == c["code"], (c["name"], abi.STATUS_NAMES.get(m.status), abi.STATUS_NAMES[c["code"]])


def test_decode_bad_message_fails_alone(codec):
    good = ob.encode_messages([tfs.TensorProto(dtype=3, int_val=np.arange(5000, dtype=np.int32))], abi.MSG_TENSOR_PROTO)[0]
    long_run = b"\x3a\x0c" + b"\x80" * 11 + b"\x01"
    blobs = [good, b"\x3a\x05\x01", good, long_run, b"", good]
    st, got = gpu_decode_both(codec, blobs, abi.MSG_TENSOR_PROTO)
    st_o, want = ob.decode_bytes(blobs, abi.MSG_TENSOR_PROTO)
    assert st == st_o == abi.TFS_ERR_BATCH
    helpers.assert_same_decoded(want, got, "tensor", "mixed good/bad")
    assert [m.status for m in got] == [0, abi.TFS_ERR_BUFFER_UNDERFLOW, 0, abi.TFS_ERR_INVALID_VARINT, 0, 0]


def test_decode_plan_outgrows_its_first_tables():
    """A small batch is planned optimistically from the context's hints (k_plan_small); one message with far more
    tensors, spans, filters and key bytes than a fresh context guesses has to take the grow-and-retry path, and the
    next, ordinary batch on the same context must still be right."""
    fresh = tfs.GpuCodec(tfs.GpuContext(0))
    try:
        rng = np.random.default_rng(77)
        outs = [("output_with_a_rather_long_name_%04d_" % i + "x" * 40,
                 tfs.TensorProto(dtype=3, tensor_shape=tfs.TensorShapeProto([3]), int_val=rng.integers(-5, 5, 3, dtype=np.int32),
                                 float_val=rng.standard_normal(2).astype(np.float32)))
                for i in range(150)]
        big = tfs.PredictRequest(tfs.ModelSpec("m"), outs, output_filter=["f%d" % i for i in range(90)])
        for msgs in ([big], [workloads.c1_request()], [big, workloads.c1_request(), big]):
            blobs = ob.encode_messages(msgs, abi.MSG_PREDICT_REQUEST)
            st, got = gpu_decode_both(fresh, blobs, abi.MSG_PREDICT_REQUEST)
            st_o, want = ob.decode_bytes(blobs, abi.MSG_PREDICT_REQUEST)
            assert st == st_o == 0
            helpers.assert_same_decoded(want, got, "request", "plan retry")
    finally:
        fresh.ctx.close()


def test_decode_reports_the_first_error_in_stream_order(codec):
    """An over-long varint INSIDE a packed body comes before a structural error later in the message: prost fails at the
    first one (found by a longer fuzz run: the walk used to report the later one)."""
    long_run = b"\x3a\x0c" + b"\x80" * 11 + b"\x01"
    blobs = [long_run + b"\x2a\x05\x00",                     # ... then float_val longer than the buffer
             b"\x08\x01" + long_run + b"\x28\x01",            # ... then a wrong wire type
             b"\x3a\x02\x01\x02" + long_run + b"\x0a",        # a good body, the bad one, then a truncated key/len
             b"\x2a\x09\x00" + long_run[:4]]                 # the structural error comes first here
    st, got = gpu_decode_both(codec, blobs, abi.MSG_TENSOR_PROTO)
    st_o, want = ob.decode_bytes(blobs, abi.MSG_TENSOR_PROTO)
    assert [m.status for m in got] == [m.status for m in want]
    assert [m.status for m in got] == [abi.TFS_ERR_INVALID_VARINT] * 3 + [abi.TFS_ERR_BUFFER_UNDERFLOW]  # (len 9 > 4 bytes left)


def test_decode_long_varint_runs_inside_big_span(codec):
    # a >10-byte varint deep inside a multi-tile packed body is only visible to the terminator kernel
    for pos in (5, 4090, 4096, 8191, 20000):
        body = bytearray(b"\x01" * 30000)
        body[pos: pos + 11] = b"\xff" * 11
        ok_body = bytearray(body)
        ok_body[pos + 9] = 0x01  # 10-byte varint: legal
        blobs = [b"\x52" + helpers_varint(len(body)) + bytes(body), b"\x52" + helpers_varint(len(ok_body)) + bytes(ok_body)]
        st, got = gpu_decode_both(codec, blobs, abi.MSG_TENSOR_PROTO)
        st_o, want = ob.decode_bytes(blobs, abi.MSG_TENSOR_PROTO)
        helpers.assert_same_decoded(want, got, "tensor", f"run at {pos}")
        assert got[0].status == abi.TFS_ERR_INVALID_VARINT and got[1].status == 0


def test_string_run_edges_decode_like_the_oracle(codec):
    """The walk takes a run of string_val elements with one-byte lengths in a tight loop; whatever ends the run - the
    tensor's end, a longer length, another field, an element that overshoots the tensor - goes the general way and
    must come out exactly as prost has it."""
    el = lambda b: b"\x42" + helpers_varint(len(b)) + b  # noqa: E731
    run = b"".join(el(bytes([65 + i]) * i) for i in range(20))
    def entry(tensor, key=b"k"):
        body = b"\x0a" + helpers_varint(len(key)) + key + b"\x12" + helpers_varint(len(tensor)) + tensor
        return b"\x12" + helpers_varint(len(body)) + body
    tensors = [
        b"\x08\x07" + run,                                              # the run ends with the tensor
        b"\x08\x07" + run + b"\x42",                                    # a key with nothing behind it
        b"\x08\x07" + run + b"\x42\x05ab",                               # the last element overshoots
        b"\x08\x07" + run + el(b"x" * 127) + el(b"y" * 128) + el(b"z" * 3),  # a two-byte length inside the run
        b"\x08\x07" + run + b"\x42\x80\x00" + el(b"after"),              # a non-canonical zero length
        b"\x08\x07" + run + b"\x3a\x02\x01\x02" + run,                   # another field between two runs
        b"\x08\x07" + run + b"\x45\x00\x00\x00\x00",                    # field 8 with a wrong wire type
        run[:-3],                                                        # cut inside the last element
    ]
    st, got = gpu_decode_both(codec, tensors, abi.MSG_TENSOR_PROTO)
    st_o, want = ob.decode_bytes(tensors, abi.MSG_TENSOR_PROTO)
    assert [m.status for m in got] == [m.status for m in want]
    helpers.assert_same_decoded(want, got, "tensor", "string run edges (tensor)")
    assert [m.status for m in got][:2] == [0, abi.TFS_ERR_INVALID_VARINT]
    # inside a request: what follows the tensor begins with the same byte (an unknown field 8 of the request), and an
    # element that overshoots its tensor but not the request is "delimited length exceeded"
    reqs = [entry(t) + b"\x42\x02hi" + entry(tensors[0], b"second") + b"\x1a\x01f" for t in tensors]
    reqs.append(entry(b"\x08\x07" + run + b"\x42\x05ab") + b"\x42\x03abc")
    st, got = gpu_decode_both(codec, reqs, abi.MSG_PREDICT_REQUEST)
    st_o, want = ob.decode_bytes(reqs, abi.MSG_PREDICT_REQUEST)
    assert [m.status for m in got] == [m.status for m in want]
    helpers.assert_same_decoded(want, got, "request", "string run edges (request)")
    assert got[0].status == 0 and got[-1].status == abi.TFS_ERR_LENGTH_EXCEEDED


def helpers_varint(v):
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def test_decode_flags_widen_half(codec):
    x = workloads.c4_payload(10_000)
    for dt, ty in ((19, abi.T_F16), (14, abi.T_BF16)):
        msgs = [tfs.TensorProto(dtype=dt, half_val=tfs.Payload(x, abi.T_F32, ty)),
                tfs.TensorProto(dtype=dt, tensor_content=tfs.Payload(x, abi.T_F32, ty)),
                tfs.TensorProto(dtype=1, half_val=np.arange(100, dtype=np.int32)),
                tfs.TensorProto(dtype=dt, tensor_content=np.arange(7, dtype=np.uint8))]
        blobs = ob.encode_messages(msgs, abi.MSG_TENSOR_PROTO)
        flags = abi.F_HALF_TO_F32 | abi.F_CONTENT_HALF_TO_F32
        st, got = gpu_decode_both(codec, blobs, abi.MSG_TENSOR_PROTO, flags)
        st_o, want = ob.decode_bytes(blobs, abi.MSG_TENSOR_PROTO, flags)
        assert st == st_o == 0
        for g, w in zip(got, want):
            for name in ("half_val", "tensor_content"):
                a, b = getattr(g.message, name), getattr(w.message, name)
                assert (a is None) == (b is None)
                if a is not None:
                    assert a.dtype == b.dtype and a.tobytes() == b.tobytes(), name


def test_grpc_prefix_roundtrip(codec):
    msgs = workloads.c5_responses(40, seed=21, max_log2=10)
    blobs = gpu_encode_both(codec, msgs, abi.MSG_PREDICT_RESPONSE, abi.F_GRPC_PREFIX)
    assert_bytes_equal(blobs, ob.encode_messages(msgs, abi.MSG_PREDICT_RESPONSE, abi.F_GRPC_PREFIX))
    st, got = gpu_decode_both(codec, blobs, abi.MSG_PREDICT_RESPONSE, abi.F_GRPC_PREFIX)
    st_o, want = ob.decode_bytes(blobs, abi.MSG_PREDICT_RESPONSE, abi.F_GRPC_PREFIX)
    assert st == st_o == 0
    helpers.assert_same_decoded(want, got, "response", "grpc")
    bad = [b"\x01" + blobs[0][1:], blobs[1][:3], blobs[2]]
    st, got = gpu_decode_both(codec, bad, abi.MSG_PREDICT_RESPONSE, abi.F_GRPC_PREFIX)
    st_o, want = ob.decode_bytes(bad, abi.MSG_PREDICT_RESPONSE, abi.F_GRPC_PREFIX)
    helpers.assert_same_decoded(want, got, "response", "grpc bad")


@pytest.mark.parametrize("kind", ["response", "request"])
def test_decode_fuzz_against_oracle(codec, kind):
    """Random truncations / byte flips of valid messages: same status class, same values when OK.  The corpus is the one
    tests/test_oracle_upb_messages.py runs through upb for accept / reject agreement (helpers.mutation_corpus)."""
    import os
    blobs = helpers.mutation_corpus(kind, int(os.environ.get("TFS_FUZZ_CASES", "400")), int(os.environ.get("TFS_FUZZ_SEED", "99")))
    mk = helpers.KIND_OF[kind]
    st_o, want = ob.decode_bytes(blobs, mk)
    st, got = gpu_decode_both(codec, blobs, mk)
    mism = []
    for i, (w, g) in enumerate(zip(want, got)):
        if w.status != g.status:
            mism.append((i, blobs[i].hex(), abi.STATUS_NAMES.get(w.status), abi.STATUS_NAMES.get(g.status)))
        elif w.status == 0:
            helpers.assert_same_decoded([w], [g], kind, f"fuzz {i}: {blobs[i].hex()}")
    assert not mism, mism[:5]


# ---- string_val: repeated bytes -------------------------------------------------
def _strings(rng, sizes):
    return [rng.integers(0, 256, int(n), dtype=np.uint8).tobytes() for n in sizes]


@pytest.mark.parametrize("shape", ["one_empty", "tiny_2047", "tiny_2048", "tiny_2049", "tiny_5000", "mixed_sizes", "big_1MiB", "edges_16k"])
def test_string_val_roundtrip(codec, shape):
    rng = np.random.default_rng(hash(shape) % 1000)
    sizes = {
        "one_empty": [0],
        "tiny_2047": rng.integers(0, 12, 2047), "tiny_2048": rng.integers(0, 12, 2048), "tiny_2049": rng.integers(0, 12, 2049),
        "tiny_5000": rng.integers(0, 40, 5000),
        "mixed_sizes": [0, 1, 127, 128, 129, 16383, 16384, 16385, 70000, 3, 0, 0, 5, 200000, 15, 16, 17],
        "big_1MiB": [1 << 20, 1, (1 << 20) + 7],
        "edges_16k": [16384 - 2, 16384 - 3, 16384 - 1, 2 * 16384, 1],
    }[shape]
    strs = _strings(rng, sizes)
    t = tfs.TensorProto(dtype=7, tensor_shape=tfs.TensorShapeProto([len(strs)]), string_val=strs)
    mixed = tfs.TensorProto(dtype=7, int_val=np.arange(5, dtype=np.int32), string_val=strs[:50],
                            scomplex_val=rng.standard_normal(6).astype(np.float32), tensor_content=np.arange(33, dtype=np.uint8))
    roundtrip_check(codec, [t, mixed, t], abi.MSG_TENSOR_PROTO, f"string_val {shape}")
    req = tfs.PredictRequest(tfs.ModelSpec("text", signature_name="serving_default"),
                             [("examples", t), ("x", tfs.TensorProto(dtype=1, float_val=rng.standard_normal(9000).astype(np.float32)))],
                             output_filter=["y"])
    roundtrip_check(codec, [req, req], abi.MSG_PREDICT_REQUEST, f"string_val request {shape}", abi.F_GRPC_PREFIX)


def test_string_val_decoded_block_encodes_again(codec):
    """The decoder writes string_val as the same TFS_T_RAGGED block the encoder takes: feed it straight back (on the device)."""
    import torch
    rng = np.random.default_rng(5)
    strs = _strings(rng, rng.integers(0, 300, 700))
    wire = ob.encode_messages([tfs.TensorProto(dtype=7, string_val=strs)], abi.MSG_TENSOR_PROTO)[0]
    dwire = torch.from_numpy(np.frombuffer(wire, dtype=np.uint8).copy()).cuda()
    d = codec.decoder().decode(dwire, [0], [len(wire)], abi.MSG_TENSOR_PROTO)
    codec.ctx.sync()
    to = d.view.tensors[0]
    slot = abi.SLOT_KINDS.index(abi.KIND_STRING_VAL)
    n, off = int(to.n_elems[slot]), int(to.dst_off[slot])
    assert n == len(strs) and int(to.dst_type[slot]) == abi.T_RAGGED
    total = sum(len(x) for x in strs)
    block = d.pool[off: off + 8 * (n + 1) + total]
    again = tfs.TensorProto(dtype=7, string_val=tfs.Ragged(block, n, total))
    eb = codec.encoder().encode([again], abi.MSG_TENSOR_PROTO)
    assert eb.all_bytes() == [wire]


def test_string_val_descriptor_must_match_memory(codec):
    blk = tfs.ragged_block([b"abc", b"de"])
    bad_total = tfs.TensorProto(dtype=7, string_val=tfs.Ragged(blk, 2, 6))  # the offsets end at 5
    with pytest.raises(tfs.TfsError) as e:
        codec.encoder().encode_host([bad_total], abi.MSG_TENSOR_PROTO)
    assert e.value.status == abi.TFS_ERR_INVALID_ARG
    desc = blk.copy()
    desc[8:16] = np.array([4], dtype="<u8").view(np.uint8)   # offsets 0, 4, 5 -> fine; then make them descend
    desc[16:24] = np.array([2], dtype="<u8").view(np.uint8)
    with pytest.raises(tfs.TfsError):
        codec.encoder().encode_host([tfs.TensorProto(dtype=7, string_val=tfs.Ragged(desc, 2, 2))], abi.MSG_TENSOR_PROTO)
    # and the context is still usable
    assert codec.encoder().encode_host([tfs.TensorProto(dtype=7, string_val=[b"ok"])], abi.MSG_TENSOR_PROTO).all_bytes() == [bytes.fromhex("0807" "42026f6b")]


# ---- Example / Features / Input framing, encode only ---------------------------------
@pytest.mark.parametrize("kind_name", ["input", "classify"])
def test_input_golden_vectors(codec, kind_name):
    cases = [c for c in G["input"] if c["kind"] == kind_name]
    msgs = [helpers.input_from_desc(kind_name, c["msg"]) for c in cases]
    got = gpu_encode_both(codec, msgs, helpers.KIND_OF[kind_name])
    for c, g in zip(cases, got):
        assert_bytes_equal([g], [bytes.fromhex(c["wire"])], c["name"])


def _random_input(rng, n_examples, max_feats, max_vals, with_defaults=True):
    exs = []
    for _ in range(n_examples):
        feats = []
        for j in range(int(rng.integers(0, max_feats + 1))):
            which = int(rng.integers(0, 4 if with_defaults else 3))
            n = int(rng.integers(0, max_vals + 1))
            name = "" if (with_defaults and rng.integers(0, 9) == 0) else "feature_%d" % j
            if which == 0:
                ft = tfs.Feature(int64_list=rng.integers(-(1 << 50), 1 << 50, n, dtype=np.int64))
            elif which == 1:
                ft = tfs.Feature(float_list=rng.standard_normal(n).astype(np.float32))
            elif which == 2:
                ft = tfs.Feature(bytes_list=_strings(rng, rng.integers(0, 60, n % 9)))
            else:
                ft = tfs.Feature()  # kind: None
            feats.append((name, ft))
        exs.append(feats)
    return tfs.Input(exs)


@pytest.mark.parametrize("flags", [0, abi.F_GRPC_PREFIX, abi.F_ALIGN_PAYLOAD])
def test_input_random_batches_match_the_oracle(codec, flags):
    rng = np.random.default_rng(31 + flags)
    inputs = [_random_input(rng, int(rng.integers(0, 12)), 5, 300) for _ in range(40)]
    got = gpu_encode_both(codec, inputs, abi.MSG_INPUT, flags)
    assert_bytes_equal(got, ob.encode_messages(inputs, abi.MSG_INPUT, flags & abi.F_GRPC_PREFIX), "Input batch")
    reqs = [tfs.ClassificationRequest(tfs.ModelSpec("m%d" % i, version=i if i % 3 else None, signature_name="classify" if i % 2 else ""), inp)
            for i, inp in enumerate(inputs)]
    got = gpu_encode_both(codec, reqs, abi.MSG_CLASSIFY_REQUEST, flags)
    assert_bytes_equal(got, ob.encode_messages(reqs, abi.MSG_CLASSIFY_REQUEST, flags & abi.F_GRPC_PREFIX), "ClassificationRequest batch")


def test_input_large_lists_and_many_messages(codec):
    rng = np.random.default_rng(8)
    big = tfs.Input([[("ids", tfs.Feature(int64_list=rng.integers(0, 30522, 300_000, dtype=np.int64))),
                      ("image", tfs.Feature(bytes_list=_strings(rng, [200_000, 0, 70_001]))),
                      ("x", tfs.Feature(float_list=rng.standard_normal(1_000_000).astype(np.float32)))],
                     [("ids", tfs.Feature(int64_list=rng.integers(-9, 9, 5000, dtype=np.int64)))]])
    got = gpu_encode_both(codec, [big], abi.MSG_INPUT, abi.F_ALIGN_PAYLOAD)
    assert_bytes_equal(got, ob.encode_messages([big], abi.MSG_INPUT), "big Input")
    many = [_random_input(rng, 2, 3, 20, with_defaults=False) for _ in range(700)]   # > 256 messages: the multi-kernel framing path
    got = gpu_encode_both(codec, many, abi.MSG_INPUT)
    assert_bytes_equal(got, ob.encode_messages(many, abi.MSG_INPUT), "700 Inputs")


def test_input_descriptor_errors(codec):
    enc = codec.encoder()
    bad_order = tfs.build_batch([tfs.Input([[("a", tfs.Feature())], [("b", tfs.Feature())]])], abi.MSG_INPUT)
    bad_order.tensors[0].version_number = 1
    bad_order.tensors[1].version_number = 0   # Example indices must ascend
    with pytest.raises(tfs.TfsError) as e:
        enc.encode_host(bad_order, abi.MSG_INPUT)
    assert e.value.status == abi.TFS_ERR_INVALID_ARG
    mixed = tfs.build_batch([tfs.TensorProto(dtype=1, float_val=np.ones(3, np.float32))], abi.MSG_TENSOR_PROTO)
    mixed.fields[0].kind = abi.KIND_FLOAT_LIST   # a list kind outside an Input
    with pytest.raises(tfs.TfsError):
        enc.encode_host(mixed, abi.MSG_TENSOR_PROTO)
    with pytest.raises(tfs.TfsError):            # decode of Input is not provided: the reference only builds Inputs
        codec.decoder().decode_host(np.zeros(4, np.uint8), [0], [4], abi.MSG_INPUT)


# ---- batch assembly: several source arrays, one field on the wire ----------------------
def test_same_kind_descriptors_are_concatenated(codec):
    rng = np.random.default_rng(12)
    imgs = [rng.integers(0, 256, 3 * 17 * 17 + i, dtype=np.uint8) for i in range(5)]  # five "images" of slightly different size
    n = sum(x.size for x in imgs)
    fused = tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([n]),
                            float_val=tuple(tfs.Payload(x, abi.T_U8, abi.T_F32, abi.OP_DIV, 255.0) for x in imgs))
    plain = tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([n]),
                            float_val=(np.concatenate(imgs).astype(np.float32) / np.float32(255.0)).astype(np.float32))
    want = ob.encode_messages([plain], abi.MSG_TENSOR_PROTO)
    assert ob.encode_messages([fused], abi.MSG_TENSOR_PROTO) == want          # the oracle agrees with itself
    assert_bytes_equal(gpu_encode_both(codec, [fused], abi.MSG_TENSOR_PROTO), want, "u8 images -> one float_val")
    # every class: varint (with an empty part and parts that cross tile edges), fixed, raw bytes, strings; big parts too
    parts64 = (rng.integers(-9, 1 << 33, 3000, dtype=np.int64), np.zeros(0, np.int64), rng.integers(0, 99, 2049, dtype=np.int64))
    t = tfs.TensorProto(dtype=9, int64_val=parts64, double_val=(rng.standard_normal(7), rng.standard_normal(100_000)),
                        tensor_content=(rng.integers(0, 256, 70_001, dtype=np.uint8), rng.integers(0, 256, 5, dtype=np.uint8)),
                        half_val=(np.arange(5, dtype=np.int32), np.arange(40000, 40100, dtype=np.int32)))
    flat = tfs.TensorProto(dtype=9, int64_val=np.concatenate(parts64), double_val=np.concatenate(t.double_val),
                           tensor_content=np.concatenate(t.tensor_content), half_val=np.concatenate(t.half_val))
    for flags in (0, abi.F_ALIGN_PAYLOAD):
        req = [tfs.PredictRequest(tfs.ModelSpec("m"), [("a", t), ("b", fused)]), tfs.PredictRequest(None, [("c", t)])]
        req_flat = [tfs.PredictRequest(tfs.ModelSpec("m"), [("a", flat), ("b", plain)]), tfs.PredictRequest(None, [("c", flat)])]
        want = ob.encode_messages(req_flat, abi.MSG_PREDICT_REQUEST)
        assert_bytes_equal(gpu_encode_both(codec, req, abi.MSG_PREDICT_REQUEST, flags), want, f"concatenated parts (flags {flags})")


# ---- BASELINE configs ---------------------------------------------------------------------------------
def roundtrip_check(codec, msgs, kind, ctx, flags=0):
    want = ob.encode_messages(msgs, kind, flags & abi.F_GRPC_PREFIX)
    got = gpu_encode_both(codec, msgs, kind, flags)
    assert_bytes_equal(got, want, ctx + " encode")
    st, dec = gpu_decode_both(codec, got, kind, flags & abi.F_GRPC_PREFIX)
    st_o, dec_o = ob.decode_bytes(want, kind, flags & abi.F_GRPC_PREFIX)
    assert st == st_o == 0
    helpers.assert_same_decoded(dec_o, dec, helpers_kind(kind), ctx + " decode")


def test_config_c1(codec):
    roundtrip_check(codec, [workloads.c1_request()] * 3, abi.MSG_PREDICT_REQUEST, "C1 request")
    roundtrip_check(codec, [workloads.c1_response()] * 3, abi.MSG_PREDICT_RESPONSE, "C1 response")
    fused = workloads.c1_image_request(fused=True)
    plain = workloads.c1_image_request(fused=False)
    a = codec.encoder().encode_host([fused], abi.MSG_PREDICT_REQUEST).all_bytes()
    assert_bytes_equal(a, ob.encode_messages([plain], abi.MSG_PREDICT_REQUEST), "C1 image (u8 -> f32 / 255 fused)")
    roundtrip_check(codec, [workloads.c1_request()] * 5000, abi.MSG_PREDICT_REQUEST, "C1 x5000")


def test_async_encode_is_complete_in_stream_order(codec):
    """TFS_F_ASYNC: tfs_encode returns when offsets and lengths are known, the payload kernels may still be running.  The
    bytes must be what the synchronous call writes once the stream has been waited for, and a decode queued right behind
    it (no host wait in between) must see them."""
    import torch
    rng = np.random.default_rng(2718)
    dev = torch.device(f"cuda:{codec.ctx.device}")
    msgs = [tfs.PredictRequest(tfs.ModelSpec("m", signature_name="s"), [
        ("x", tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([6_000_000 + i]), float_val=rng.standard_normal(6_000_000 + i).astype(np.float32))),
        ("ids", tfs.TensorProto(dtype=9, int64_val=rng.integers(-5, 1 << 33, 900_001, dtype=np.int64))),
        ("s", tfs.TensorProto(dtype=7, string_val=[b"abc", b"", b"z" * 5000]))]) for i in range(6)]
    kind = abi.MSG_PREDICT_REQUEST
    want = ob.encode_messages(msgs, kind)
    dmsgs = workloads.to_device(msgs, kind, dev)
    enc, dec = codec.encoder(), codec.decoder()
    built = tfs.build_batch(dmsgs, kind)
    for flags in (abi.F_ASYNC, abi.F_ASYNC | abi.F_ALIGN_PAYLOAD):
        wire = torch.full((enc.bound(built, flags) + 64,), 0x5A, dtype=torch.uint8, device=dev)
        pool = torch.empty(sum(len(w) for w in want) * 2 + (1 << 20), dtype=torch.uint8, device=dev)
        for _ in range(3):
            eb = enc.encode(built, kind, flags, dst=wire)
            d = dec.decode_into(eb.wire, eb.offsets, eb.lengths, kind, pool)  # queued behind the encode on the same stream
            codec.ctx.sync()
            assert_bytes_equal(eb.all_bytes(), want, f"async encode (flags {flags})")
            st_o, dec_o = ob.decode_bytes(want, kind)
            helpers.assert_same_decoded(dec_o, d.messages(), "request", "decode behind an async encode")
            d.close()
            wire.fill_(0x5A)


def test_many_small_messages_compile_on_several_threads(codec):
    """Above ~10^5 descriptors (messages + tensors + fields) the host compiles them on several threads (csrc/tfs_encode.cu compile_batch): the
    names of every range land in their own block and are joined afterwards.  Distinct names, one to three tensors, some
    filters and version choices per message, against the oracle byte for byte; then a descriptor error in a late message
    must come back as the one-thread walk reports it."""
    from tfs_b200._lib import TfsError
    rng = np.random.default_rng(811)
    msgs = []
    for i in range(30_000):
        nt = 1 + i % 3
        tensors = [("in_%d_%d" % (i, k) + "y" * (i % 7),
                    tfs.TensorProto(dtype=abi.DT_FLOAT if k != 1 else abi.DT_INT64, tensor_shape=tfs.TensorShapeProto([2 + k]),
                                    float_val=rng.standard_normal(2 + k).astype(np.float32) if k != 1 else None,
                                    int64_val=rng.integers(-3, 1 << 20, 2 + k, dtype=np.int64) if k == 1 else None))
                   for k in range(nt)]
        spec = None if i % 11 == 0 else tfs.ModelSpec("model_%d" % (i % 97), version=i if i % 5 == 0 else None,
                                                     version_label="stable" if i % 5 == 1 else None, signature_name="sig%d" % (i % 3))
        msgs.append(tfs.PredictRequest(spec, tensors, output_filter=["out%d" % i] if i % 13 == 0 else []))
    roundtrip_check(codec, msgs, abi.MSG_PREDICT_REQUEST, "30 000 small requests")
    built = tfs.build_batch(msgs, abi.MSG_PREDICT_REQUEST)
    built.msgs[17_000].version_choice = 9
    built.msgs[29_000].kind = 77
    with pytest.raises(TfsError) as e:
        codec.encoder().encode_host(built, abi.MSG_PREDICT_REQUEST)
    assert e.value.status == abi.TFS_ERR_INVALID_ARG and "version_choice" in str(e.value)


@pytest.mark.parametrize("variant", ["float_val", "tensor_content", "tensor_content_bf16"])
def test_config_c2_full_size(codec, variant):
    x = workloads.c2_payload()
    roundtrip_check(codec, [workloads.c2_request(x, variant)], abi.MSG_PREDICT_REQUEST, f"C2 {variant}", abi.F_ALIGN_PAYLOAD)


@pytest.mark.parametrize("flags", [0, abi.F_ALIGN_PAYLOAD])
def test_pinned_buffers_take_the_chunked_copy_engine_path(codec, flags):
    """Pinned sources / destinations above 8 MB are brought in by the copy engine in 4 MB chunks while the copy kernel
    follows chunk by chunk (one gated launch per chunk); pageable and device paths must give the same bytes and values."""
    rng = np.random.default_rng(4242)
    msgs = []
    for i in range(2):
        msgs.append(tfs.PredictRequest(tfs.ModelSpec("big", signature_name="serving_default"), [
            ("a", tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([3_000_001]), float_val=rng.standard_normal(3_000_001).astype(np.float32))),
            ("b", tfs.TensorProto(dtype=4, tensor_content=rng.integers(0, 256, 9_000_003 + i, dtype=np.uint8))),
            ("ids", tfs.TensorProto(dtype=9, int64_val=rng.integers(-5, 1 << 40, 1_500_000, dtype=np.int64))),
            ("small", tfs.TensorProto(dtype=2, double_val=rng.standard_normal(7))),
            ("c", tfs.TensorProto(dtype=2, double_val=rng.standard_normal(1_200_000)))]))
    roundtrip_check(codec, msgs, abi.MSG_PREDICT_REQUEST, "chunked pinned path", flags)


@pytest.mark.parametrize("flags", [0, abi.F_ALIGN_PAYLOAD])
def test_pageable_buffers_cross_the_pinned_ring(codec, flags):
    """Pageable payloads / wire / pools of 1 MB and more are copied through a two-slot pinned ring by helper threads
    (csrc/tfs_host.cu staged_h2d / staged_d2h): several ring turns, fields that end inside a slot, a pinned destination
    (the copy kernel then follows the chunks) and a pageable one; bytes against the oracle, values back."""
    import torch
    rng = np.random.default_rng(9091)
    msgs = [tfs.PredictRequest(tfs.ModelSpec("ring", signature_name="serving_default"), [
        ("x", tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([10_000_003 + i]), float_val=rng.standard_normal(10_000_003 + i).astype(np.float32))),
        ("ids", tfs.TensorProto(dtype=9, int64_val=rng.integers(-5, 1 << 33, 700_001, dtype=np.int64))),
        ("raw", tfs.TensorProto(dtype=4, tensor_content=rng.integers(0, 256, 1_048_577, dtype=np.uint8))),
        ("tiny", tfs.TensorProto(dtype=1, float_val=rng.standard_normal(3).astype(np.float32)))]) for i in range(3)]
    kind = abi.MSG_PREDICT_REQUEST
    want = ob.encode_messages(msgs, kind)
    enc, dec = codec.encoder(), codec.decoder()
    built = tfs.build_batch(msgs, kind)
    # pageable payloads -> pageable wire, and -> pinned wire
    got = enc.encode_host(built, kind, flags)
    assert_bytes_equal(got.all_bytes(), want, "pageable -> pageable")
    pdst = torch.empty(enc.bound(built, flags) + 16, dtype=torch.uint8).pin_memory()
    got_p = enc.encode_host(built, kind, flags, dst=pdst.numpy())
    assert_bytes_equal(got_p.all_bytes(), want, "pageable -> pinned")
    # pageable wire -> pageable pool, pinned pool
    st_o, dec_o = ob.decode_bytes(want, kind)
    wire = np.array(got.wire[: got.total], copy=True)
    d = dec.decode_host(wire, got.offsets, got.lengths, kind)
    pool = np.empty(d.pool_bytes + 64, np.uint8)
    d.close()
    for pool_buf in (pool, torch.empty(pool.size, dtype=torch.uint8).pin_memory().numpy()):
        pool_buf[:] = 0xEE
        d = dec.decode_host_into(wire, got.offsets, got.lengths, kind, pool_buf)
        helpers.assert_same_decoded(dec_o, d.messages(), "request", "pageable wire decode")
        d.close()


def test_chunked_pinned_decode_with_bad_messages_and_prefix(codec):
    """The chunked copy-engine path of tfs_decode_host_into (pinned wire >= 8 MB, walk reading host memory while the wire
    is still arriving) with the gRPC prefix and with malformed messages between good ones."""
    import torch
    rng = np.random.default_rng(77)
    good = [tfs.PredictResponse([("y", tfs.TensorProto(dtype=1, float_val=rng.standard_normal(1_500_000).astype(np.float32))),
                                 ("i", tfs.TensorProto(dtype=9, int64_val=rng.integers(-3, 1 << 20, 200_000, dtype=np.int64))),
                                 ("s", tfs.TensorProto(dtype=7, string_val=[b"abc" * 1000, b"", b"z" * 70000]))],
                                tfs.ModelSpec("m", version=i)) for i in range(3)]
    for flags in (0, abi.F_GRPC_PREFIX):
        blobs = ob.encode_messages(good, abi.MSG_PREDICT_RESPONSE, flags)
        bad1 = blobs[1][: len(blobs[1]) // 2]                      # truncated in the middle of a payload
        bad2 = bytearray(blobs[2]); bad2[9 if flags else 4] ^= 0x07  # a damaged key near the front
        msgs = [blobs[0], bad1, blobs[1], bytes(bad2), blobs[2]]
        offs, lens, cur = [], [], 0
        for b in msgs:
            offs.append(cur)
            lens.append(len(b))
            cur += len(b)
        assert cur >= (8 << 20)
        st_o, want = ob.decode_bytes(msgs, abi.MSG_PREDICT_RESPONSE, flags)
        pwire = torch.empty(cur, dtype=torch.uint8).pin_memory()
        pwire.numpy()[:] = np.frombuffer(b"".join(msgs), dtype=np.uint8)
        ppool = torch.zeros(cur * 2 + 4096, dtype=torch.uint8).pin_memory()
        d = codec.decoder().decode_host_into(pwire.numpy(), offs, lens, abi.MSG_PREDICT_RESPONSE, ppool.numpy(), flags)
        assert d.status == st_o == abi.TFS_ERR_BATCH
        helpers.assert_same_decoded(want, d.messages(), "response", f"chunked pinned decode, flags {flags}")
        assert [m.status for m in d.messages()] == [m.status for m in want]
        d.close()


def test_config_c2_batched_unaligned(codec):
    x = workloads.c2_payload(batch=4)
    msgs = [workloads.c2_request(x, v, batch=4) for v in ("float_val", "tensor_content", "float_val")]
    roundtrip_check(codec, msgs, abi.MSG_PREDICT_REQUEST, "C2 batch, packed back to back")


@pytest.mark.parametrize("negatives", [0.0, 0.01])
def test_config_c3_full_size(codec, negatives):
    ids, mask = workloads.c3_arrays(negatives=negatives)
    roundtrip_check(codec, [workloads.c3_request(ids, mask)] * 4, abi.MSG_PREDICT_REQUEST, f"C3 neg={negatives}")


@pytest.mark.parametrize("fmt", ["f16", "bf16"])
def test_config_c4_full_size(codec, fmt):
    x = workloads.c4_payload()
    roundtrip_check(codec, [workloads.c4_tensor(x, fmt)], abi.MSG_TENSOR_PROTO, f"C4 {fmt}")


def test_config_c5_mixed_responses(codec):
    msgs = workloads.c5_responses(512, seed=5001, max_log2=14)
    roundtrip_check(codec, msgs, abi.MSG_PREDICT_RESPONSE, "C5")


@pytest.mark.parametrize("flags", [abi.F_ALIGN_PAYLOAD, abi.F_GRPC_PREFIX, abi.F_GRPC_PREFIX | abi.F_ALIGN_PAYLOAD])
@pytest.mark.parametrize("seed", [11, 12])
def test_mixed_batches_under_every_flag(codec, flags, seed):
    """Mixed-dtype responses and requests (element counts log-uniform from 1: most payloads are a few bytes and take the
    warp-per-segment copy, some are large) under the flags the C5 test leaves out; host, device and pinned paths."""
    msgs = workloads.c5_responses(160, seed=seed, max_log2=13)
    roundtrip_check(codec, msgs, abi.MSG_PREDICT_RESPONSE, f"mixed responses (flags {flags})", flags)
    reqs = [tfs.PredictRequest(m.model_spec if i % 3 else None, m.outputs, ["f%d" % i] * (i % 3)) for i, m in enumerate(msgs[:96])]
    roundtrip_check(codec, reqs, abi.MSG_PREDICT_REQUEST, f"mixed requests (flags {flags})", flags)


def test_varint_encode_around_length_prefix_boundaries(codec):
    """Element counts around the powers of 128 where a length prefix changes size, every varint source type, gRPC prefix on
    and off, two tensors and filters per request: bytes against the oracle."""
    import torch
    dev = torch.device(f"cuda:{codec.ctx.device}")
    rng = np.random.default_rng(128)
    for flags in (0, abi.F_GRPC_PREFIX):
        msgs = []
        for i in range(96):
            n = int(rng.choice([1, 2, 12, 13, 42, 43, 127, 128, 129, 1638, 1639, 2047, 2048, 2049, 5461, 5462, 16383, 16384, 16385,
                                40000, 209715, 209716, 300000])) + int(rng.integers(0, 3))
            which = i % 6
            if which == 0:
                t = tfs.TensorProto(dtype=9, int64_val=rng.integers(0, 1 << int(rng.integers(1, 40)), n, dtype=np.int64))
            elif which == 1:
                t = tfs.TensorProto(dtype=3, int_val=rng.integers(-5, 300, n).astype(np.int32))
            elif which == 2:
                t = tfs.TensorProto(dtype=19, half_val=tfs.Payload(rng.standard_normal(n).astype(np.float32), abi.T_F32, abi.T_F16))
            elif which == 3:
                t = tfs.TensorProto(dtype=22, uint32_val=rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32))
            elif which == 4:
                t = tfs.TensorProto(dtype=9, int64_val=np.zeros(n, np.int64), float_val=rng.standard_normal(7).astype(np.float32),
                                    uint64_val=rng.integers(0, 1 << 63, n // 3 + 1, dtype=np.uint64))
            else:
                t = tfs.TensorProto(dtype=14, half_val=rng.integers(0, 1 << 16, n).astype(np.uint16))
            msgs.append(tfs.PredictRequest(tfs.ModelSpec("m" * int(rng.integers(0, 200)), signature_name="s"), [("k%d" % i, t), ("z", tfs.TensorProto(dtype=1))],
                                           ["f"] * int(rng.integers(0, 3))))
        want = ob.encode_messages(msgs, abi.MSG_PREDICT_REQUEST, flags)
        for rep in range(2):
            eb = codec.encoder().encode(workloads.to_device(msgs, abi.MSG_PREDICT_REQUEST, dev), abi.MSG_PREDICT_REQUEST, flags)
            got = eb.all_bytes()
            for i, (g, w) in enumerate(zip(got, want)):
                assert g == w, f"flags {flags} rep {rep} message {i}"


def test_config_c5_at_benchmarked_scale(codec):
    """C5 exactly as bench.py times it: one GPU's share (1 024 of the 8 192 responses, seed 5001, element counts
    log-uniform up to 2^18, ~200 MB of wire), the wire produced by the oracle, decoded in one call into a device
    pool; every status and every value against the oracle's decode."""
    import torch
    msgs = workloads.c5_responses(1024, seed=5001, max_log2=18)
    built = tfs.build_batch(msgs, abi.MSG_PREDICT_RESPONSE)
    wire, offs, lens = ob.encode(built)
    st_o, want = ob.decode(wire, offs, lens, abi.MSG_PREDICT_RESPONSE)
    assert st_o == 0
    dev = torch.device(f"cuda:{codec.ctx.device}")
    dwire = torch.from_numpy(wire).to(dev)
    d = codec.decoder().scan(dwire, offs, lens, abi.MSG_PREDICT_RESPONSE)
    pool = torch.empty(d.pool_bytes + 64, dtype=torch.uint8, device=dev)
    d.close()
    d = codec.decoder().decode_into(dwire, offs, lens, abi.MSG_PREDICT_RESPONSE, pool)
    codec.ctx.sync()
    assert d.status == 0
    helpers.assert_same_decoded_fast(want, d.messages(), "response", "C5 at bench scale")
    d.close()
    # and the GPU's own encoding of the same responses is the oracle's wire, byte for byte
    eb = codec.encoder().encode(workloads.to_device(msgs, abi.MSG_PREDICT_RESPONSE, dev), abi.MSG_PREDICT_RESPONSE)
    assert eb.lengths == lens
    got = eb.wire[: eb.total].cpu().numpy()
    assert eb.offsets == offs and got.tobytes() == wire.tobytes()


def test_full_size_properties(codec):
    """Size-independent checks at full BASELINE scale, without the oracle: encode -> decode is the identity
    and the wire has the length prost's encoded_len rule gives."""
    import torch
    dev = torch.device(f"cuda:{codec.ctx.device}")
    x = torch.randn(4, 64 * 224 * 224 * 3, device=dev)
    msgs = [tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"),
                               [("input", tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([64, 224, 224, 3]), float_val=x[i]))])
            for i in range(4)]
    eb = codec.encoder().encode(msgs, abi.MSG_PREDICT_REQUEST, abi.F_ALIGN_PAYLOAD)
    assert eb.lengths == [38535239] * 4
    d = codec.decoder().decode(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST)
    view = d.view
    for i in range(4):
        t = view.tensors[view.msgs[i].first_tensor]
        assert t.n_elems[1] == x.shape[1] and t.dtype == 1 and [t.dims[j] for j in range(4)] == [64, 224, 224, 3]
        got = d.pool[t.dst_off[1]: t.dst_off[1] + 4 * x.shape[1]].view(torch.float32)
        assert torch.equal(got, x[i])
    ids = torch.randint(-5, 30522, (64, 256 * 512), device=dev, dtype=torch.int64)
    msgs = [tfs.TensorProto(dtype=9, int64_val=ids[i]) for i in range(64)]
    eb = codec.encoder().encode(msgs, abi.MSG_TENSOR_PROTO)
    d = codec.decoder().decode(eb.wire, eb.offsets, eb.lengths, abi.MSG_TENSOR_PROTO)
    for i in range(64):
        t = d.view.tensors[d.view.msgs[i].first_tensor]
        assert t.n_elems[6] == ids.shape[1]
        got = d.pool[t.dst_off[6]: t.dst_off[6] + 8 * ids.shape[1]].view(torch.int64)
        assert torch.equal(got, ids[i])
