"""Out-of-bounds WRITE check of our own (compute-sanitizer is not available on the GPU pool): every output buffer sits
inside a larger allocation filled with a canary byte; after encode / decode the canaries on both sides must be intact
and the bytes between them must equal the oracle's."""
import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import abi, workloads

import helpers
import oracle_bind as ob

pytestmark = pytest.mark.gpu
CANARY = 0xA5
GUARD = 4096


@pytest.fixture(scope="module")
def codec():
    c = tfs.GpuCodec(tfs.GpuContext(0))
    yield c
    c.ctx.close()


def _cases():
    rng = np.random.default_rng(2024)
    strs = [rng.integers(0, 256, int(n), dtype=np.uint8).tobytes() for n in [0, 1, 15, 16, 17, 4095, 70001, 3, 0, 65536 - 3, 9]]
    tensors = [
        tfs.TensorProto(dtype=9, int64_val=rng.integers(-3, 1 << 45, 6145, dtype=np.int64)),
        tfs.TensorProto(dtype=3, int_val=rng.integers(-100, 100, 2049, dtype=np.int32), float_val=rng.standard_normal(8191).astype(np.float32)),
        tfs.TensorProto(dtype=7, string_val=strs, tensor_content=rng.integers(0, 256, 33333, dtype=np.uint8)),
        tfs.TensorProto(dtype=19, half_val=rng.integers(0, 65536, 4097).astype(np.int32), bool_val=rng.integers(0, 2, 1001).astype(np.bool_)),
        tfs.TensorProto(dtype=2, double_val=rng.standard_normal(5000)),
        # many short elements: the decode walk's run loop, one table entry and one offset word per element
        tfs.TensorProto(dtype=7, string_val=[rng.integers(0, 256, int(n), dtype=np.uint8).tobytes() for n in rng.integers(0, 41, 700)]),
    ]
    reqs = [tfs.PredictRequest(tfs.ModelSpec("m", version=3, signature_name="s"), [("a", tensors[0]), ("b", tensors[2])], ["y"]),
            tfs.PredictRequest(None, [("", tensors[1]), ("k", tensors[3]), ("d", tensors[4])])]
    inputs = [tfs.Input([[("ids", tfs.Feature(int64_list=rng.integers(0, 9999, 3000, dtype=np.int64))),
                          ("img", tfs.Feature(bytes_list=strs)), ("w", tfs.Feature(float_list=rng.standard_normal(777).astype(np.float32)))], []])]
    return [("tensor", abi.MSG_TENSOR_PROTO, tensors), ("request", abi.MSG_PREDICT_REQUEST, reqs), ("input", abi.MSG_INPUT, inputs)]


@pytest.mark.parametrize("flags", [0, abi.F_GRPC_PREFIX, abi.F_ALIGN_PAYLOAD])
@pytest.mark.parametrize("skew", [0, 1, 7])
def test_encode_writes_only_its_own_bytes(codec, flags, skew):
    import torch
    enc = codec.encoder()
    for name, kind, msgs in _cases():
        want = ob.encode_messages(msgs, kind, flags & abi.F_GRPC_PREFIX)
        dev_msgs = workloads.to_device(msgs, kind, torch.device("cuda:0"))
        built = tfs.build_batch(dev_msgs, kind)
        # learn the exact total at the same address skew (TFS_F_ALIGN_PAYLOAD pads according to dst mod 16)
        probe_buf = torch.empty(GUARD + skew + enc.bound(built, flags), dtype=torch.uint8, device="cuda:0")
        total = enc.encode(built, kind, flags, dst=probe_buf[GUARD + skew:]).total
        buf = torch.full((GUARD + skew + total + GUARD,), CANARY, dtype=torch.uint8, device="cuda:0")
        dst = buf[GUARD + skew: GUARD + skew + total]     # capacity == exactly what is needed
        eb = enc.encode(built, kind, flags, dst=dst)
        codec.ctx.sync()
        host = buf.cpu().numpy()
        assert (host[: GUARD + skew] == CANARY).all(), f"{name}: bytes written BEFORE dst (flags {flags}, skew {skew})"
        assert (host[GUARD + skew + total:] == CANARY).all(), f"{name}: bytes written AFTER dst (flags {flags}, skew {skew})"
        got = [bytes(host[GUARD + skew + o: GUARD + skew + o + n + (5 if flags & abi.F_GRPC_PREFIX else 0)]) for o, n in zip(eb.offsets, eb.lengths)]
        assert [len(g) for g in got] == [len(w) for w in want], name
        assert got == want, f"{name}: encoded bytes differ (flags {flags}, skew {skew})"


def test_decode_rejects_an_unaligned_device_pool(codec):
    import torch
    wire = torch.from_numpy(np.frombuffer(bytes.fromhex("3a03010203"), dtype=np.uint8).copy()).cuda()
    buf = torch.empty(4096, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(tfs.TfsError) as e:
        codec.decoder().decode_into(wire, [0], [5], abi.MSG_TENSOR_PROTO, buf[3:])
    assert e.value.status == abi.TFS_ERR_INVALID_ARG


@pytest.mark.parametrize("skew", [0, 16, 48])
def test_decode_writes_only_its_own_pool(codec, skew):
    import torch
    dec = codec.decoder()
    for name, kind, msgs in _cases():
        if kind == abi.MSG_INPUT:
            continue  # encode only
        blobs = ob.encode_messages(msgs, kind)
        offs, lens, cur = [], [], 0
        for b in blobs:
            offs.append(cur)
            lens.append(len(b))
            cur += len(b)
        wire = torch.from_numpy(np.frombuffer(b"".join(blobs), dtype=np.uint8).copy()).cuda()
        d0 = dec.scan(wire, offs, lens, kind)
        need = d0.pool_bytes
        d0.close()
        buf = torch.full((GUARD + skew + need + GUARD,), CANARY, dtype=torch.uint8, device="cuda:0")
        pool = buf[GUARD + skew: GUARD + skew + need]
        d = dec.decode_into(wire, offs, lens, kind, pool)
        codec.ctx.sync()
        host = buf.cpu().numpy()
        assert (host[: GUARD + skew] == CANARY).all(), f"{name}: bytes written BEFORE the pool (skew {skew})"
        assert (host[GUARD + skew + need:] == CANARY).all(), f"{name}: bytes written AFTER the pool (skew {skew})"
        st, want = ob.decode_bytes(blobs, kind)
        helpers.assert_same_decoded(want, d.messages(), "tensor" if kind == abi.MSG_TENSOR_PROTO else "request", f"guarded decode {name}")
        d.close()


def test_encode_rejects_a_misaligned_varint_source(codec):
    """int64 / int32 arrays are loaded element-wise by the varint kernels: an address that is not a multiple of the
    element size is refused (INVALID_ARG), not dereferenced."""
    import torch
    x = torch.arange(64, dtype=torch.int64, device="cuda:0")
    built = tfs.build_batch([tfs.TensorProto(dtype=9, int64_val=x)], abi.MSG_TENSOR_PROTO)
    built.fields[0].data += 4
    built.fields[0].n_elems = 60
    with pytest.raises(tfs.TfsError) as e:
        codec.encoder().encode(built, abi.MSG_TENSOR_PROTO)
    assert e.value.status == abi.TFS_ERR_INVALID_ARG
    # byte-copied fields may sit anywhere
    f = torch.arange(64, dtype=torch.float32, device="cuda:0")
    built = tfs.build_batch([tfs.TensorProto(dtype=1, float_val=f)], abi.MSG_TENSOR_PROTO)
    built.fields[0].data += 1
    built.fields[0].n_elems = 60
    got = codec.encoder().encode(built, abi.MSG_TENSOR_PROTO).all_bytes()[0]
    raw = f.cpu().numpy().tobytes()[1: 1 + 240]
    assert got == bytes.fromhex("08012af001") + raw


def test_encode_into_a_too_small_buffer_writes_nothing(codec):
    """dst_capacity one byte short: TFS_ERR_NOMEM, and not a single byte of (or around) the buffer is touched - the
    framing, varint, string and copy kernels are all gated on the layout kernel's overflow flag."""
    import torch
    enc = codec.encoder()
    for name, kind, msgs in _cases():
        dev_msgs = workloads.to_device(msgs, kind, torch.device("cuda:0"))
        built = tfs.build_batch(dev_msgs, kind)
        total = enc.encode(built, kind, 0).total
        buf = torch.full((GUARD + total + GUARD,), CANARY, dtype=torch.uint8, device="cuda:0")
        for flags in (0, abi.F_ASYNC):  # the asynchronous form returns before the (gated) payload kernels have run
            with pytest.raises(tfs.TfsError) as e:
                enc.encode(built, kind, flags, dst=buf[GUARD: GUARD + total - 1])
            assert e.value.status == abi.TFS_ERR_NOMEM, name
            codec.ctx.sync()
            assert bool((buf == CANARY).all()), f"{name}: bytes were written although the buffer was too small (flags {flags})"
        # the context is fine afterwards
        assert enc.encode(built, kind, 0).all_bytes() == ob.encode_messages(msgs, kind)


def test_decode_into_a_too_small_pool_writes_nothing(codec):
    import torch
    dec = codec.decoder()
    for name, kind, msgs in _cases():
        if kind == abi.MSG_INPUT:
            continue
        blobs = ob.encode_messages(msgs, kind)
        offs, lens, cur = [], [], 0
        for b in blobs:
            offs.append(cur)
            lens.append(len(b))
            cur += len(b)
        wire = torch.from_numpy(np.frombuffer(b"".join(blobs), dtype=np.uint8).copy()).cuda()
        d0 = dec.scan(wire, offs, lens, kind)
        need = d0.pool_bytes
        d0.close()
        buf = torch.full((GUARD + need + GUARD,), CANARY, dtype=torch.uint8, device="cuda:0")
        with pytest.raises(tfs.TfsError) as e:
            dec.decode_into(wire, offs, lens, kind, buf[GUARD: GUARD + need - 64])
        assert e.value.status == abi.TFS_ERR_NOMEM, name
        codec.ctx.sync()
        assert bool((buf == CANARY).all()), f"{name}: values were written although the pool was too small"


def test_tma_ring_queued_ahead_of_the_counts(codec):
    """After a decode of bulk payloads the next plan queues the TMA copy kernel BEHIND the plan kernel, before the host has
    seen the counts (tfs_decode.cu: launch_copy_tma_deferred); the kernel reads them from device memory and must stay idle
    unless the plan ended well.  Sequences on ONE context: bulk -> bulk (the ring runs ahead), -> a pool that is too small
    (nothing may be written), -> varint bodies (the plan stops early), -> tiny messages (no bulk at all), -> malformed
    messages between good ones, -> bulk again; values against the oracle every time, canaries around every pool."""
    import torch
    rng = np.random.default_rng(4711)
    fresh = tfs.GpuCodec(tfs.GpuContext(0))
    dec = fresh.decoder()

    def bulk(n_msgs, n):
        return [tfs.PredictResponse([("y", tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([n + i]),
                                                            float_val=rng.standard_normal(n + i).astype(np.float32)))], tfs.ModelSpec("m", version=i))
                for i in range(n_msgs)]

    varint = [tfs.PredictResponse([("i", tfs.TensorProto(dtype=9, int64_val=rng.integers(-9, 1 << 40, 70_000, dtype=np.int64))),
                                   ("y", tfs.TensorProto(dtype=1, float_val=rng.standard_normal(50_000).astype(np.float32)))]) for _ in range(5)]
    tiny = [tfs.PredictResponse([("t", tfs.TensorProto(dtype=1, float_val=rng.standard_normal(3).astype(np.float32)))]) for _ in range(40)]
    kind = abi.MSG_PREDICT_RESPONSE

    def run(msgs, shrink=0, corrupt=()):
        blobs = ob.encode_messages(msgs, kind)
        blobs = [b[: len(b) // 2] if i in corrupt else b for i, b in enumerate(blobs)]
        offs, lens, cur = [], [], 0
        for b in blobs:
            offs.append(cur)
            lens.append(len(b))
            cur += len(b)
        wire = torch.from_numpy(np.frombuffer(b"".join(blobs), dtype=np.uint8).copy()).cuda()
        d0 = dec.scan(wire, offs, lens, kind)
        need = d0.pool_bytes
        d0.close()
        buf = torch.full((GUARD + need + GUARD,), CANARY, dtype=torch.uint8, device="cuda:0")
        if shrink:
            with pytest.raises(tfs.TfsError) as e:
                dec.decode_into(wire, offs, lens, kind, buf[GUARD: GUARD + need - shrink])
            assert e.value.status == abi.TFS_ERR_NOMEM
            fresh.ctx.sync()
            assert bool((buf == CANARY).all()), "values were written although the pool was too small"
            return
        d = dec.decode_into(wire, offs, lens, kind, buf[GUARD: GUARD + need])
        fresh.ctx.sync()
        host = buf.cpu().numpy()
        assert (host[:GUARD] == CANARY).all() and (host[GUARD + need:] == CANARY).all()
        st, want = ob.decode_bytes(blobs, kind)
        helpers.assert_same_decoded_fast(want, d.messages(), "response", "ring ahead")
        d.close()

    try:
        run(bulk(6, 300_000))
        run(bulk(6, 300_000))                 # the ring is queued ahead from here on
        run(bulk(5, 250_000), shrink=4096)    # ... and must not write into a pool that is too small
        run(bulk(4, 200_000))
        run(varint)                           # the plan kernel stops before the layout: no go
        run(bulk(70, 40_000))                 # more than 64 segments: the warp-cooperative variant
        run(bulk(3, 100_000), corrupt=(1,))   # a truncated message between good ones
        run(tiny)                             # nothing for the ring to do
        run(bulk(2, 2_000_000))
    finally:
        fresh.ctx.close()
