"""Randomised differential test on the GPU: many messages with random mixes of every payload kind (strings, parts that are
concatenated, empty fields, casts), random flags; encode bytes and decoded values against the oracle.  Sized to run in a
few seconds."""
import os

import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import abi, workloads

import helpers
import oracle_bind as ob

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def codec():
    c = tfs.GpuCodec(tfs.GpuContext(0))
    yield c
    c.ctx.close()


def _size(rng):
    r = rng.random()
    if r < 0.15:
        return 0
    if r < 0.7:
        return int(rng.integers(1, 40))
    if r < 0.95:
        return int(rng.integers(40, 5000))
    return int(rng.integers(5000, 120_000))


def _ints(rng, n, dtype):
    info = np.iinfo(dtype)
    mode = rng.integers(0, 4)
    if mode == 0:
        return rng.integers(0, 128, n).astype(dtype)
    if mode == 1:
        return rng.integers(max(info.min, -(1 << 20)), min(info.max, 1 << 20), n).astype(dtype)
    if mode == 2:
        return rng.integers(info.min, info.max, n, dtype=dtype, endpoint=True)
    a = rng.integers(0, 1 << 14, n).astype(dtype)
    if n and info.min < 0:
        a[rng.integers(0, n, max(1, n // 50))] = -1
    return a


def random_tensor(rng):
    t = tfs.TensorProto(dtype=int(rng.integers(0, 24)), version_number=int(rng.integers(0, 3)))
    if rng.random() < 0.7:
        t.tensor_shape = tfs.TensorShapeProto([int(x) for x in rng.integers(-1, 1 << 33, int(rng.integers(0, 5)))], bool(rng.random() < 0.1))
    kinds = rng.permutation(12)[: int(rng.integers(0, 5))]
    for k in kinds:
        n = _size(rng)
        if k == 0:
            t.tensor_content = rng.integers(0, 256, n, dtype=np.uint8)
        elif k == 1:
            t.float_val = rng.standard_normal(n).astype(np.float32)
            if rng.random() < 0.3:  # two parts, one field
                t.float_val = (t.float_val, rng.standard_normal(_size(rng)).astype(np.float32))
        elif k == 2:
            t.double_val = rng.standard_normal(n)
        elif k == 3:
            t.int_val = _ints(rng, n, np.int32)
        elif k == 4:
            t.string_val = [rng.integers(0, 256, _size(rng) % 3000, dtype=np.uint8).tobytes() for _ in range(n % 300)]
        elif k == 5:
            t.scomplex_val = rng.standard_normal(2 * (n // 2)).astype(np.float32)
        elif k == 6:
            t.int64_val = _ints(rng, n, np.int64)
            if rng.random() < 0.3:
                t.int64_val = (t.int64_val, _ints(rng, _size(rng), np.int64))
        elif k == 7:
            t.bool_val = rng.integers(0, 2, n).astype(np.bool_)
        elif k == 8:
            t.dcomplex_val = rng.standard_normal(2 * (n // 2))
        elif k == 9:
            t.half_val = rng.integers(0, 1 << 16, n).astype(np.int32)
        elif k == 10:
            t.uint32_val = _ints(rng, n, np.uint32)
        else:
            t.uint64_val = _ints(rng, n, np.uint64)
    return t


def flatten_parts(t):
    """The same tensor with concatenated parts joined (what the wire and the decoder see)."""
    import copy
    u = copy.copy(t)
    for k in abi.SLOT_KINDS:
        name = abi.KIND_FIELD_NAMES[k]
        v = getattr(u, name)
        if isinstance(v, tuple):
            setattr(u, name, np.concatenate(v))
    return u


SEEDS = list(range(1, 1 + int(os.environ.get("TFS_STRESS_SEEDS", "3"))))  # more seeds for a longer hunt


@pytest.mark.parametrize("seed", SEEDS)
def test_random_requests_encode_and_decode_like_the_oracle(codec, seed):
    rng = np.random.default_rng(1000 + seed)
    n_msgs = [60, 300, 17][(seed - 1) % 3]  # below and above the 256-message single-launch paths
    msgs = []
    for i in range(n_msgs):
        entries = [("" if rng.random() < 0.05 else "t%d" % j, random_tensor(rng)) for j in range(int(rng.integers(0, 5)))]
        ms = None if rng.random() < 0.3 else tfs.ModelSpec("m%d" % i, int(rng.integers(0, 1 << 40)) if rng.random() < 0.5 else None,
                                                           None, "sig" if rng.random() < 0.5 else "")
        msgs.append(tfs.PredictRequest(ms, entries, ["f%d" % k for k in range(int(rng.integers(0, 3)))]))
    flat = [tfs.PredictRequest(m.model_spec, [(k, flatten_parts(t)) for k, t in m.inputs], m.output_filter) for m in msgs]
    for flags in (0, abi.F_GRPC_PREFIX, abi.F_ALIGN_PAYLOAD):
        want = ob.encode_messages(flat, abi.MSG_PREDICT_REQUEST, flags & abi.F_GRPC_PREFIX)
        import torch
        dev = workloads.to_device(msgs, abi.MSG_PREDICT_REQUEST, torch.device("cuda:0"))
        got = codec.encoder().encode(dev, abi.MSG_PREDICT_REQUEST, flags).all_bytes()
        assert [len(g) for g in got] == [len(w) for w in want]
        bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
        assert not bad, f"seed {seed} flags {flags}: messages {bad[:5]} differ"
        host = codec.encoder().encode_host(msgs, abi.MSG_PREDICT_REQUEST, flags).all_bytes()
        assert host == want
    # decode what was encoded (device and host paths) and compare values with the oracle's decode
    blobs = ob.encode_messages(flat, abi.MSG_PREDICT_REQUEST)
    offs, lens, cur = [], [], 0
    for b in blobs:
        offs.append(cur)
        lens.append(len(b))
        cur += len(b)
    st_o, want_dec = ob.decode_bytes(blobs, abi.MSG_PREDICT_REQUEST)
    wire_np = np.frombuffer(b"".join(blobs), dtype=np.uint8)
    d = codec.decoder().decode_host(wire_np, offs, lens, abi.MSG_PREDICT_REQUEST)
    assert d.status == st_o == 0
    helpers.assert_same_decoded(want_dec, d.messages(), "request", f"stress seed {seed} host")
    import torch
    dw = torch.from_numpy(wire_np.copy()).cuda()
    d2 = codec.decoder().decode(dw, offs, lens, abi.MSG_PREDICT_REQUEST)
    codec.ctx.sync()
    helpers.assert_same_decoded(want_dec, d2.messages(), "request", f"stress seed {seed} device")


def _mutate(rng, b):
    b = bytearray(b)
    op = rng.integers(0, 5)
    if op == 0 and len(b) > 1:
        b = b[: rng.integers(0, len(b))]
    elif op == 1 and len(b):
        for _ in range(int(rng.integers(1, 4))):
            b[rng.integers(0, len(b))] = rng.integers(0, 256)
    elif op == 2 and len(b):
        j = rng.integers(0, len(b))
        b[j:j] = bytes(rng.integers(0, 256, rng.integers(1, 5), dtype=np.uint8))
    elif op == 3 and len(b) > 2:
        j = rng.integers(0, min(len(b), 60))
        b[j] ^= 1 << int(rng.integers(0, 8))
    elif len(b) > 4:  # a run of continuation bytes somewhere (over-long varints, lengths that swallow what follows)
        j = rng.integers(0, len(b) - 1)
        n = int(rng.integers(1, 13))
        b[j: j + n] = b"\xff" * min(n, len(b) - j)
    return bytes(b)


@pytest.mark.parametrize("kind", [abi.MSG_TENSOR_PROTO, abi.MSG_PREDICT_REQUEST])
def test_mutated_messages_fail_like_the_oracle(codec, kind):
    """Byte-level mutations of valid messages with every payload kind (strings included): the same status per message as
    the oracle, the same values where the message still parses."""
    rng = np.random.default_rng(int(os.environ.get("TFS_FUZZ_SEED", "7")) + kind)
    small = []
    for i in range(30):
        t = random_tensor(rng)
        for k in abi.SLOT_KINDS:  # keep the fuzz corpus small: mutations should hit structure, not megabytes of payload
            name = abi.KIND_FIELD_NAMES[k]
            v = getattr(t, name)
            if isinstance(v, tuple):
                v = np.concatenate(v)
            if v is not None:
                setattr(t, name, v[:50])
        small.append(t)
    if kind == abi.MSG_TENSOR_PROTO:
        base = ob.encode_messages(small, kind)
    else:
        reqs = [tfs.PredictRequest(tfs.ModelSpec("m", signature_name="s") if i % 2 else None,
                                   [("k%d" % j, small[(i + j) % len(small)]) for j in range(i % 4)], ["f"] * (i % 3)) for i in range(20)]
        base = ob.encode_messages(reqs, kind)
    n_cases = int(os.environ.get("TFS_FUZZ_CASES", "600"))
    blobs = [_mutate(rng, base[i % len(base)]) for i in range(n_cases)]
    st_o, want = ob.decode_bytes(blobs, kind)
    offs, lens, cur = [], [], 0
    for b in blobs:
        offs.append(cur)
        lens.append(len(b))
        cur += len(b)
    wire_np = np.frombuffer(b"".join(blobs), dtype=np.uint8)
    d = codec.decoder().decode_host(wire_np, offs, lens, kind)
    got = d.messages()
    mism = [(i, blobs[i].hex()[:400], abi.STATUS_NAMES.get(w.status), abi.STATUS_NAMES.get(g.status))
            for i, (w, g) in enumerate(zip(want, got)) if w.status != g.status]
    assert not mism, mism[:4]
    name = "tensor" if kind == abi.MSG_TENSOR_PROTO else "request"
    for i, (w, g) in enumerate(zip(want, got)):
        if w.status == 0:
            helpers.assert_same_decoded([w], [g], name, f"mutation {i}: {blobs[i].hex()[:200]}")


def test_contexts_on_separate_threads_do_not_disturb_each_other():
    """The reference serves requests from a multi-threaded runtime; the contract here is one ctx per thread.  Four threads,
    each with its own ctx (its own streams, tables and staging buffers), encode and decode different batches at the same
    time (ctypes releases the GIL for the duration of a call): every result equals the oracle's."""
    import threading

    import torch
    errors = []

    def worker(tid):
        try:
            c = tfs.GpuCodec(tfs.GpuContext(0))
            try:
                rng = np.random.default_rng(500 + tid)
                for it in range(6):
                    msgs = [tfs.PredictRequest(tfs.ModelSpec("m%d" % tid, signature_name="s"),
                                               [("t%d" % j, random_tensor(rng)) for j in range(int(rng.integers(1, 4)))])
                            for _ in range(int(rng.integers(1, 40)))]
                    flat = [tfs.PredictRequest(m.model_spec, [(k, flatten_parts(t)) for k, t in m.inputs], m.output_filter) for m in msgs]
                    want = ob.encode_messages(flat, abi.MSG_PREDICT_REQUEST)
                    got = c.encoder().encode_host(msgs, abi.MSG_PREDICT_REQUEST).all_bytes()
                    if got != want:
                        raise AssertionError(f"thread {tid} iteration {it}: encode differs")
                    dev = workloads.to_device(msgs, abi.MSG_PREDICT_REQUEST, torch.device("cuda:0"))
                    if c.encoder().encode(dev, abi.MSG_PREDICT_REQUEST).all_bytes() != want:
                        raise AssertionError(f"thread {tid} iteration {it}: device encode differs")
                    offs, lens, cur = [], [], 0
                    for b in want:
                        offs.append(cur)
                        lens.append(len(b))
                        cur += len(b)
                    st_o, want_dec = ob.decode_bytes(want, abi.MSG_PREDICT_REQUEST)
                    d = c.decoder().decode_host(np.frombuffer(b"".join(want), dtype=np.uint8), offs, lens, abi.MSG_PREDICT_REQUEST)
                    helpers.assert_same_decoded(want_dec, d.messages(), "request", f"thread {tid} iteration {it}")
            finally:
                c.ctx.close()
        except BaseException as e:  # noqa: BLE001 - reported in the main thread
            errors.append((tid, repr(e)))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
