"""GPU test of tfs_pack_int64_noalloc, the one entry point that never allocates device memory.

It serves callers that cannot call cudaMalloc — a codec inside a serving poll loop (an allocation would
synchronise the device and stall other requests), or a client whose GPU memory has already been claimed by a
co-located TensorFlow Serving process — by keeping its tables in a fixed block inside the library image.
Here it is called through the C ABI (ctypes) and its output is compared byte for byte with what the regular
entry point tfs_encode produces for the same values AND with the CPU oracle's packed `int64_val` body (the
restatement of prost's int64::encode_packed), on consecutive calls that reuse the shared block.

The file name sorts last on purpose: with `pytest -x` every other test has already run if this one fails.
"""
import ctypes as C

import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import abi
from tfs_b200._lib import lib

pytestmark = pytest.mark.gpu


def varint(v):
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def pack_noalloc(values, capacity=None):
    """values: torch int64 CUDA tensor -> (status, body bytes or None, reported length)"""
    import torch
    n = values.numel()
    cap = n * 10 + 16 if capacity is None else capacity
    dst = torch.full((cap + 32,), 0xCD, dtype=torch.uint8, device=values.device)
    length = torch.zeros(1, dtype=torch.int64, device=values.device)
    stream = torch.cuda.current_stream(values.device).cuda_stream
    st = lib().tfs_pack_int64_noalloc(C.c_void_p(values.data_ptr() if n else 0), n, C.c_void_p(dst.data_ptr() + 1), cap,
                                      C.c_void_p(length.data_ptr()), C.c_void_p(stream))
    if st != abi.TFS_OK:
        return st, None, None
    torch.cuda.synchronize()
    blen = int(length.item())
    host = dst.cpu().numpy()
    assert host[0] == 0xCD, "byte before dst was overwritten"
    if blen <= cap:
        assert (host[1 + blen:] == 0xCD).all(), "bytes past the body were overwritten"
        return st, bytes(host[1: 1 + blen]), blen
    assert (host == 0xCD).all(), "an over-long body must not be written at all"
    return st, None, blen


def strip_key_and_length(wire, n):
    """A bare TensorProto holding only int64_val is `52 varint(len) body`: returns the body."""
    if n == 0:
        assert wire == b""
        return b""
    assert wire[0] == 0x52
    # length varint
    i, blen, shift = 1, 0, 0
    while True:
        b = wire[i]
        blen |= (b & 0x7F) << shift
        shift += 7
        i += 1
        if b < 0x80:
            break
    assert len(wire) == i + blen
    return wire[i:]


def body_from_regular_entry_point(codec, values):
    """The same values through tfs_encode (a bare TensorProto holding only int64_val)."""
    eb = codec.encoder().encode([tfs.TensorProto(int64_val=values)], abi.MSG_TENSOR_PROTO)
    return strip_key_and_length(eb.all_bytes()[0], values.numel())


def body_from_oracle(values):
    """The same values through the CPU oracle (oracle/cpu_ref.cpp, prost's int64::encode_packed restated)."""
    import oracle_bind as ob
    wire = ob.encode_messages([tfs.TensorProto(int64_val=values.cpu().numpy())], abi.MSG_TENSOR_PROTO)[0]
    return strip_key_and_length(wire, values.numel())


def test_no_alloc_matches_regular_entry_point_on_consecutive_calls():
    import torch
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    codec = tfs.GpuCodec(tfs.GpuContext(0))
    g = torch.Generator(device=dev)
    g.manual_seed(20260)
    # consecutive calls share the library's fixed block: a large input first, then smaller ones whose
    # tables must not pick up anything the large call left behind, then tile-edge sizes and an empty input
    sizes = [300_001, 5, 2048, 2049, 70_000, 0, 1, 4097]
    for call, n in enumerate(sizes):
        shift = torch.randint(0, 63, (n,), device=dev, generator=g)
        vals = torch.randint(-(1 << 62), 1 << 62, (n,), device=dev, dtype=torch.int64, generator=g) >> shift
        if n > 10:
            vals[::7] = -1          # ten-byte varints
            vals[3::11] = 0
        st, body, blen = pack_noalloc(vals)
        assert st == abi.TFS_OK, (call, n, st)
        want = body_from_regular_entry_point(codec, vals)
        assert blen == len(want), (call, n, blen, len(want))
        assert body == want, f"call {call} (n={n}): packed bytes differ from tfs_encode"
        assert body == body_from_oracle(vals), f"call {call} (n={n}): packed bytes differ from the oracle"
    # twice in a row on the very same input: identical output both times
    vals = torch.arange(-5000, 5000, device=dev, dtype=torch.int64) * 123457
    first = pack_noalloc(vals)
    second = pack_noalloc(vals)
    assert first == second and first[1] == body_from_regular_entry_point(codec, vals) == body_from_oracle(vals)
    codec.ctx.close()


def test_no_alloc_limits_are_reported():
    import torch
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    # beyond the fixed block: refused with a status of its own, before anything is launched
    length = torch.zeros(1, dtype=torch.int64, device=dev)
    dummy = torch.zeros(16, dtype=torch.int64, device=dev)
    st = lib().tfs_pack_int64_noalloc(C.c_void_p(dummy.data_ptr()), abi.TFS_NOALLOC_MAX_ELEMS + 1, C.c_void_p(dummy.data_ptr()), 128,
                                      C.c_void_p(length.data_ptr()), None)
    assert st == abi.TFS_ERR_TOO_LARGE
    assert b"TFS_NOALLOC_MAX_ELEMS" in lib().tfs_strerror(st)
    assert lib().tfs_pack_int64_noalloc(None, 4, C.c_void_p(dummy.data_ptr()), 128, C.c_void_p(length.data_ptr()), None) == abi.TFS_ERR_INVALID_ARG
    # a destination that is too small: nothing written, the needed length is still reported
    vals = torch.full((1000,), -1, dtype=torch.int64, device=dev)
    st, body, blen = pack_noalloc(vals, capacity=9_999)
    assert st == abi.TFS_OK and body is None and blen == 10_000
    st, body, blen = pack_noalloc(vals, capacity=10_000)
    assert st == abi.TFS_OK and blen == 10_000 and body == b"\xff\xff\xff\xff\xff\xff\xff\xff\xff\x01" * 1000
    # the largest accepted size works (128 MiB of input) and agrees with the closed form for small values
    big = torch.full((abi.TFS_NOALLOC_MAX_ELEMS,), 300, dtype=torch.int64, device=dev)
    dst = torch.empty(2 * abi.TFS_NOALLOC_MAX_ELEMS, dtype=torch.uint8, device=dev)
    st = lib().tfs_pack_int64_noalloc(C.c_void_p(big.data_ptr()), big.numel(), C.c_void_p(dst.data_ptr()), dst.numel(),
                                      C.c_void_p(length.data_ptr()), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert st == abi.TFS_OK
    torch.cuda.synchronize()
    assert int(length.item()) == 2 * big.numel()
    assert torch.equal(dst.view(torch.int16), torch.full((big.numel(),), int(np.frombuffer(varint(300), "<i2")[0]), dtype=torch.int16, device=dev))
