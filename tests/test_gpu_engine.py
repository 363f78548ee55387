"""GPU tests of the staging pool, the engine (asynchronous jobs on worker threads, one or several GPUs from one process)
and the coalescer - the pieces between the codec and the transport (reference call site src/lib.rs:265; the client is
cloned across tokio tasks, src/lib.rs:148-156, examples/async.rs:29-46).  Everything goes through the C ABI; results are
checked against the oracle."""
import ctypes as C
import threading

import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import abi, workloads
from tfs_b200._lib import lib

import helpers
import oracle_bind as ob

pytestmark = pytest.mark.gpu


def _devices():
    import torch
    return list(range(min(torch.cuda.device_count(), 8)))


def test_staging_pool_slabs_skew_recycling_and_fallback():
    ctx = tfs.GpuContext(0, pinned_bytes=1 << 20)
    h = ctx.handle
    in_use, size, fb = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib().tfs_stage_stats(h, C.byref(in_use), C.byref(size), C.byref(fb))
    assert size.value == 1 << 20 and in_use.value == 0
    slabs = [tfs.Slab(h, n, skew) for n, skew in ((100, 0), (5000, 3), (4096, 15), (70_000, 9))]
    for s, skew in zip(slabs, (0, 3, 15, 9)):
        assert s.address % 16 == skew and s.raw.capacity >= 100
        s.array[:] = 0xA5  # the whole slab is writable
    lib().tfs_stage_stats(h, C.byref(in_use), C.byref(size), C.byref(fb))
    assert in_use.value == 4096 + 8192 + 8192 + 131072 and fb.value == 0
    first = slabs[0].address
    for s in slabs:
        s.release()
    again = tfs.Slab(h, 50, 0)
    assert again.address == first  # recycled, not carved anew
    again.release()
    big = tfs.Slab(h, 3 << 20, 5)  # larger than the pool: a one-off pinned allocation
    assert big.address % 16 == 5
    big.array[: 3 << 20] = 1
    lib().tfs_stage_stats(h, C.byref(in_use), C.byref(size), C.byref(fb))
    assert fb.value == 1
    big.release()
    # a slab is pinned memory: the host entry points use it in place
    codec = tfs.GpuCodec(ctx)
    msg = workloads.c1_request()
    dst = tfs.Slab(h, 4096, 0)
    eb = codec.encoder().encode_host([msg], abi.MSG_PREDICT_REQUEST, 0, dst=dst.array)
    assert eb.all_bytes() == ob.encode_messages([msg], abi.MSG_PREDICT_REQUEST)
    dst.release()
    ctx.close()


def test_decode_from_a_skewed_slab_takes_the_aligned_path():
    """A response received into a slab whose skew puts its bulk payload on a 16-byte boundary decodes to the same values."""
    ctx = tfs.GpuContext(0, pinned_bytes=64 << 20)
    codec = tfs.GpuCodec(ctx)
    x = np.random.default_rng(5).standard_normal(1_000_003).astype(np.float32)
    resp = tfs.PredictResponse([("probabilities", tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([x.size]), float_val=x))],
                               tfs.ModelSpec("m", version=3, signature_name="serving_default"))
    wire = ob.encode_messages([resp], abi.MSG_PREDICT_RESPONSE)[0]
    h = wire.index(x[:4].tobytes())  # where the payload starts inside the message
    slab = tfs.Slab(ctx.handle, len(wire), (16 - h % 16) % 16)
    assert (slab.address + h) % 16 == 0
    slab.array[: len(wire)] = np.frombuffer(wire, np.uint8)
    pool = tfs.Slab(ctx.handle, 4 * x.size + 4096, 0)
    n0 = ctx.launch_count()
    d = codec.decoder().decode_host_into(slab.array[: len(wire)], [0], [len(wire)], abi.MSG_PREDICT_RESPONSE, pool.array)
    st, want = ob.decode_bytes([wire], abi.MSG_PREDICT_RESPONSE)
    helpers.assert_same_decoded(want, d.messages(), "response", "skewed slab")
    assert ctx.launch_count() - n0 <= 3  # plan + TMA ring (+ nothing for the generic copy)
    d.close()
    slab.release()
    pool.release()
    ctx.close()


def test_engine_async_jobs_match_the_synchronous_entry_points():
    eng = tfs.Engine([0], lanes_per_device=2, pinned_bytes_per_lane=32 << 20)
    assert eng.lanes == 2
    rng = np.random.default_rng(11)
    kind = abi.MSG_PREDICT_REQUEST
    batches = []
    for j in range(6):
        msgs = [tfs.PredictRequest(tfs.ModelSpec("m%d" % j, signature_name="s"),
                                   [("x", tfs.TensorProto(dtype=1, float_val=rng.standard_normal(50_000 + 17 * i).astype(np.float32))),
                                    ("ids", tfs.TensorProto(dtype=9, int64_val=rng.integers(-3, 1 << 30, 30_000 + i, dtype=np.int64)))])
                for i in range(4)]
        batches.append(msgs)
    builts = [tfs.build_batch(m, kind) for m in batches]
    dsts = [np.empty(int(lib().tfs_encode_bound(C.byref(b.batch), 0)), np.uint8) for b in builts]
    tickets = [eng.encode_host(b, d) for b, d in zip(builts, dsts)]          # all six in flight
    encs = [eng.wait(t) for t in reversed(tickets)][::-1]                       # waited for in reverse order
    for msgs, eb in zip(batches, encs):
        assert eb.all_bytes() == ob.encode_messages(msgs, kind)
    pools = [np.empty(2 << 20, np.uint8) for _ in encs]
    dt = [eng.decode_host_into(eb.wire[: eb.total], eb.offsets, eb.lengths, kind, p) for eb, p in zip(encs, pools)]
    for msgs, t in zip(batches, dt):
        d = eng.wait(t)
        st, want = ob.decode_bytes(ob.encode_messages(msgs, kind), kind)
        helpers.assert_same_decoded(want, d.messages(), "request", "engine decode")
        d.close()
    # a job that fails reports ITS status through its ticket, and the lane keeps working
    bad = tfs.build_batch([tfs.TensorProto(dtype=1, float_val=np.ones(100_000, np.float32))], abi.MSG_TENSOR_PROTO)
    t = eng.encode_host(bad, np.empty(100, np.uint8))
    with pytest.raises(tfs.TfsError) as e:
        eng.wait(t)
    assert e.value.status == abi.TFS_ERR_NOMEM
    t = eng.encode_host(builts[0], dsts[0])
    assert eng.wait(t).all_bytes() == ob.encode_messages(batches[0], kind)
    eng.close()


def test_engine_sharded_decode_over_all_gpus():
    """BASELINE config 5 in one process: a batch of independent mixed-dtype responses cut into runs of about equal bytes,
    one run per lane (every visible GPU), checked against the oracle."""
    devs = _devices()
    eng = tfs.Engine(devs, lanes_per_device=1, pinned_bytes_per_lane=0)
    msgs = workloads.c5_responses(96, seed=5001, max_log2=14)
    wire, offs, lens = ob.encode(tfs.build_batch(msgs, abi.MSG_PREDICT_RESPONSE))
    st_o, want = ob.decode(wire, offs, lens, abi.MSG_PREDICT_RESPONSE)
    pools = [np.empty(int(wire.size) * 3 + 65536, np.uint8) for _ in range(eng.lanes)]
    st, res, first = eng.decode_host_sharded(wire, offs, lens, abi.MSG_PREDICT_RESPONSE, pools)
    assert st == 0 and first[0] == 0 and first == sorted(first)
    got = []
    for r in res:
        if r is not None:
            got.extend(r.messages())
            r.close()
    helpers.assert_same_decoded(want, got, "response", f"sharded over {eng.lanes} lanes")
    # the runs are balanced by bytes
    if eng.lanes > 1:
        ends = first[1:] + [len(msgs)]
        per = [sum(lens[a:b]) for a, b in zip(first, ends)]
        assert max(per) < 2.5 * (sum(per) / len(per)) + max(lens)
    eng.close()


def test_coalescer_gathers_single_message_calls_into_batches():
    eng = tfs.Engine([0], lanes_per_device=2, pinned_bytes_per_lane=16 << 20)
    co = tfs.Coalescer(eng, window_us=20_000, max_msgs=8)
    rng = np.random.default_rng(23)
    msgs = [tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"),
                               [("input", tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([1, 8, 8, 3]),
                                                          float_val=rng.standard_normal(192 + i).astype(np.float32)))], ["out%d" % i])
            for i in range(24)]
    want = ob.encode_messages(msgs, abi.MSG_PREDICT_REQUEST, abi.F_GRPC_PREFIX)
    got = [None] * len(msgs)

    def call(i):
        got[i] = co.encode(msgs[i], abi.MSG_PREDICT_REQUEST, abi.F_GRPC_PREFIX)

    threads = [threading.Thread(target=call, args=(i,)) for i in range(len(msgs))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert got == want
    batches, messages = co.stats()
    assert messages == len(msgs) and batches < len(msgs)  # they did share launches
    # everything was released: the pools are empty again
    for lane in range(eng.lanes):
        in_use = C.c_uint64()
        lib().tfs_stage_stats(eng.lane_ctx(lane), C.byref(in_use), None, None)
        assert in_use.value == 0
    # a lone caller is not held longer than the window
    assert co.encode(msgs[0], abi.MSG_PREDICT_REQUEST, abi.F_GRPC_PREFIX) == want[0]
    co.close()
    eng.close()
