"""GpuCodec — host-side mirror of the seam the reference uses for this path.

In the reference every RPC stub builds `tonic::codec::ProstCodec::default()` and hands it to
`Grpc::unary(request, path, codec)` (generated code behind `self.prediction_client.predict(request)`,
reference src/lib.rs:265).  tonic's Codec contract is `encoder() -> Encoder`, `decoder() -> Decoder`,
with `Encoder::encode(item, &mut BytesMut)` appending exactly the protobuf bytes and
`Decoder::decode(&mut BytesMut) -> Option<Item>` consuming one complete message.

`GpuCodec` keeps those names; because a GPU wants batches, `encode` / `decode` take a list of messages
(the Rust shim queues the requests of a short window, rust/codec.rs).  Everything below only drives the
C ABI (include/tfs_gpu.h) through ctypes — no arithmetic happens in Python.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import numpy as np

from . import abi
from ._lib import TfsError, lib
from .messages import BuiltBatch, DecodedMessage, build_batch, messages_from_view

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


class GpuContext:
    """One tfs_ctx_t: one GPU, one stream (one process per GPU; the reference creates its channels in
    TensorflowServingBuilder::build, src/lib.rs:117-145 — this is created alongside)."""

    def __init__(self, device: Optional[int] = None, pinned_bytes: int = 0, use_torch_stream: bool = True):
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
        self.device = device
        self._h = C.c_void_p()
        st = lib().tfs_ctx_create(device, pinned_bytes, C.byref(self._h))
        if st != abi.TFS_OK:
            raise TfsError(st, "tfs_ctx_create failed: no usable CUDA device (this package has no CPU fallback)")
        if use_torch_stream and torch is not None and torch.cuda.is_available():
            torch.cuda.set_device(device)
            # torch reports its default stream as handle 0, which the ABI reads as "the ctx's own stream": name the
            # legacy default stream explicitly (cudaStreamLegacy = 1) so that the codec is ordered after torch's work
            self.set_stream(torch.cuda.current_stream(device).cuda_stream or 1)

    @property
    def handle(self):
        return self._h

    def check(self, st, ok=(abi.TFS_OK,)):
        if st not in ok:
            raise TfsError(st, (lib().tfs_last_error_detail(self._h) or b"").decode("utf-8", "replace"))
        return st

    def set_stream(self, cuda_stream: int):
        self.check(lib().tfs_ctx_set_stream(self._h, C.c_void_p(cuda_stream)))

    def sync(self):
        self.check(lib().tfs_sync(self._h))

    def launch_count(self) -> int:
        return int(lib().tfs_launch_count(self._h))

    def prof_enable(self, on: bool = True):
        self.check(lib().tfs_prof_enable(self._h, 1 if on else 0))

    def prof_read(self, kernel_id: int, reset: bool = False):
        """-> (launches, total milliseconds) of one kernel since profiling was enabled / last reset."""
        n, ms = C.c_uint64(0), C.c_double(0.0)
        self.check(lib().tfs_prof_read(self._h, kernel_id, C.byref(n), C.byref(ms), 1 if reset else 0))
        return int(n.value), float(ms.value)

    def close(self):
        if self._h:
            lib().tfs_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class EncodedBatch:
    """Result of an encode call.  `offsets` / `lengths` are Python lists built on first use; `offsets_np` / `lengths_np`
    are the uint64 arrays the library filled (a batch of 10^6 tiny messages should never pay for the lists)."""

    def __init__(self, wire, offsets, lengths, total: int, flags: int = 0):
        self.wire = wire        # torch.uint8 CUDA tensor (device path) or numpy uint8 array (host path)
        self.offsets_np = np.asarray(offsets, dtype=np.uint64)  # first byte of each message (of its gRPC prefix with F_GRPC_PREFIX)
        self.lengths_np = np.asarray(lengths, dtype=np.uint64)  # protobuf length of each message (prefix excluded)
        self.total = total
        self.flags = flags
        self._offsets = self._lengths = None

    @property
    def offsets(self) -> List[int]:
        if self._offsets is None:
            self._offsets = self.offsets_np.astype(np.int64).tolist()
        return self._offsets

    @property
    def lengths(self) -> List[int]:
        if self._lengths is None:
            self._lengths = self.lengths_np.astype(np.int64).tolist()
        return self._lengths

    def message_bytes(self, i: int) -> bytes:
        pre = 5 if self.flags & abi.F_GRPC_PREFIX else 0
        w = self.wire
        lo, hi = self.offsets[i], self.offsets[i] + self.lengths[i] + pre
        if isinstance(w, np.ndarray):
            return bytes(w[lo:hi])
        return bytes(w[lo:hi].cpu().numpy())

    def all_bytes(self) -> List[bytes]:
        w = self.wire if isinstance(self.wire, np.ndarray) else self.wire[: self.total].cpu().numpy()
        pre = 5 if self.flags & abi.F_GRPC_PREFIX else 0
        return [bytes(w[o: o + l + pre]) for o, l in zip(self.offsets, self.lengths)]


class GpuEncoder:
    """tonic `Encoder` for PredictRequest / PredictResponse / TensorProto batches."""

    def __init__(self, ctx: GpuContext):
        self.ctx = ctx

    def encoded_len(self, messages: Sequence, kind: int) -> List[int]:
        """prost `Message::encoded_len` of every message (payload arrays on the device)."""
        built = messages if isinstance(messages, BuiltBatch) else build_batch(messages, kind)
        out = (C.c_uint64 * max(1, built.n_msgs))()
        self.ctx.check(lib().tfs_encode_size(self.ctx.handle, C.byref(built.batch), out))
        return [int(out[i]) for i in range(built.n_msgs)]

    def bound(self, built: BuiltBatch, flags: int = 0) -> int:
        return int(lib().tfs_encode_bound(C.byref(built.batch), flags))

    def encode(self, messages: Sequence, kind: int, flags: int = 0, dst=None) -> EncodedBatch:
        """Device path: payload arrays are CUDA tensors, the wire bytes stay in HBM."""
        built = messages if isinstance(messages, BuiltBatch) else build_batch(messages, kind)
        if dst is None:
            dst = torch.empty(self.bound(built, flags), dtype=torch.uint8, device=f"cuda:{self.ctx.device}")
        n = built.n_msgs
        off, ln, total = (C.c_uint64 * max(1, n))(), (C.c_uint64 * max(1, n))(), C.c_uint64(0)
        self.ctx.check(lib().tfs_encode(self.ctx.handle, C.byref(built.batch), C.c_void_p(dst.data_ptr()), dst.numel(), flags,
                                        off, ln, C.byref(total)))
        return EncodedBatch(dst, np.ctypeslib.as_array(off)[:n].copy(), np.ctypeslib.as_array(ln)[:n].copy(),
                            int(total.value), flags)

    def encode_host(self, messages: Sequence, kind: int, flags: int = 0, dst: Optional[np.ndarray] = None) -> EncodedBatch:
        """End to end: payload arrays in host memory, wire bytes returned in host memory."""
        built = messages if isinstance(messages, BuiltBatch) else build_batch(messages, kind)
        if dst is None:
            dst = np.empty(self.bound(built, flags), dtype=np.uint8)
        n = built.n_msgs
        off, ln, total = (C.c_uint64 * max(1, n))(), (C.c_uint64 * max(1, n))(), C.c_uint64(0)
        self.ctx.check(lib().tfs_encode_host(self.ctx.handle, C.byref(built.batch), C.c_void_p(dst.ctypes.data), dst.size, flags,
                                             off, ln, C.byref(total)))
        return EncodedBatch(dst, np.ctypeslib.as_array(off)[:n].copy(), np.ctypeslib.as_array(ln)[:n].copy(),
                            int(total.value), flags)


class DecodedBatch:
    """Result of the decode scan: plain-data view + (after `values`) the value pool."""

    def __init__(self, ctx: GpuContext, handle, status: int, kind: int):
        self.ctx, self._h, self.status, self.kind = ctx, handle, status, kind
        self.view = abi.DecodedView()
        ctx.check(lib().tfs_decoded_view(self._h, C.byref(self.view)))
        self.pool = None

    @property
    def pool_bytes(self) -> int:
        return int(self.view.pool_bytes)

    def messages(self) -> List[DecodedMessage]:
        pool = self.pool
        if pool is None:
            raise RuntimeError("values have not been decoded yet")
        if not isinstance(pool, np.ndarray):
            pool = pool.cpu().numpy()
        return messages_from_view(self.view, pool, self.kind)

    def close(self):
        if self._h:
            lib().tfs_decoded_free(self.ctx.handle, self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GpuDecoder:
    """tonic `Decoder` for PredictResponse / PredictRequest / TensorProto batches."""

    def __init__(self, ctx: GpuContext):
        self.ctx = ctx

    @staticmethod
    def _arrays(offsets, lengths):
        """(n, uint64 ctypes array of offsets, of lengths): through numpy, a Python-level loop costs milliseconds for 10^4 messages"""
        n = len(offsets)
        o = np.ascontiguousarray(offsets, dtype=np.uint64) if n else np.zeros(1, np.uint64)
        ln = np.ascontiguousarray(lengths, dtype=np.uint64) if n else np.zeros(1, np.uint64)
        return n, np.ctypeslib.as_ctypes(o), np.ctypeslib.as_ctypes(ln)

    def scan(self, wire, offsets, lengths, kind: int, flags: int = 0) -> DecodedBatch:
        """Pass 1 on device-resident wire bytes (torch.uint8 CUDA tensor)."""
        n, off, ln = self._arrays(offsets, lengths)
        h = C.c_void_p()
        st = lib().tfs_decode_scan(self.ctx.handle, C.c_void_p(wire.data_ptr() if wire.numel() else 0), wire.numel(), off, ln, n, kind,
                                   flags, C.byref(h))
        self.ctx.check(st, ok=(abi.TFS_OK, abi.TFS_ERR_BATCH))
        return DecodedBatch(self.ctx, h, st, kind)

    def values(self, dec: DecodedBatch, pool=None) -> DecodedBatch:
        """Pass 2: write every value into a device pool."""
        if pool is None:
            pool = torch.empty(max(1, dec.pool_bytes), dtype=torch.uint8, device=f"cuda:{self.ctx.device}")
        self.ctx.check(lib().tfs_decode(self.ctx.handle, dec._h, C.c_void_p(pool.data_ptr()), pool.numel()))
        dec.pool = pool
        return dec

    def decode(self, wire, offsets, lengths, kind: int, flags: int = 0, pool=None) -> DecodedBatch:
        return self.values(self.scan(wire, offsets, lengths, kind, flags), pool)

    def decode_into(self, wire, offsets, lengths, kind: int, pool, flags: int = 0) -> DecodedBatch:
        """Both passes in one call into a device pool of known capacity (tfs_decode_into)."""
        n, off, ln = self._arrays(offsets, lengths)
        h = C.c_void_p()
        st = lib().tfs_decode_into(self.ctx.handle, C.c_void_p(wire.data_ptr() if wire.numel() else 0), wire.numel(), off, ln, n, kind,
                                   flags, C.c_void_p(pool.data_ptr()), pool.numel(), C.byref(h))
        if st == abi.TFS_ERR_NOMEM and h:
            lib().tfs_decoded_free(self.ctx.handle, h)
        self.ctx.check(st, ok=(abi.TFS_OK, abi.TFS_ERR_BATCH))
        dec = DecodedBatch(self.ctx, h, st, kind)
        dec.pool = pool
        return dec

    def decode_host_into(self, wire: np.ndarray, offsets, lengths, kind: int, pool: np.ndarray, flags: int = 0) -> DecodedBatch:
        """Host-memory form of decode_into: pinned buffers are read / written in place by the kernels."""
        n, off, ln = self._arrays(offsets, lengths)
        h = C.c_void_p()
        st = lib().tfs_decode_host_into(self.ctx.handle, C.c_void_p(wire.ctypes.data if wire.size else 0), wire.size, off, ln, n, kind,
                                        flags, C.c_void_p(pool.ctypes.data), pool.size, C.byref(h))
        if st == abi.TFS_ERR_NOMEM and h:
            lib().tfs_decoded_free(self.ctx.handle, h)
        self.ctx.check(st, ok=(abi.TFS_OK, abi.TFS_ERR_BATCH))
        dec = DecodedBatch(self.ctx, h, st, kind)
        dec.pool = pool
        return dec

    def decode_host(self, wire: np.ndarray, offsets, lengths, kind: int, flags: int = 0, pool: Optional[np.ndarray] = None) -> DecodedBatch:
        """End to end: wire bytes in host memory, values returned in a host pool."""
        wire = np.ascontiguousarray(wire, dtype=np.uint8)
        n, off, ln = self._arrays(offsets, lengths)
        h = C.c_void_p()
        st = lib().tfs_decode_host(self.ctx.handle, C.c_void_p(wire.ctypes.data if wire.size else 0), wire.size, off, ln, n, kind, flags,
                                   C.byref(h))
        self.ctx.check(st, ok=(abi.TFS_OK, abi.TFS_ERR_BATCH))
        dec = DecodedBatch(self.ctx, h, st, kind)
        if pool is None:
            pool = np.empty(max(1, dec.pool_bytes), dtype=np.uint8)
        self.ctx.check(lib().tfs_decode_host_finish(self.ctx.handle, dec._h, C.c_void_p(pool.ctypes.data), pool.size))
        dec.pool = pool
        return dec


class GpuCodec:
    """`impl tonic::codec::Codec for GpuCodec { fn encoder(..); fn decoder(..) }`"""

    def __init__(self, ctx: Optional[GpuContext] = None):
        self.ctx = ctx or GpuContext()

    def encoder(self) -> GpuEncoder:
        return GpuEncoder(self.ctx)

    def decoder(self) -> GpuDecoder:
        return GpuDecoder(self.ctx)


# ------------------------------------------------------------------------------------------------
# staging pool, engine (worker threads over one or several GPUs), coalescer
# ------------------------------------------------------------------------------------------------
class Slab:
    """A pinned slab out of a context's staging pool (tfs_stage_acquire): `.array` is a numpy uint8 view of it."""

    def __init__(self, ctx_handle, nbytes: int, align_skew: int = 0):
        self._ctx = ctx_handle
        self.raw = abi.Slab()
        st = lib().tfs_stage_acquire(ctx_handle, nbytes, align_skew, C.byref(self.raw))
        if st != abi.TFS_OK:
            raise TfsError(st, "tfs_stage_acquire")
        self.array = np.ctypeslib.as_array(C.cast(self.raw.ptr, C.POINTER(C.c_uint8)), shape=(int(self.raw.capacity),))

    @property
    def address(self) -> int:
        return int(self.raw.ptr)

    def release(self):
        if self.raw.ptr:
            lib().tfs_stage_release(self._ctx, C.byref(self.raw))
            self.raw = abi.Slab()
            self.array = None


class Engine:
    """tfs_engine_t: `lanes_per_device` contexts (stream + staging pool + host worker thread each) on every listed GPU;
    jobs are submitted asynchronously and waited for by ticket."""

    def __init__(self, devices: Sequence[int], lanes_per_device: int = 2, pinned_bytes_per_lane: int = 0):
        self._h = C.c_void_p()
        arr = (C.c_int32 * len(devices))(*devices)
        st = lib().tfs_engine_create(arr, len(devices), lanes_per_device, pinned_bytes_per_lane, C.byref(self._h))
        if st != abi.TFS_OK:
            raise TfsError(st, "tfs_engine_create")
        self._keep = {}

    @property
    def lanes(self) -> int:
        return int(lib().tfs_engine_lanes(self._h))

    def lane_ctx(self, lane: int):
        return C.c_void_p(lib().tfs_engine_ctx(self._h, lane))

    def slab(self, lane: int, nbytes: int, align_skew: int = 0) -> Slab:
        return Slab(self.lane_ctx(lane), nbytes, align_skew)

    def encode_host(self, built: BuiltBatch, dst: np.ndarray, flags: int = 0, lane: int = -1) -> int:
        """-> ticket; `wait_encode(ticket)` returns the EncodedBatch"""
        n = built.n_msgs
        off, ln, total = (C.c_uint64 * max(1, n))(), (C.c_uint64 * max(1, n))(), C.c_uint64(0)
        t = C.c_uint64(0)
        st = lib().tfs_engine_encode_host(self._h, lane, C.byref(built.batch), C.c_void_p(dst.ctypes.data), dst.size, flags, off, ln, C.byref(total), C.byref(t))
        if st != abi.TFS_OK:
            raise TfsError(st, "tfs_engine_encode_host")
        self._keep[t.value] = ("enc", built, dst, off, ln, total, flags, n)
        return int(t.value)

    def decode_host_into(self, wire: np.ndarray, offsets, lengths, kind: int, pool: np.ndarray, flags: int = 0, lane: int = -1) -> int:
        n, off, ln = GpuDecoder._arrays(offsets, lengths)
        h, t = C.c_void_p(), C.c_uint64(0)
        hp = C.pointer(h)
        st = lib().tfs_engine_decode_host_into(self._h, lane, C.c_void_p(wire.ctypes.data if wire.size else 0), wire.size, off, ln, n, kind, flags,
                                               C.c_void_p(pool.ctypes.data), pool.size, hp, C.byref(t))
        if st != abi.TFS_OK:
            raise TfsError(st, "tfs_engine_decode_host_into")
        self._keep[t.value] = ("dec", wire, off, ln, pool, h, hp, kind)
        return int(t.value)

    def wait(self, ticket: int):
        """Blocks until the job is done: an EncodedBatch for an encode ticket, an EngineDecoded for a decode ticket."""
        detail = C.create_string_buffer(256)
        st = lib().tfs_engine_wait(self._h, ticket, detail)
        rec = self._keep.pop(ticket)
        if rec[0] == "enc":
            _, built, dst, off, ln, total, flags, n = rec
            if st != abi.TFS_OK:
                raise TfsError(st, detail.value.decode("utf-8", "replace"))
            return EncodedBatch(dst, np.ctypeslib.as_array(off)[:n].copy(), np.ctypeslib.as_array(ln)[:n].copy(),
                                int(total.value), flags)
        _, wire, off, ln, pool, h, hp, kind = rec
        if st not in (abi.TFS_OK, abi.TFS_ERR_BATCH):
            if h:
                lib().tfs_engine_decoded_free(self._h, h)
            raise TfsError(st, detail.value.decode("utf-8", "replace"))
        return EngineDecoded(self, h, st, kind, pool)

    def decode_host_sharded(self, wire: np.ndarray, offsets, lengths, kind: int, pools: Sequence[np.ndarray], flags: int = 0):
        """One batch over all lanes (tfs_engine_decode_host_sharded) -> (status, [EngineDecoded or None per lane], first message per lane)"""
        n, off, ln = GpuDecoder._arrays(offsets, lengths)
        L = self.lanes
        pp = (C.c_void_p * L)(*[p.ctypes.data for p in pools])
        pc = (C.c_uint64 * L)(*[p.size for p in pools])
        outs = (C.c_void_p * L)()
        first = (C.c_uint32 * L)()
        st = lib().tfs_engine_decode_host_sharded(self._h, C.c_void_p(wire.ctypes.data if wire.size else 0), wire.size, off, ln, n, kind, flags, pp, pc, outs, first)
        if st not in (abi.TFS_OK, abi.TFS_ERR_BATCH):
            for i in range(L):
                if outs[i]:
                    lib().tfs_engine_decoded_free(self._h, C.c_void_p(outs[i]))
            raise TfsError(st, "tfs_engine_decode_host_sharded")
        res = [EngineDecoded(self, C.c_void_p(outs[i]), st, kind, pools[i]) if outs[i] else None for i in range(L)]
        return st, res, [int(first[i]) for i in range(L)]

    def close(self):
        if self._h:
            lib().tfs_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class EngineDecoded:
    """Decode result of an engine job (the handle belongs to the lane that made it)."""

    def __init__(self, eng: Engine, handle, status: int, kind: int, pool: np.ndarray):
        self.eng, self._h, self.status, self.kind, self.pool = eng, handle, status, kind, pool
        self.view = abi.DecodedView()
        st = lib().tfs_decoded_view(self._h, C.byref(self.view))
        if st != abi.TFS_OK:
            raise TfsError(st, "tfs_decoded_view")

    @property
    def pool_bytes(self) -> int:
        return int(self.view.pool_bytes)

    def messages(self) -> List[DecodedMessage]:
        return messages_from_view(self.view, self.pool, self.kind)

    def close(self):
        if self._h:
            lib().tfs_engine_decoded_free(self.eng._h, self._h)
            self._h = C.c_void_p()


class Coalescer:
    """tfs_coalescer_t: single-message encode calls from many threads, one batch (one launch) per window."""

    def __init__(self, eng: Engine, window_us: int = 200, max_msgs: int = 64):
        self.eng = eng
        self._h = C.c_void_p()
        st = lib().tfs_coalescer_create(eng._h, window_us, max_msgs, C.byref(self._h))
        if st != abi.TFS_OK:
            raise TfsError(st, "tfs_coalescer_create")

    def encode(self, message, kind: int, flags: int = 0) -> bytes:
        """Blocks until the batch this message joined is encoded; returns a COPY of its bytes (and releases the lease: a
        transport would send straight from the pinned slab and release afterwards)."""
        built = build_batch([message], kind)
        wire, ln, lease = C.c_void_p(), C.c_uint64(0), C.c_uint64(0)
        st = lib().tfs_coalescer_encode(self._h, C.byref(built.batch), flags, C.byref(wire), C.byref(ln), C.byref(lease))
        if st != abi.TFS_OK:
            raise TfsError(st, "tfs_coalescer_encode")
        out = C.string_at(wire, ln.value)
        lib().tfs_coalescer_release(self._h, lease)
        return out

    def stats(self):
        b, m = C.c_uint64(0), C.c_uint64(0)
        lib().tfs_coalescer_stats(self._h, C.byref(b), C.byref(m))
        return int(b.value), int(m.value)

    def close(self):
        if self._h:
            lib().tfs_coalescer_destroy(self._h)
            self._h = C.c_void_p()
