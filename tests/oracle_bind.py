"""ctypes binding of the CPU oracle (oracle/libtfs_oracle.so) — checker only.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

import tfs_b200 as tfs
from tfs_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libtfs_oracle.so")

_lib = None


def build(force=False):
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("cpu_ref.cpp", "cpu_ref_capi.cpp", "cpu_ref.hpp")]
    srcs.append(os.path.join(ROOT, "include", "tfs_gpu.h"))
    have_src = all(os.path.exists(s) for s in srcs)
    stale = (not os.path.exists(ORACLE_SO)) or (have_src and any(os.path.getmtime(s) > os.path.getmtime(ORACLE_SO) for s in srcs))
    if force or stale:
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(ORACLE_SO)
        L.oracle_encode_size.argtypes = [C.POINTER(abi.Batch), C.POINTER(C.c_uint64)]
        L.oracle_encode.argtypes = [C.POINTER(abi.Batch), C.c_void_p, C.c_uint64, C.c_uint32, C.POINTER(C.c_uint64),
                                    C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.oracle_decode.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_uint32, C.c_int32,
                                    C.c_uint32, C.POINTER(C.c_void_p)]
        L.oracle_decoded_view.argtypes = [C.c_void_p, C.POINTER(abi.DecodedView), C.POINTER(C.c_void_p)]
        L.oracle_decoded_free.argtypes = [C.c_void_p]
        L.oracle_prepare.argtypes = [C.POINTER(abi.Batch), C.POINTER(C.c_void_p)]
        L.oracle_prepared_free.argtypes = [C.c_void_p]
        L.oracle_run_encode.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_uint64)]
        L.oracle_run_encode.restype = C.c_double
        L.oracle_run_decode.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.oracle_run_decode.restype = C.c_double
        L.oracle_encoded_len_varint.argtypes = [C.c_uint64]
        L.oracle_encoded_len_varint.restype = C.c_uint64
        for n in ("oracle_f32_to_f16", "oracle_f32_to_bf16"):
            getattr(L, n).argtypes = [C.c_float]
            getattr(L, n).restype = C.c_uint32
        for n in ("oracle_f16_to_f32", "oracle_bf16_to_f32"):
            getattr(L, n).argtypes = [C.c_uint32]
            getattr(L, n).restype = C.c_float
        L.oracle_is_utf8.argtypes = [C.c_char_p, C.c_uint64]
        _lib = L
    return _lib


def encode_size(built):
    out = (C.c_uint64 * max(1, built.n_msgs))()
    st = lib().oracle_encode_size(C.byref(built.batch), out)
    if st != 0:
        raise ValueError(f"oracle_encode_size: {abi.STATUS_NAMES.get(st, st)}")
    return [int(out[i]) for i in range(built.n_msgs)]


def encode(built, flags=0):
    """-> (wire: np.uint8 array, offsets, lengths)"""
    sizes = encode_size(built)
    cap = sum(sizes) + 5 * built.n_msgs + 16
    dst = np.zeros(cap, dtype=np.uint8)
    off = (C.c_uint64 * max(1, built.n_msgs))()
    ln = (C.c_uint64 * max(1, built.n_msgs))()
    total = C.c_uint64(0)
    st = lib().oracle_encode(C.byref(built.batch), dst.ctypes.data, cap, flags, off, ln, C.byref(total))
    if st != 0:
        raise ValueError(f"oracle_encode: {abi.STATUS_NAMES.get(st, st)}")
    n = built.n_msgs
    return dst[: total.value], [int(off[i]) for i in range(n)], [int(ln[i]) for i in range(n)]


def encode_messages(messages, kind, flags=0):
    """-> list of bytes, one per message"""
    built = tfs.build_batch(messages, kind)
    wire, off, ln = encode(built, flags)
    pre = 5 if flags & abi.F_GRPC_PREFIX else 0
    return [bytes(wire[o: o + l + pre]) for o, l in zip(off, ln)]


def decode(wire, offsets, lengths, kind, flags=0):
    """-> (status, list[DecodedMessage]) from the oracle."""
    wire = np.ascontiguousarray(wire, dtype=np.uint8)
    n = len(offsets)
    off = (C.c_uint64 * max(1, n))(*offsets)
    ln = (C.c_uint64 * max(1, n))(*lengths)
    h = C.c_void_p()
    st = lib().oracle_decode(wire.ctypes.data if wire.size else None, off, ln, n, kind, flags, C.byref(h))
    if st not in (0, abi.TFS_ERR_BATCH):
        raise ValueError(f"oracle_decode: {abi.STATUS_NAMES.get(st, st)}")
    try:
        view = abi.DecodedView()
        pool_p = C.c_void_p()
        lib().oracle_decoded_view(h, C.byref(view), C.byref(pool_p))
        pool = np.frombuffer(C.string_at(pool_p, view.pool_bytes), dtype=np.uint8) if view.pool_bytes else np.zeros(0, np.uint8)
        msgs = tfs.messages_from_view(view, pool.copy(), kind)
    finally:
        lib().oracle_decoded_free(h)
    return st, msgs


def decode_bytes(blobs, kind, flags=0):
    """Decode a list of independent byte strings."""
    offs, lens, cur = [], [], 0
    for b in blobs:
        offs.append(cur)
        lens.append(len(b))
        cur += len(b)
    wire = np.frombuffer(b"".join(blobs), dtype=np.uint8) if cur else np.zeros(0, np.uint8)
    return decode(wire, offs, lens, kind, flags)


class Prepared:
    """Timed CPU baseline handle (prost structs built once, untimed)."""

    def __init__(self, built):
        self.h = C.c_void_p()
        st = lib().oracle_prepare(C.byref(built.batch), C.byref(self.h))
        if st != 0:
            raise ValueError(f"oracle_prepare: {abi.STATUS_NAMES.get(st, st)}")

    def run_encode(self, threads):
        wb = C.c_uint64(0)
        s = lib().oracle_run_encode(self.h, threads, C.byref(wb))
        return s, wb.value

    def run_decode(self, threads):
        wb, ck = C.c_uint64(0), C.c_uint64(0)
        s = lib().oracle_run_decode(self.h, threads, C.byref(wb), C.byref(ck))
        return s, wb.value, ck.value

    def close(self):
        if self.h:
            lib().oracle_prepared_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
