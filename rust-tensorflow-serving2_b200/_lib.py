"""ctypes loader of csrc/libtfs_gpu.so (the C ABI of include/tfs_gpu.h).

Fails loudly: no built library, or no CUDA device at context creation, raises — there is no CPU
fallback anywhere in this package.
"""
import ctypes as C
import os
import subprocess

from . import abi

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.environ.get("TFS_LIB_PATH") or os.path.join(CSRC, "libtfs_gpu.so")  # (the override is for A/B runs of two builds)

_lib = None


class TfsError(RuntimeError):
    def __init__(self, status, detail=""):
        self.status = status
        super().__init__(f"tfs status {status} ({abi.STATUS_NAMES.get(status, '?')}){': ' + detail if detail else ''}")


def build(force=False):
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    srcs.append(os.path.join(os.path.dirname(HERE), "include", "tfs_gpu.h"))
    stale = not os.path.exists(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs if os.path.exists(s))
    if force or stale:
        subprocess.check_call(["make", "-C", CSRC, "-s"] + (["-B"] if force else []))
    return LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: the CUDA extension is not built (run __graft_entry__.build()); "
                           "this package has no CPU fallback")
    L = C.CDLL(LIB_PATH)
    u64p, u32p, vp = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.c_void_p
    L.tfs_abi_version.restype = C.c_int32
    L.tfs_strerror.restype = C.c_char_p
    L.tfs_strerror.argtypes = [C.c_int32]
    L.tfs_ctx_create.argtypes = [C.c_int32, C.c_uint64, C.POINTER(vp)]
    L.tfs_ctx_destroy.argtypes = [vp]
    L.tfs_last_error_detail.restype = C.c_char_p
    L.tfs_last_error_detail.argtypes = [vp]
    L.tfs_last_error_location.argtypes = [vp, u32p, u64p]
    L.tfs_ctx_set_stream.argtypes = [vp, vp]
    L.tfs_sync.argtypes = [vp]
    L.tfs_launch_count.restype = C.c_uint64
    L.tfs_launch_count.argtypes = [vp]
    L.tfs_host_alloc.argtypes = [vp, C.c_uint64, C.POINTER(vp)]
    L.tfs_host_free.argtypes = [vp, vp]
    L.tfs_dev_alloc.argtypes = [vp, C.c_uint64, C.POINTER(vp)]
    L.tfs_dev_free.argtypes = [vp, vp]
    L.tfs_copy_h2d.argtypes = [vp, vp, vp, C.c_uint64]
    L.tfs_copy_d2h.argtypes = [vp, vp, vp, C.c_uint64]
    L.tfs_encode_size.argtypes = [vp, C.POINTER(abi.Batch), u64p]
    L.tfs_encode_bound.restype = C.c_uint64
    L.tfs_encode_bound.argtypes = [C.POINTER(abi.Batch), C.c_uint32]
    L.tfs_encode.argtypes = [vp, C.POINTER(abi.Batch), vp, C.c_uint64, C.c_uint32, u64p, u64p, u64p]
    L.tfs_encode_host.argtypes = [vp, C.POINTER(abi.Batch), vp, C.c_uint64, C.c_uint32, u64p, u64p, u64p]
    L.tfs_decode_scan.argtypes = [vp, vp, C.c_uint64, u64p, u64p, C.c_uint32, C.c_int32, C.c_uint32, C.POINTER(vp)]
    L.tfs_decoded_view.argtypes = [vp, C.POINTER(abi.DecodedView)]
    L.tfs_decode.argtypes = [vp, vp, vp, C.c_uint64]
    L.tfs_decoded_free.argtypes = [vp, vp]
    L.tfs_decode_host.argtypes = [vp, vp, C.c_uint64, u64p, u64p, C.c_uint32, C.c_int32, C.c_uint32, C.POINTER(vp)]
    L.tfs_decode_host_finish.argtypes = [vp, vp, vp, C.c_uint64]
    L.tfs_decode_into.argtypes = [vp, vp, C.c_uint64, u64p, u64p, C.c_uint32, C.c_int32, C.c_uint32, vp, C.c_uint64, C.POINTER(vp)]
    L.tfs_decode_host_into.argtypes = [vp, vp, C.c_uint64, u64p, u64p, C.c_uint32, C.c_int32, C.c_uint32, vp, C.c_uint64, C.POINTER(vp)]
    L.tfs_shard_plan.argtypes = [u64p, C.c_uint32, C.c_uint32, u32p]
    L.tfs_slot_of_kind.argtypes = [C.c_int32]
    L.tfs_kind_of_slot.argtypes = [C.c_int32]
    L.tfs_type_size.restype = C.c_uint32
    L.tfs_type_size.argtypes = [C.c_int32]
    L.tfs_pack_int64_noalloc.argtypes = [vp, C.c_uint64, vp, C.c_uint64, vp, vp]
    L.tfs_stage_acquire.argtypes = [vp, C.c_uint64, C.c_uint32, C.POINTER(abi.Slab)]
    L.tfs_stage_release.argtypes = [vp, C.POINTER(abi.Slab)]
    L.tfs_stage_stats.argtypes = [vp, u64p, u64p, u64p]
    L.tfs_engine_create.argtypes = [C.POINTER(C.c_int32), C.c_uint32, C.c_uint32, C.c_uint64, C.POINTER(vp)]
    L.tfs_engine_destroy.argtypes = [vp]
    L.tfs_engine_lanes.restype = C.c_uint32
    L.tfs_engine_lanes.argtypes = [vp]
    L.tfs_engine_ctx.restype = vp
    L.tfs_engine_ctx.argtypes = [vp, C.c_uint32]
    L.tfs_engine_encode_host.argtypes = [vp, C.c_int32, C.POINTER(abi.Batch), vp, C.c_uint64, C.c_uint32, u64p, u64p, u64p, u64p]
    L.tfs_engine_decode_host_into.argtypes = [vp, C.c_int32, vp, C.c_uint64, u64p, u64p, C.c_uint32, C.c_int32, C.c_uint32, vp, C.c_uint64,
                                              C.POINTER(vp), u64p]
    L.tfs_engine_wait.argtypes = [vp, C.c_uint64, C.c_char_p]
    L.tfs_engine_decoded_free.argtypes = [vp, vp]
    L.tfs_engine_decode_host_sharded.argtypes = [vp, vp, C.c_uint64, u64p, u64p, C.c_uint32, C.c_int32, C.c_uint32, C.POINTER(vp), u64p,
                                                 C.POINTER(vp), u32p]
    L.tfs_shard_ranges.argtypes = [u64p, C.c_uint32, C.c_uint32, u32p]
    L.tfs_coalescer_create.argtypes = [vp, C.c_uint32, C.c_uint32, C.POINTER(vp)]
    L.tfs_coalescer_destroy.argtypes = [vp]
    L.tfs_coalescer_encode.argtypes = [vp, C.POINTER(abi.Batch), C.c_uint32, C.POINTER(vp), u64p, u64p]
    L.tfs_coalescer_release.argtypes = [vp, C.c_uint64]
    L.tfs_coalescer_stats.argtypes = [vp, u64p, u64p]
    L.tfs_prof_enable.argtypes = [vp, C.c_int32]
    L.tfs_prof_read.argtypes = [vp, C.c_int32, u64p, C.POINTER(C.c_double), C.c_int32]
    for name in abi.EXPORTED_SYMBOLS:
        fn = getattr(L, name)
        if fn.restype is C.c_int and name not in ("tfs_abi_version", "tfs_slot_of_kind", "tfs_kind_of_slot", "tfs_engine_lanes", "tfs_engine_ctx"):
            fn.restype = C.c_int32
    _lib = L
    return L


def shard_plan(msg_bytes, n_shards):
    """Greedy largest-first assignment of independent messages to shards (host only)."""
    n = len(msg_bytes)
    arr = (C.c_uint64 * max(1, n))(*[int(b) for b in msg_bytes])
    out = (C.c_uint32 * max(1, n))()
    st = lib().tfs_shard_plan(arr, n, n_shards, out)
    if st != abi.TFS_OK:
        raise TfsError(st)
    return [int(out[i]) for i in range(n)]
