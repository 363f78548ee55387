"""CPU tests (no GPU): the C-ABI library builds, loads and exports exactly what include/tfs_gpu.h declares;
host-only entry points work; compute entry points fail loudly without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import _lib, abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "tfs_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tfs_[a-z0-9_]+)\s*\(", src)))


def test_library_is_built_and_loads():
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"
    L = _lib.lib()
    assert L.tfs_abi_version() == abi.TFS_ABI_VERSION


def test_exports_match_header():
    declared = header_functions()
    assert sorted(abi.EXPORTED_SYMBOLS) == declared
    L = C.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/tfs_gpu.h but not exported"


def test_struct_layouts_match_header():
    # sizes the C compiler gives the structs (spot-checked constants; a layout drift breaks the ABI)
    assert C.sizeof(abi.Span) == 16 and C.sizeof(abi.Str) == 16
    assert C.sizeof(abi.FieldDesc) == 48
    assert C.sizeof(abi.TensorDesc) == 16 + 4 * 8 + 8 * abi.TFS_MAX_DIMS
    assert C.sizeof(abi.MsgDesc) == 8 + 3 * 16 + 8 + 8 + 16
    assert C.sizeof(abi.TensorOut) == 16 + 6 * 4 + 8 * abi.TFS_MAX_DIMS + 20 * abi.TFS_N_SLOTS
    assert C.sizeof(abi.MsgOut) == 16 + 3 * 16 + 8 + 8 + 16


def test_host_only_helpers():
    L = _lib.lib()
    assert [L.tfs_slot_of_kind(k) for k in abi.SLOT_KINDS] == list(range(abi.TFS_N_SLOTS))
    assert [L.tfs_kind_of_slot(s) for s in range(abi.TFS_N_SLOTS)] == list(abi.SLOT_KINDS)
    assert L.tfs_slot_of_kind(18) == -1 and L.tfs_slot_of_kind(14) == 12 and L.tfs_slot_of_kind(15) == 13
    for t, size in abi.TYPE_SIZE.items():
        assert L.tfs_type_size(t) == size
    assert L.tfs_strerror(abi.TFS_ERR_INVALID_VARINT) == b"invalid varint"
    assert L.tfs_strerror(abi.TFS_ERR_LENGTH_EXCEEDED) == b"delimited length exceeded"


def test_encode_bound_is_an_upper_bound():
    import oracle_bind as ob
    msgs = tfs.workloads.c5_responses(50, seed=3, max_log2=9) if hasattr(tfs, "workloads") else None
    from tfs_b200 import workloads
    msgs = workloads.c5_responses(50, seed=3, max_log2=9)
    built = tfs.build_batch(msgs, abi.MSG_PREDICT_RESPONSE)
    sizes = ob.encode_size(built)
    L = _lib.lib()
    for flags in (0, abi.F_GRPC_PREFIX, abi.F_ALIGN_PAYLOAD | abi.F_GRPC_PREFIX):
        bound = L.tfs_encode_bound(C.byref(built.batch), flags)
        assert bound >= sum(sizes) + (5 * len(sizes) if flags & abi.F_GRPC_PREFIX else 0) + (16 * len(sizes) if flags & abi.F_ALIGN_PAYLOAD else 0)
        assert bound < 12 * sum(sizes) + 4096 * len(sizes)
    bad = tfs.build_batch([tfs.TensorProto(int64_val=np.zeros(3, np.float32))], abi.MSG_TENSOR_PROTO)
    assert L.tfs_encode_bound(C.byref(bad.batch), 0) == 0


def test_shard_plan_balances_bytes():
    rng = np.random.default_rng(0)
    sizes = (2 ** rng.uniform(0, 18, 1000)).astype(np.int64).tolist()
    for n in (1, 2, 4, 8):
        assign = _lib.shard_plan(sizes, n)
        loads = [sum(s for s, a in zip(sizes, assign) if a == k) for k in range(n)]
        assert set(assign) <= set(range(n))
        assert max(loads) - min(loads) <= max(sizes)
    assert _lib.shard_plan([], 4) == []
    assert sorted(_lib.shard_plan([0, 0, 0, 0], 4)) == [0, 1, 2, 3]


def test_no_cpu_fallback_without_a_device():
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is visible")
    except ImportError:
        pass
    from tfs_b200.codec import GpuContext
    with pytest.raises(_lib.TfsError) as e:
        GpuContext(0)
    assert e.value.status == abi.TFS_ERR_CUDA


def test_product_package_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "rust-tensorflow-serving2_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp", ".rs")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in text.lower() or f == "workloads.py" and "oracle" not in text, os.path.join(dirpath, f)


def test_constants_match_header():
    """Every enum constant of include/tfs_gpu.h that the Python harness mirrors has the header's value."""
    import re
    text = open(os.path.join(ROOT, "include", "tfs_gpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    found = {}
    for name, val in re.findall(r"\b(TFS_[A-Z0-9_]+)\s*=\s*(1u\s*<<\s*\d+|\d+)", text):
        found[name] = eval(val.replace("u", ""))
    for m in re.findall(r"#define\s+(TFS_[A-Z0-9_]+)\s+(\d+)\b", text):
        found[m[0]] = int(m[1])
    assert len(found) > 60
    checked = 0
    for name, val in found.items():
        for cand in (name, name[4:]):  # TFS_ERR_* keep their prefix in abi.py, the others drop "TFS_"
            if hasattr(abi, cand):
                assert getattr(abi, cand) == val, (name, getattr(abi, cand), val)
                checked += 1
                break
    assert checked > 50, checked


def test_cpp_mirror_header_compiles():
    """include/tfs_codec.hpp (the C++ mirror of the reference's codec seam) and its test program compile cleanly."""
    import shutil
    import subprocess
    if not shutil.which("g++"):
        pytest.skip("no g++")
    src = os.path.join(ROOT, "tests", "cpp", "test_codec.cpp")
    r = subprocess.run(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-fsyntax-only", src], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[:2000]


def test_shard_ranges_are_contiguous_and_balanced():
    L = _lib.lib()
    rng = np.random.default_rng(8)
    for n_msgs, n_shards in ((0, 3), (1, 4), (17, 4), (1000, 8), (8192, 8)):
        lens = (2 ** rng.uniform(0, 18, n_msgs)).astype(np.uint64)
        arr = (C.c_uint64 * max(1, n_msgs))(*[int(x) for x in lens])
        first = (C.c_uint32 * (n_shards + 1))()
        assert L.tfs_shard_ranges(arr, n_msgs, n_shards, first) == abi.TFS_OK
        f = list(first)
        assert f[0] == 0 and f[-1] == n_msgs and f == sorted(f)
        if n_msgs >= 1000:
            per = [int(lens[a:b].sum()) for a, b in zip(f[:-1], f[1:])]
            assert max(per) - min(per) <= 2 * int(lens.max())
    assert L.tfs_shard_ranges(None, 0, 0, (C.c_uint32 * 1)()) == abi.TFS_ERR_INVALID_ARG
