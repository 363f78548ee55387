"""Shared test helpers: golden-vector loading and message comparison."""
import json
import os

import numpy as np

import tfs_b200 as tfs
from tfs_b200 import abi

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "upb_vectors.json")

FIELD_NP = {
    "tensor_content": "u1", "float_val": "<f4", "double_val": "<f8", "int_val": "<i4", "scomplex_val": "<f4",
    "int64_val": "<i8", "bool_val": "u1", "dcomplex_val": "<f8", "half_val": "<i4", "uint32_val": "<u4",
    "uint64_val": "<u8",
}
RAGGED_FIELDS = ("string_val", "resource_handle_val", "variant_val")
KIND_OF = {"tensor": abi.MSG_TENSOR_PROTO, "request": abi.MSG_PREDICT_REQUEST, "response": abi.MSG_PREDICT_RESPONSE,
           "input": abi.MSG_INPUT, "classify": abi.MSG_CLASSIFY_REQUEST}


def feature_from_desc(d) -> tfs.Feature:
    if "bytes_list" in d:
        return tfs.Feature(bytes_list=[bytes.fromhex(x) for x in d["bytes_list"]])
    if "float_list" in d:
        return tfs.Feature(float_list=np.frombuffer(bytes.fromhex(d["float_list"]), dtype="<f4").copy())
    if "int64_list" in d:
        return tfs.Feature(int64_list=np.frombuffer(bytes.fromhex(d["int64_list"]), dtype="<i8").copy())
    return tfs.Feature()


def input_from_desc(kind: str, m):
    """Golden "input" section -> tfs.Input / tfs.ClassificationRequest."""
    inp = tfs.Input([[(k, feature_from_desc(ft)) for k, ft in ex] for ex in m["examples"]])
    if kind == "input":
        return inp
    ms = None
    if m["model_spec"] is not None:
        s = m["model_spec"]
        ms = tfs.ModelSpec(s["name"], s.get("version"), s.get("version_label"), s["signature_name"])
    return tfs.ClassificationRequest(ms, inp)

_golden = None


def golden():
    global _golden
    if _golden is None:
        with open(GOLDEN) as f:
            _golden = json.load(f)
    return _golden


def tensor_from_desc(d) -> tfs.TensorProto:
    t = tfs.TensorProto(dtype=d["dtype"], version_number=d["version_number"])
    if d["shape"] is not None:
        t.tensor_shape = tfs.TensorShapeProto(list(d["shape"]), d["unknown_rank"])
    for k, hx in d["fields"].items():
        if k in RAGGED_FIELDS:
            setattr(t, k, [bytes.fromhex(x) for x in hx])
            continue
        a = np.frombuffer(bytes.fromhex(hx), dtype=FIELD_NP[k]).copy()
        if k == "bool_val":
            a = a.astype(np.bool_)
        setattr(t, k, a)
    return t


def message_from_desc(kind: str, m):
    if kind == "tensor":
        return tensor_from_desc(m)
    ms = None
    if m["model_spec"] is not None:
        s = m["model_spec"]
        ms = tfs.ModelSpec(s["name"], s.get("version"), s.get("version_label"), s["signature_name"])
    entries = [(k, tensor_from_desc(td)) for k, td in m["entries"]]
    if kind == "request":
        return tfs.PredictRequest(ms, entries, list(m.get("output_filter", [])))
    return tfs.PredictResponse(entries, ms)


def tensor_to_desc(t: tfs.TensorProto):
    """Canonical dict of a decoded TensorProto: raw little-endian hex per non-empty field."""
    d = {"dtype": int(t.dtype), "shape": None, "unknown_rank": False, "version_number": int(t.version_number), "fields": {}}
    if t.tensor_shape is not None:
        d["shape"] = [int(x) for x in t.tensor_shape.dim]
        d["unknown_rank"] = bool(t.tensor_shape.unknown_rank)
    for kind, v in t.payloads():
        name = abi.KIND_FIELD_NAMES[kind]
        if name in RAGGED_FIELDS:
            if len(v):
                d["fields"][name] = [bytes(x).hex() for x in v]
            continue
        a = np.ascontiguousarray(v)
        if a.size:
            d["fields"][name] = a.astype(FIELD_NP[name], copy=False).tobytes().hex() if a.dtype.kind != "b" else a.astype("u1").tobytes().hex()
    return d


def message_to_desc(kind: str, msg, last_wins=True):
    if kind == "tensor":
        return tensor_to_desc(msg)
    entries = msg.inputs if kind == "request" else msg.outputs
    if last_wins:  # HashMap::insert semantics, keys sorted like the golden generator does
        dd = {}
        for k, t in entries:
            dd[k] = t
        ent = [[k, tensor_to_desc(dd[k])] for k in sorted(dd)]
    else:
        ent = [[k, tensor_to_desc(t)] for k, t in entries]
    ms = None
    if msg.model_spec is not None:
        s = msg.model_spec
        ms = {"name": s.name, "signature_name": s.signature_name, "version": s.version, "version_label": s.version_label}
    out = {"model_spec": ms, "entries": ent}
    if kind == "request":
        out["output_filter"] = list(msg.output_filter)
    return out


def norm_spec(d):
    """Golden encode descriptions omit version_label; normalise for comparison."""
    if isinstance(d, dict) and d.get("model_spec") is not None:
        d = dict(d)
        ms = dict(d["model_spec"])
        ms.setdefault("version_label", None)
        ms.setdefault("version", None)
        d["model_spec"] = ms
    return d


def assert_same_decoded(a, b, kind: str, ctx=""):
    """a, b: lists of DecodedMessage (e.g. oracle vs GPU).  Entries are compared in wire order."""
    assert len(a) == len(b), f"{ctx}: message count {len(a)} != {len(b)}"
    for i, (x, y) in enumerate(zip(a, b)):
        assert x.status == y.status, f"{ctx}: msg {i} status {x.status} != {y.status} (offsets {x.err_offset}, {y.err_offset})"
        if x.status != 0:
            continue
        dx, dy = message_to_desc(kind, x.message, last_wins=False), message_to_desc(kind, y.message, last_wins=False)
        if dx != dy:
            raise AssertionError(f"{ctx}: msg {i} differs:\n{_first_diff(dx, dy)}")


def _first_diff(dx, dy, path=""):
    if type(dx) is not type(dy):
        return f"{path}: type {type(dx)} vs {type(dy)}"
    if isinstance(dx, dict):
        for k in sorted(set(dx) | set(dy)):
            if dx.get(k) != dy.get(k):
                return _first_diff(dx.get(k), dy.get(k), f"{path}.{k}")
    if isinstance(dx, list):
        if len(dx) != len(dy):
            return f"{path}: len {len(dx)} vs {len(dy)}"
        for i, (p, q) in enumerate(zip(dx, dy)):
            if p != q:
                return _first_diff(p, q, f"{path}[{i}]")
    if isinstance(dx, str) and len(dx) > 64:
        n = next((i for i in range(min(len(dx), len(dy))) if dx[i] != dy[i]), min(len(dx), len(dy)))
        return f"{path}: hex differs at nibble {n} (byte {n // 2}) of {len(dx) // 2}/{len(dy) // 2}: ...{dx[max(0, n - 16): n + 32]} vs ...{dy[max(0, n - 16): n + 32]}"
    return f"{path}: {dx!r} vs {dy!r}"


def assert_same_decoded_fast(a, b, kind: str, ctx=""):
    """Same contract as assert_same_decoded for batches too large for hex descriptions (hundreds of MB): statuses,
    model specs, keys, dtypes, shapes, and every payload compared as raw little-endian bytes."""
    assert len(a) == len(b), f"{ctx}: message count {len(a)} != {len(b)}"
    for i, (x, y) in enumerate(zip(a, b)):
        assert x.status == y.status, f"{ctx}: msg {i} status {x.status} != {y.status}"
        if x.status != 0:
            continue
        mx, my = x.message, y.message
        if kind == "tensor":
            ex, ey = [("", mx)], [("", my)]
        else:
            assert mx.model_spec == my.model_spec, f"{ctx}: msg {i} model_spec {mx.model_spec} != {my.model_spec}"
            ex, ey = (mx.inputs, my.inputs) if kind == "request" else (mx.outputs, my.outputs)
            if kind == "request":
                assert list(mx.output_filter) == list(my.output_filter), f"{ctx}: msg {i} output_filter"
        assert len(ex) == len(ey), f"{ctx}: msg {i} has {len(ex)} vs {len(ey)} entries"
        for j, ((kx, tx), (ky, ty)) in enumerate(zip(ex, ey)):
            where = f"{ctx}: msg {i} entry {j} ({kx!r})"
            assert kx == ky and tx.dtype == ty.dtype and tx.version_number == ty.version_number and tx.tensor_shape == ty.tensor_shape, where
            px, py = dict(tx.payloads()), dict(ty.payloads())
            px = {k: v for k, v in px.items() if len(v)}
            py = {k: v for k, v in py.items() if len(v)}
            assert sorted(px) == sorted(py), f"{where}: fields {sorted(px)} vs {sorted(py)}"
            for k in px:
                name = abi.KIND_FIELD_NAMES[k]
                if name in RAGGED_FIELDS:
                    assert [bytes(e) for e in px[k]] == [bytes(e) for e in py[k]], f"{where}: {name}"
                    continue
                ax = np.ascontiguousarray(px[k]).astype(FIELD_NP[name], copy=False) if np.asarray(px[k]).dtype.kind != "b" else np.asarray(px[k]).astype("u1")
                ay = np.ascontiguousarray(py[k]).astype(FIELD_NP[name], copy=False) if np.asarray(py[k]).dtype.kind != "b" else np.asarray(py[k]).astype("u1")
                if ax.size != ay.size or ax.tobytes() != ay.tobytes():
                    bad = -1
                    if ax.size == ay.size:
                        d = np.nonzero(ax.view("u1") != ay.view("u1"))[0]
                        bad = int(d[0]) if d.size else -1
                    raise AssertionError(f"{where}: {name} differs ({ax.size} vs {ay.size} elements, first differing byte {bad})")


def mutation_corpus(kind: str, n_cases: int, seed: int = 99):
    """Byte-level mutations (truncation, byte replacement, insertion, a bit flip near the front) of valid
    PredictResponses / PredictRequests: the corpus the GPU fuzz test and the upb accept / reject test share."""
    from tfs_b200 import workloads
    import oracle_bind as ob
    rng = np.random.default_rng(seed)
    resp = workloads.c5_responses(24, seed=31, max_log2=7)
    if kind == "response":
        base = ob.encode_messages(resp, abi.MSG_PREDICT_RESPONSE)
    else:
        reqs = [tfs.PredictRequest(r.model_spec, r.outputs, ["a", "filter-%d" % i][: i % 3]) for i, r in enumerate(resp)]
        base = ob.encode_messages(reqs, abi.MSG_PREDICT_REQUEST)
    blobs = []
    for i in range(n_cases):
        b = bytearray(base[i % len(base)])
        op = rng.integers(0, 4)
        if op == 0 and len(b) > 1:
            b = b[: rng.integers(0, len(b))]
        elif op == 1 and len(b):
            for _ in range(int(rng.integers(1, 4))):
                b[rng.integers(0, len(b))] = rng.integers(0, 256)
        elif op == 2 and len(b):
            j = rng.integers(0, len(b))
            b[j:j] = bytes(rng.integers(0, 256, rng.integers(1, 5), dtype=np.uint8))
        elif len(b) > 2:
            j = rng.integers(0, min(len(b), 40))
            b[j] ^= 1 << int(rng.integers(0, 8))
        blobs.append(bytes(b))
    return blobs
