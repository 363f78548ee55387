"""Host-side mirror of the prost structs the reference sends and receives, plus the conversion
between them and the C ABI's plain-data descriptors.

Field names follow the generated Rust structs (`TensorProto { dtype, tensor_shape, float_val, .. }`,
`PredictRequest { model_spec, inputs, output_filter }`, reference src/lib.rs:249-263); map fields are
ordered lists of (key, TensorProto) because the ABI emits entries in the caller's order (prost's
HashMap order is unspecified).

Payload values are 1-D contiguous numpy arrays (host) or torch tensors (host or CUDA); nothing here
computes on them — only pointers and sizes go into the descriptors.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Any, List, Optional, Tuple

import numpy as np

from . import abi

try:  # torch is only plumbing (device memory, streams)
    import torch
except Exception:  # pragma: no cover
    torch = None


@dataclass
class Payload:
    """A payload array with an explicit in-memory type and (optionally) a fused conversion."""
    data: Any
    src_type: int
    dst_type: int = abi.T_NONE
    op: int = abi.OP_NONE
    op_param: float = 0.0


@dataclass
class TensorShapeProto:
    dim: List[int] = field(default_factory=list)  # Dim.size; Dim.name is always "" (src/lib.rs:233)
    unknown_rank: bool = False


@dataclass
class TensorProto:
    dtype: int = 0
    tensor_shape: Optional[TensorShapeProto] = None
    version_number: int = 0
    tensor_content: Any = None
    half_val: Any = None
    float_val: Any = None
    double_val: Any = None
    int_val: Any = None
    string_val: Any = None
    scomplex_val: Any = None
    int64_val: Any = None
    bool_val: Any = None
    dcomplex_val: Any = None
    uint32_val: Any = None
    uint64_val: Any = None
    # repeated ResourceHandleProto / VariantTensorDataProto (tensor.proto:72-76): a list whose element i is the ENCODED BODY of
    # message i (bytes-like; `resource_handle_body` / `variant_body` below build them), or a ready Ragged block
    resource_handle_val: Any = None
    variant_val: Any = None

    def payloads(self):
        """(kind, value) for every payload field that is set."""
        for kind in abi.SLOT_KINDS:
            v = getattr(self, abi.KIND_FIELD_NAMES[kind])
            if v is not None:
                yield kind, v


@dataclass
class ModelSpec:
    name: str = ""
    version: Optional[int] = None        # oneof version_choice: Int64Value version = 2
    version_label: Optional[str] = None  #                        string version_label = 4
    signature_name: str = ""


@dataclass
class AnyProto:
    """google.protobuf.Any as prost_types::Any holds it: the value stays opaque bytes."""
    type_url: str = ""
    value: bytes = b""


@dataclass
class GetModelMetadataRequest:
    """get_model_metadata.proto:16-23 (built by the reference at src/lib.rs:305-308)."""
    model_spec: Optional[ModelSpec] = None
    metadata_field: List[str] = field(default_factory=list)


@dataclass
class GetModelMetadataResponse:
    """get_model_metadata.proto:25-35; `metadata` in wire order (the reference keeps a HashMap: the last key wins)."""
    model_spec: Optional[ModelSpec] = None
    metadata: List[Tuple[str, AnyProto]] = field(default_factory=list)


@dataclass
class PredictRequest:
    model_spec: Optional[ModelSpec] = None
    inputs: List[Tuple[str, TensorProto]] = field(default_factory=list)
    output_filter: List[str] = field(default_factory=list)


@dataclass
class PredictResponse:
    outputs: List[Tuple[str, TensorProto]] = field(default_factory=list)
    model_spec: Optional[ModelSpec] = None


@dataclass
class Feature:
    """feature.proto:76-83: oneof kind { bytes_list = 1, float_list = 2, int64_list = 3 } (`Payload` at src/lib.rs:497-512).
    bytes_list: list of bytes-like (or a Ragged block); float_list: float32 array; int64_list: int64 array; all None =
    Feature { kind: None }."""
    bytes_list: Any = None
    float_list: Any = None
    int64_list: Any = None


@dataclass
class Input:
    """tensorflow.serving.Input with an ExampleList (input.proto:77-82), as build_input makes it (src/lib.rs:413-431):
    `examples` is a list of Examples, each an ordered list of (feature name, Feature)."""
    examples: List[List[Tuple[str, Feature]]] = field(default_factory=list)


@dataclass
class ClassificationRequest:
    """classification.proto:33-40 (RegressionRequest is field-for-field the same, regression.proto:23-30)."""
    model_spec: Optional[ModelSpec] = None
    input: Input = field(default_factory=Input)


_NP_TYPES = {
    "uint8": abi.T_U8, "int8": abi.T_I8, "uint16": abi.T_U16, "int16": abi.T_I16, "int32": abi.T_I32,
    "uint32": abi.T_U32, "int64": abi.T_I64, "uint64": abi.T_U64, "float16": abi.T_F16, "float32": abi.T_F32,
    "float64": abi.T_F64, "bool": abi.T_BOOL,
}


def _array_info(data):
    """(pointer, n_elems, TFS_T_*, keepalive) of a 1-D contiguous numpy array or torch tensor."""
    if isinstance(data, np.ndarray):
        if not data.flags["C_CONTIGUOUS"]:
            raise ValueError("payload arrays must be C-contiguous")
        t = _NP_TYPES.get(data.dtype.name)
        if t is None:
            raise TypeError(f"unsupported numpy dtype {data.dtype}")
        return (data.ctypes.data if data.size else 0), int(data.size), t, data
    if torch is not None and isinstance(data, torch.Tensor):
        if not data.is_contiguous():
            raise ValueError("payload tensors must be contiguous")
        name = str(data.dtype).replace("torch.", "")
        t = abi.T_BF16 if name == "bfloat16" else _NP_TYPES.get(name)
        if t is None:
            raise TypeError(f"unsupported torch dtype {data.dtype}")
        return (data.data_ptr() if data.numel() else 0), int(data.numel()), t, data
    raise TypeError(f"payload must be a numpy array or torch tensor, got {type(data)}")


class BuiltBatch:
    """Owns the ctypes arrays of one tfs_batch_t and everything they point at."""

    def __init__(self, msgs, tensors, fields, filters, keep):
        self.msgs, self.tensors, self.fields, self.filters = msgs, tensors, fields, filters
        self._keep = keep
        self.batch = abi.Batch(
            C.cast(msgs, C.POINTER(abi.MsgDesc)), C.cast(tensors, C.POINTER(abi.TensorDesc)),
            C.cast(fields, C.POINTER(abi.FieldDesc)), C.cast(filters, C.POINTER(abi.Str)),
            len(msgs), len(tensors), len(fields), len(filters))

    @property
    def n_msgs(self):
        return len(self.msgs)


def _mk_str(s, keep):
    b = s if isinstance(s, (bytes, bytearray)) else s.encode("utf-8")
    if not b:
        return abi.Str(None, 0, 0)
    buf = C.create_string_buffer(bytes(b), len(b))
    keep.append(buf)
    return abi.Str(C.cast(buf, C.c_void_p), len(b), 0)


def ragged_block(strings) -> np.ndarray:
    """`repeated bytes` (a list of bytes-like elements) flattened into the TFS_T_RAGGED block the ABI takes and the
    decoder returns: (n + 1) little-endian u64 offsets, then the bytes of all elements."""
    n = len(strings)
    offs = np.zeros(n + 1, dtype="<u8")
    if n:
        offs[1:] = np.cumsum([len(s) for s in strings], dtype=np.uint64)
    blob = np.empty(8 * (n + 1) + int(offs[n]), dtype=np.uint8)
    blob[: 8 * (n + 1)] = offs.view(np.uint8)
    blob[8 * (n + 1):] = np.frombuffer(b"".join(bytes(s) for s in strings), dtype=np.uint8)
    return blob


@dataclass
class Ragged:
    """A TFS_T_RAGGED block that already lives where the encoder needs it (host numpy array or torch tensor of uint8)."""
    block: Any
    n_elems: int
    total_bytes: int


def build_batch(messages, kind) -> BuiltBatch:
    """Turn a list of TensorProto / PredictRequest / PredictResponse into a tfs_batch_t."""
    keep: list = []
    msg_l, ten_l, fld_l, flt_l = [], [], [], []

    def add_tensor(name, t: TensorProto):
        td = abi.TensorDesc()
        td.name = _mk_str(name, keep)
        td.dtype = int(t.dtype)
        td.version_number = int(t.version_number)
        if t.tensor_shape is not None:
            td.has_shape = 1
            td.unknown_rank = 1 if t.tensor_shape.unknown_rank else 0
            td.n_dims = len(t.tensor_shape.dim)
            if td.n_dims > abi.TFS_MAX_DIMS:
                raise ValueError("rank exceeds TFS_MAX_DIMS")
            for i, d in enumerate(t.tensor_shape.dim):
                td.dims[i] = int(d)
        td.first_field = len(fld_l)
        n = 0
        for k, v in t.payloads():
            if k in abi.RAGGED_KINDS:  # repeated bytes / repeated messages: a list of bytes-like elements, or a ready block
                if not isinstance(v, Ragged):
                    blk = ragged_block(v)
                    v = Ragged(blk, len(v), blk.size - 8 * (len(v) + 1))
                ptr, _, ty, ka = _array_info(v.block)
                if ty != abi.T_U8:
                    raise TypeError("a ragged block is a uint8 array")
                keep.append(ka)
                fd = abi.FieldDesc()
                fd.kind, fd.src_type, fd.dst_type, fd.op = k, abi.T_RAGGED, abi.T_NONE, abi.OP_NONE
                fd.n_elems, fd.data, fd.aux = v.n_elems, ptr, v.total_bytes
                fld_l.append(fd)
                n += 1
                continue
            if isinstance(v, tuple):  # several arrays = ONE field on the wire, concatenated in order (batch assembly)
                parts = v
            else:
                parts = (v,)
            for part in parts:
                n += _add_field(fld_l, keep, k, part)
            continue
        td.n_fields = n
        ten_l.append(td)

    def _add_field(fld_l, keep, k, v):
        p = v if isinstance(v, Payload) else Payload(v, abi.T_NONE)
        ptr, count, ty, ka = _array_info(p.data)
        keep.append(ka)
        fd = abi.FieldDesc()
        fd.kind = k
        fd.src_type = p.src_type if p.src_type != abi.T_NONE else ty
        fd.dst_type = p.dst_type
        fd.op = p.op
        fd.op_param = float(p.op_param)
        # a source array viewed as a wider / narrower type than it is stored in
        fd.n_elems = count * abi.TYPE_SIZE[ty] // abi.TYPE_SIZE[fd.src_type]
        fd.data = ptr
        fld_l.append(fd)
        return 1

    def add_feature(name, ft: Feature, example_index: int):
        td = abi.TensorDesc()
        td.name = _mk_str(name, keep)
        td.version_number = example_index  # which Example of the list the feature belongs to
        td.first_field = len(fld_l)
        set_ = [(k, v) for k, v in ((abi.KIND_BYTES_LIST, ft.bytes_list), (abi.KIND_FLOAT_LIST, ft.float_list),
                                    (abi.KIND_INT64_LIST, ft.int64_list)) if v is not None]
        if len(set_) > 1:
            raise ValueError("Feature.kind is a oneof: set at most one list")
        for k, v in set_:
            fd = abi.FieldDesc()
            fd.kind, fd.dst_type, fd.op = k, abi.T_NONE, abi.OP_NONE
            if k == abi.KIND_BYTES_LIST:
                if not isinstance(v, Ragged):
                    blk = ragged_block(v)
                    v = Ragged(blk, len(v), blk.size - 8 * (len(v) + 1))
                ptr, _, ty, ka = _array_info(v.block)
                fd.src_type, fd.n_elems, fd.data, fd.aux = abi.T_RAGGED, v.n_elems, ptr, v.total_bytes
            else:
                ptr, count, ty, ka = _array_info(v)
                want = abi.T_F32 if k == abi.KIND_FLOAT_LIST else abi.T_I64
                if ty != want:
                    raise TypeError("float_list takes float32, int64_list takes int64")
                fd.src_type, fd.n_elems, fd.data = ty, count, ptr
            keep.append(ka)
            fld_l.append(fd)
        td.n_fields = len(set_)
        ten_l.append(td)

    for m in messages:
        md = abi.MsgDesc()
        md.kind = kind
        md.first_tensor = len(ten_l)
        md.first_filter = len(flt_l)
        if kind in (abi.MSG_INPUT, abi.MSG_CLASSIFY_REQUEST):
            inp = m if kind == abi.MSG_INPUT else m.input
            for ei, ex in enumerate(inp.examples):
                for name, ft in ex:
                    add_feature(name, ft, ei)
            if kind == abi.MSG_CLASSIFY_REQUEST and m.model_spec is not None:
                ms = m.model_spec
                md.has_model_spec = 1
                md.model_name = _mk_str(ms.name, keep)
                md.signature_name = _mk_str(ms.signature_name, keep)
                if ms.version is not None:
                    md.version_choice = abi.VERSION_NUMBER
                    md.version = int(ms.version)
                elif ms.version_label is not None:
                    md.version_choice = abi.VERSION_LABEL
                    md.version_label = _mk_str(ms.version_label, keep)
            md.n_tensors = len(ten_l) - md.first_tensor
            md.first_filter = 0
            md.n_filters = len(inp.examples)  # number of Examples (see TFS_MSG_INPUT in tfs_gpu.h)
            msg_l.append(md)
            continue
        if kind == abi.MSG_TENSOR_PROTO:
            add_tensor("", m)
        elif kind == abi.MSG_METADATA_RESPONSE:
            raise ValueError("GetModelMetadataResponse is decode only")
        else:
            entries = [] if kind == abi.MSG_METADATA_REQUEST else m.inputs if kind == abi.MSG_PREDICT_REQUEST else m.outputs
            for name, t in entries:
                add_tensor(name, t)
            if m.model_spec is not None:
                ms = m.model_spec
                md.has_model_spec = 1
                md.model_name = _mk_str(ms.name, keep)
                md.signature_name = _mk_str(ms.signature_name, keep)
                if ms.version is not None:
                    md.version_choice = abi.VERSION_NUMBER
                    md.version = int(ms.version)
                elif ms.version_label is not None:
                    md.version_choice = abi.VERSION_LABEL
                    md.version_label = _mk_str(ms.version_label, keep)
            if kind in (abi.MSG_PREDICT_REQUEST, abi.MSG_METADATA_REQUEST):
                for s in (m.output_filter if kind == abi.MSG_PREDICT_REQUEST else m.metadata_field):
                    flt_l.append(_mk_str(s, keep))
        md.n_tensors = len(ten_l) - md.first_tensor
        md.n_filters = len(flt_l) - md.first_filter
        msg_l.append(md)

    return BuiltBatch((abi.MsgDesc * len(msg_l))(*msg_l), (abi.TensorDesc * len(ten_l))(*ten_l),
                      (abi.FieldDesc * len(fld_l))(*fld_l), (abi.Str * len(flt_l))(*flt_l), keep)


# ---- the two message-typed TensorProto fields: element bodies are framed on the host (a handful of strings and scalars),
#      the TensorProtos INSIDE a variant are encoded / decoded by the library like any other tensor ----
def _varint(v: int) -> bytes:
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _len_field(tag: int, payload: bytes) -> bytes:
    return _varint((tag << 3) | 2) + _varint(len(payload)) + bytes(payload)


@dataclass
class ResourceHandleProto:
    """resource_handle.proto:13-30"""
    device: str = ""
    container: str = ""
    name: str = ""
    hash_code: int = 0
    maybe_type_name: str = ""


@dataclass
class VariantTensorDataProto:
    """tensor.proto:86-94; `tensors` are TensorProto objects (recursive)"""
    type_name: str = ""
    metadata: bytes = b""
    tensors: List[Any] = field(default_factory=list)


def resource_handle_body(h: ResourceHandleProto) -> bytes:
    """Encoded body of one resource_handle_val element: proto3 defaults skipped, ascending tags (prost's encode_raw)."""
    out = b""
    for tag, sv in ((1, h.device), (2, h.container), (3, h.name)):
        if sv:
            out += _len_field(tag, sv.encode("utf-8"))
    if h.hash_code:
        out += _varint((4 << 3) | 0) + _varint(h.hash_code)
    if h.maybe_type_name:
        out += _len_field(5, h.maybe_type_name.encode("utf-8"))
    return out


def variant_body(v: VariantTensorDataProto, encode_tensor) -> bytes:
    """Encoded body of one variant_val element.  `encode_tensor(TensorProto) -> bytes` encodes the tensors it holds (the
    codec itself: `lambda t: encoder.encode_host([t], abi.MSG_TENSOR_PROTO).all_bytes()[0]`)."""
    out = b""
    if v.type_name:
        out += _len_field(1, v.type_name.encode("utf-8"))
    if v.metadata:
        out += _len_field(2, bytes(v.metadata))
    for t in v.tensors:
        out += _len_field(3, encode_tensor(t))
    return out


def parse_resource_handle(body: bytes) -> ResourceHandleProto:
    """Opens one decoded resource_handle_val element (the library has validated it: no error handling needed here)."""
    h, i = ResourceHandleProto(), 0
    while i < len(body):
        key, i = _read_varint(body, i)
        tag, wt = key >> 3, key & 7
        if wt == 2:
            n, i = _read_varint(body, i)
            val = bytes(body[i: i + n])
            i += n
            if tag in (1, 2, 3, 5):
                setattr(h, {1: "device", 2: "container", 3: "name", 5: "maybe_type_name"}[tag], val.decode("utf-8"))
        elif wt == 0:
            val, i = _read_varint(body, i)
            if tag == 4:
                h.hash_code = val
        elif wt == 1:
            i += 8
        elif wt == 5:
            i += 4
        else:
            raise ValueError("groups inside a validated element are not expected here")
    return h


def parse_variant(body: bytes):
    """-> (type_name, metadata, [encoded TensorProto bytes, ...]) of one decoded variant_val element; the tensors are
    decoded by the library (TFS_MSG_TENSOR_PROTO)."""
    type_name, metadata, tensors, i = "", b"", [], 0
    while i < len(body):
        key, i = _read_varint(body, i)
        tag, wt = key >> 3, key & 7
        if wt == 2:
            n, i = _read_varint(body, i)
            val = bytes(body[i: i + n])
            i += n
            if tag == 1:
                type_name = val.decode("utf-8")
            elif tag == 2:
                metadata = val
            elif tag == 3:
                tensors.append(val)
        elif wt == 0:
            _, i = _read_varint(body, i)
        elif wt == 1:
            i += 8
        elif wt == 5:
            i += 4
        else:
            raise ValueError("groups inside a validated element are not expected here")
    return type_name, metadata, tensors


def _read_varint(b, i):
    v, shift = 0, 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7F) << shift
        shift += 7
        if c < 0x80:
            return v & ((1 << 64) - 1), i


def replicate_batch(built: BuiltBatch, n: int) -> BuiltBatch:
    """`n` copies of a ONE-message batch as one batch of n messages sharing the payload arrays and strings (the C1 shape:
    2^20 copies of a tiny request).  The descriptor arrays are replicated with numpy; a Python loop per message would
    take minutes at this size."""
    if built.n_msgs != 1:
        raise ValueError("replicate_batch takes a batch of exactly one message")
    nt, nf, nfl = len(built.tensors), len(built.fields), len(built.filters)

    def rep(arr, ctype, count):
        if count == 0:
            return (ctype * 0)(), None
        raw = np.frombuffer(bytes(arr), dtype=np.uint8).reshape(count, C.sizeof(ctype))
        out = (ctype * (count * n))()
        view = np.frombuffer(out, dtype=np.uint8).reshape(n, count, C.sizeof(ctype))
        view[:] = raw[None]
        return out, view

    msgs, mv = rep(built.msgs, abi.MsgDesc, 1)
    tens, tv = rep(built.tensors, abi.TensorDesc, nt)
    flds, _ = rep(built.fields, abi.FieldDesc, nf)
    flts, _ = rep(built.filters, abi.Str, nfl)
    idx = np.arange(n, dtype=np.uint32)

    def patch(view, ctype, field, stride, inner=None):
        off = getattr(ctype, field).offset
        base = np.frombuffer(bytes(view[0].tobytes()), dtype=np.uint8)  # (keeps the template's own values)
        for j in range(view.shape[1]):
            tmpl = int(np.frombuffer(view[0, j, off: off + 4].tobytes(), dtype="<u4")[0])
            vals = (tmpl + idx * stride).astype("<u4")
            view[:, j, off: off + 4] = vals.view(np.uint8).reshape(n, 4)
        del base

    patch(mv, abi.MsgDesc, "first_tensor", nt)
    patch(mv, abi.MsgDesc, "first_filter", nfl)
    if tv is not None:
        patch(tv, abi.TensorDesc, "first_field", nf)
    return BuiltBatch(msgs, tens, flds, flts, [built])


# ------------------------------------------------------------------------------------------------
# decode results -> message objects
# ------------------------------------------------------------------------------------------------

@dataclass
class DecodedMessage:
    status: int
    err_offset: int
    message: Any  # TensorProto / PredictRequest / PredictResponse, or None when status != 0


def _span_bytes(strings: bytes, sp) -> bytes:
    return strings[sp.off: sp.off + sp.len]


def _tensor_from_out(to, strings: bytes, pool: np.ndarray) -> Tuple[str, TensorProto]:
    t = TensorProto(dtype=to.dtype, version_number=to.version_number)
    if to.has_shape:
        t.tensor_shape = TensorShapeProto([int(to.dims[i]) for i in range(to.n_dims)], bool(to.unknown_rank))
    for slot, kind in enumerate(abi.SLOT_KINDS):
        n = int(to.n_elems[slot])
        if n == 0:
            continue
        ty = int(to.dst_type[slot])
        off = int(to.dst_off[slot])
        if kind in abi.RAGGED_KINDS:
            offs = pool[off: off + 8 * (n + 1)].view("<u8")
            base = off + 8 * (n + 1)
            val = [bytes(pool[base + int(offs[i]): base + int(offs[i + 1])]) for i in range(n)]
        elif kind == abi.KIND_TENSOR_CONTENT and ty == abi.T_F32:
            val = pool[off: off + (n // 2) * 4].view("<f4")  # widened 16-bit content
        else:
            val = pool[off: off + n * abi.TYPE_SIZE[ty]].view(abi.TYPE_NUMPY[ty])
        setattr(t, abi.KIND_FIELD_NAMES[kind], val)
    return _span_bytes(strings, to.key).decode("utf-8"), t


def messages_from_view(view: abi.DecodedView, pool: np.ndarray, kind: int) -> List[DecodedMessage]:
    """Rebuild message objects from a tfs_decoded_view_t and the (host copy of the) value pool."""
    strings = C.string_at(view.strings, view.strings_bytes) if view.strings_bytes else b""
    out = []
    for i in range(view.n_msgs):
        mo = view.msgs[i]
        if mo.status != abi.TFS_OK:
            out.append(DecodedMessage(mo.status, int(mo.err_offset), None))
            continue
        entries = [_tensor_from_out(view.tensors[mo.first_tensor + j], strings, pool) for j in range(mo.n_tensors)]
        ms = None
        if mo.has_model_spec:
            ms = ModelSpec(_span_bytes(strings, mo.model_name).decode("utf-8"),
                           int(mo.version) if mo.version_choice == abi.VERSION_NUMBER else None,
                           _span_bytes(strings, mo.version_label).decode("utf-8")
                           if mo.version_choice == abi.VERSION_LABEL else None,
                           _span_bytes(strings, mo.signature_name).decode("utf-8"))
        if kind == abi.MSG_TENSOR_PROTO:
            msg = entries[0][1] if entries else TensorProto()
        elif kind == abi.MSG_PREDICT_REQUEST:
            flt = [_span_bytes(strings, view.filters[mo.first_filter + j]).decode("utf-8") for j in range(mo.n_filters)]
            msg = PredictRequest(ms, entries, flt)
        elif kind == abi.MSG_METADATA_REQUEST:
            flt = [_span_bytes(strings, view.filters[mo.first_filter + j]).decode("utf-8") for j in range(mo.n_filters)]
            msg = GetModelMetadataRequest(ms, flt)
        elif kind == abi.MSG_METADATA_RESPONSE:
            # one record per map entry: key, string_val = [type_url, value]
            meta = []
            for key, t in entries:
                url, value = (list(t.string_val) + [b"", b""])[:2]
                meta.append((key, AnyProto(bytes(url).decode("utf-8"), bytes(value))))
            msg = GetModelMetadataResponse(ms, meta)
        else:
            msg = PredictResponse(entries, ms)
        out.append(DecodedMessage(0, 0, msg))
    return out
