"""ctypes mirror of include/tfs_gpu.h (constants + plain-data structs).

This is the same picture of the ABI the Rust shim's `#[repr(C)]` structs give (rust/ffi.rs); the
Python harness (tests, bench.py) drives the C entry points exactly as `GpuCodec` in the reference's
src/lib.rs would.
"""
import ctypes as C

TFS_ABI_VERSION = 3
TFS_MAX_DIMS = 8
TFS_N_SLOTS = 14

# status codes
TFS_OK = 0
TFS_ERR_INVALID_VARINT = 1
TFS_ERR_BUFFER_UNDERFLOW = 2
TFS_ERR_LENGTH_EXCEEDED = 3
TFS_ERR_WIRE_TYPE = 4
TFS_ERR_WIRE_TYPE_VALUE = 5
TFS_ERR_INVALID_KEY = 6
TFS_ERR_INVALID_TAG = 7
TFS_ERR_RECURSION = 8
TFS_ERR_END_GROUP = 9
TFS_ERR_UTF8 = 10
TFS_ERR_GRPC_FLAG = 11
TFS_ERR_UNSUPPORTED = 12
TFS_ERR_INVALID_ARG = 20
TFS_ERR_NOMEM = 21
TFS_ERR_CUDA = 22
TFS_ERR_BATCH = 23

STATUS_NAMES = {
    0: "ok", 1: "invalid varint", 2: "buffer underflow", 3: "delimited length exceeded",
    4: "invalid wire type", 5: "invalid wire type value", 6: "invalid key value", 7: "invalid tag value: 0",
    8: "recursion limit reached", 9: "unexpected end group tag",
    10: "invalid string value: data is not UTF-8 encoded", 11: "grpc compressed flag", 12: "unsupported",
    20: "invalid argument", 21: "out of memory / capacity", 22: "cuda error", 23: "some messages failed",
}

# payload kinds (= TensorProto field numbers, tensor.proto:36-83)
KIND_NONE = 0
KIND_TENSOR_CONTENT = 4
KIND_FLOAT_VAL = 5
KIND_DOUBLE_VAL = 6
KIND_INT_VAL = 7
KIND_STRING_VAL = 8
KIND_SCOMPLEX_VAL = 9
KIND_INT64_VAL = 10
KIND_BOOL_VAL = 11
KIND_DCOMPLEX_VAL = 12
KIND_HALF_VAL = 13
KIND_RESOURCE_HANDLE_VAL = 14
KIND_VARIANT_VAL = 15
KIND_UINT32_VAL = 16
KIND_UINT64_VAL = 17
SLOT_KINDS = (4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 16, 17, 14, 15)
RAGGED_KINDS = (8, 14, 15)
KIND_FIELD_NAMES = {
    4: "tensor_content", 5: "float_val", 6: "double_val", 7: "int_val", 8: "string_val", 9: "scomplex_val",
    10: "int64_val", 11: "bool_val", 12: "dcomplex_val", 13: "half_val", 16: "uint32_val", 17: "uint64_val",
    14: "resource_handle_val", 15: "variant_val",
}

# element types
T_NONE, T_U8, T_I8, T_U16, T_I16, T_I32, T_U32, T_I64, T_U64, T_F16, T_BF16, T_F32, T_F64, T_BOOL, T_RAGGED = range(15)
TYPE_SIZE = {T_NONE: 0, T_U8: 1, T_I8: 1, T_U16: 2, T_I16: 2, T_I32: 4, T_U32: 4, T_I64: 8, T_U64: 8,
             T_F16: 2, T_BF16: 2, T_F32: 4, T_F64: 8, T_BOOL: 1}
# numpy view dtype of each element type (16-bit floats are carried as their bit patterns)
TYPE_NUMPY = {T_U8: "u1", T_I8: "i1", T_U16: "<u2", T_I16: "<i2", T_I32: "<i4", T_U32: "<u4", T_I64: "<i8",
              T_U64: "<u8", T_F16: "<u2", T_BF16: "<u2", T_F32: "<f4", T_F64: "<f8", T_BOOL: "u1"}

OP_NONE, OP_DIV, OP_MUL = 0, 1, 2
MSG_TENSOR_PROTO, MSG_PREDICT_REQUEST, MSG_PREDICT_RESPONSE, MSG_INPUT, MSG_CLASSIFY_REQUEST = 0, 1, 2, 3, 4
MSG_METADATA_REQUEST, MSG_METADATA_RESPONSE = 5, 6  # get_model_metadata.proto:16-35 (the response is decode only)
KIND_BYTES_LIST, KIND_FLOAT_LIST, KIND_INT64_LIST = 1, 2, 3  # Feature.kind members (feature.proto:79-81)
VERSION_NONE, VERSION_NUMBER, VERSION_LABEL = 0, 1, 2

F_GRPC_PREFIX = 1 << 0
F_ALIGN_PAYLOAD = 1 << 1
F_HALF_TO_F32 = 1 << 2
F_CONTENT_HALF_TO_F32 = 1 << 3
F_ASYNC = 1 << 4  # tfs_encode: return once the offsets are known; dst completes in stream order

# DataType values used by the path (types.proto:12-68)
DT_FLOAT, DT_DOUBLE, DT_INT32, DT_UINT8, DT_INT16, DT_INT8, DT_STRING, DT_COMPLEX64, DT_INT64, DT_BOOL = range(1, 11)
DT_BFLOAT16, DT_UINT16, DT_COMPLEX128, DT_HALF, DT_UINT32, DT_UINT64 = 14, 17, 18, 19, 22, 23


class Span(C.Structure):
    _fields_ = [("off", C.c_uint64), ("len", C.c_uint32), ("_pad", C.c_uint32)]


class Str(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("len", C.c_uint32), ("_pad", C.c_uint32)]


class FieldDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("src_type", C.c_int32), ("dst_type", C.c_int32), ("op", C.c_int32),
                ("op_param", C.c_float), ("_pad", C.c_uint32), ("n_elems", C.c_uint64), ("data", C.c_void_p),
                ("aux", C.c_uint64)]


class TensorDesc(C.Structure):
    _fields_ = [("name", Str), ("dtype", C.c_int32), ("version_number", C.c_int32), ("has_shape", C.c_int32),
                ("unknown_rank", C.c_int32), ("n_dims", C.c_uint32), ("first_field", C.c_uint32),
                ("n_fields", C.c_uint32), ("_pad", C.c_uint32), ("dims", C.c_int64 * TFS_MAX_DIMS)]


class MsgDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("has_model_spec", C.c_int32), ("model_name", Str), ("signature_name", Str),
                ("version_label", Str), ("version_choice", C.c_int32), ("_pad", C.c_uint32), ("version", C.c_int64),
                ("first_tensor", C.c_uint32), ("n_tensors", C.c_uint32), ("first_filter", C.c_uint32),
                ("n_filters", C.c_uint32)]


class Batch(C.Structure):
    _fields_ = [("msgs", C.POINTER(MsgDesc)), ("tensors", C.POINTER(TensorDesc)), ("fields", C.POINTER(FieldDesc)),
                ("filters", C.POINTER(Str)), ("n_msgs", C.c_uint32), ("n_tensors", C.c_uint32),
                ("n_fields", C.c_uint32), ("n_filters", C.c_uint32)]


class TensorOut(C.Structure):
    _fields_ = [("key", Span), ("msg_index", C.c_uint32), ("dtype", C.c_int32), ("version_number", C.c_int32),
                ("has_shape", C.c_int32), ("unknown_rank", C.c_int32), ("n_dims", C.c_uint32),
                ("dims", C.c_int64 * TFS_MAX_DIMS), ("n_elems", C.c_uint64 * TFS_N_SLOTS),
                ("dst_off", C.c_uint64 * TFS_N_SLOTS), ("dst_type", C.c_int32 * TFS_N_SLOTS)]


class MsgOut(C.Structure):
    _fields_ = [("status", C.c_int32), ("has_model_spec", C.c_int32), ("err_offset", C.c_uint64),
                ("model_name", Span), ("signature_name", Span), ("version_label", Span),
                ("version_choice", C.c_int32), ("_pad", C.c_uint32), ("version", C.c_int64),
                ("first_tensor", C.c_uint32), ("n_tensors", C.c_uint32), ("first_filter", C.c_uint32),
                ("n_filters", C.c_uint32)]


class DecodedView(C.Structure):
    _fields_ = [("msgs", C.POINTER(MsgOut)), ("tensors", C.POINTER(TensorOut)), ("filters", C.POINTER(Span)),
                ("strings", C.POINTER(C.c_uint8)), ("strings_bytes", C.c_uint64), ("pool_bytes", C.c_uint64),
                ("n_msgs", C.c_uint32), ("n_tensors", C.c_uint32), ("n_filters", C.c_uint32),
                ("n_failed", C.c_uint32)]


# Every symbol include/tfs_gpu.h declares (checked by the not-gpu ABI test).
EXPORTED_SYMBOLS = (
    "tfs_abi_version", "tfs_strerror", "tfs_ctx_create", "tfs_ctx_destroy", "tfs_last_error_detail",
    "tfs_ctx_set_stream", "tfs_sync", "tfs_launch_count", "tfs_host_alloc", "tfs_host_free", "tfs_dev_alloc",
    "tfs_dev_free", "tfs_copy_h2d", "tfs_copy_d2h", "tfs_encode_size", "tfs_encode_bound", "tfs_encode",
    "tfs_encode_host", "tfs_decode_scan", "tfs_decoded_view", "tfs_decode", "tfs_decoded_free",
    "tfs_decode_host", "tfs_decode_host_finish", "tfs_shard_plan", "tfs_slot_of_kind", "tfs_kind_of_slot",
    "tfs_type_size", "tfs_prof_enable", "tfs_prof_read", "tfs_pack_int64_noalloc", "tfs_decode_into", "tfs_decode_host_into",
    "tfs_last_error_location",
    "tfs_stage_acquire", "tfs_stage_release", "tfs_stage_stats", "tfs_engine_create", "tfs_engine_destroy", "tfs_engine_lanes",
    "tfs_engine_ctx", "tfs_engine_encode_host", "tfs_engine_decode_host_into", "tfs_engine_wait", "tfs_engine_decoded_free",
    "tfs_engine_decode_host_sharded", "tfs_shard_ranges", "tfs_coalescer_create", "tfs_coalescer_destroy", "tfs_coalescer_encode",
    "tfs_coalescer_release", "tfs_coalescer_stats",
)


class Slab(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("capacity", C.c_uint64), ("id", C.c_uint64)]
TFS_ERR_TOO_LARGE = 24
TFS_NOALLOC_MAX_ELEMS = 1 << 24
STATUS_NAMES[TFS_ERR_TOO_LARGE] = "input exceeds TFS_NOALLOC_MAX_ELEMS"

PK_COPY, PK_VARINT_SIZE, PK_VARINT_SCATTER, PK_VARINT_COUNT, PK_VARINT_GATHER, PK_COPY_TMA, PK_WALK, PK_FRAME = range(8)
PK_COPY_CAST = 8
PK_COUNT = 9
PK_NAMES = {PK_COPY: "k_copy", PK_VARINT_SIZE: "k_varint_size", PK_VARINT_SCATTER: "k_varint_scatter", PK_VARINT_COUNT: "k_varint_count",
            PK_VARINT_GATHER: "k_varint_gather", PK_COPY_TMA: "k_copy_tma", PK_WALK: "walk", PK_FRAME: "frame", PK_COPY_CAST: "k_copy_cast"}
