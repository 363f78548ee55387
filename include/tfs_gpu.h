/*
 * tfs_gpu.h — C ABI of the B200-native TensorProto / PredictRequest / PredictResponse wire codec.
 *
 * This is the drop-in boundary for the ONE hot path of simonrw/rust-tensorflow-serving2: the prost
 * encode / decode that tonic's ProstCodec runs around `self.prediction_client.predict(request)`
 * (reference src/lib.rs:265; message built at src/lib.rs:229-263; wire contract in
 * protos/tensorflow/core/framework/tensor.proto:14-84, tensor_shape.proto:13-46, types.proto:12-68,
 * protos/tensorflow_serving/apis/predict.proto:12-40, model.proto:9-33).  The reference has no FFI of
 * its own; these are the symbols a `GpuCodec: tonic::codec::Codec` in src/lib.rs would bind through
 * `extern "C"` (see INTEGRATION.md for the Rust stub and rust/ for the shim sources).
 *
 * Conventions
 *   - plain pointers and sizes only; no C++/torch types; every function returns a tfs_status_t
 *     and never throws or aborts across the boundary.
 *   - one tfs_ctx_t per GPU (one process per GPU); a ctx is NOT re-entrant: serialise calls on one
 *     ctx, use several ctxs for several streams/threads.
 *   - "device pointer" = memory the GPU can address (cudaMalloc'd / torch CUDA tensor storage).
 *     "host pointer" = ordinary host memory; the *_host entry points stage it through the ctx's
 *     pinned slabs (host<->device copies are inside those calls).  Pageable buffers of 1 MB and more are copied
 *     through a pinned ring by helper threads the context starts on first use (at most 7, TFS_STAGE_THREADS in the
 *     environment overrides the count; they sleep between calls and end with tfs_ctx_destroy); a batch of about
 *     10^5 descriptors (messages + tensors + fields) and more is compiled on up to 8 short-lived threads
 *     (TFS_COMPILE_THREADS overrides the count).  TFS_TRACE=1 prints host phase times to stderr.
 *   - output bytes are byte-identical to prost 0.6.0's encoding of the same message with map entries
 *     in the caller's order; decoded values are value-identical.
 *   - over-read contract: the kernels fetch whole ALIGNED 16-byte words, so up to 15 bytes before the first and
 *     after the last byte of a source range (payload arrays on encode, the wire buffer on decode) may be READ; they are
 *     never written and never influence the output.  Any cudaMalloc'd / cudaHostAlloc'd / torch-allocated buffer
 *     satisfies this (allocations are at least 256-byte granular); a hand-made sub-allocation must not end closer than
 *     that to unmapped memory, and a neighbouring sub-allocation that another stream is writing is only ever read,
 *     and the bytes read from it are discarded.
 *   - there is no CPU fallback anywhere behind this header: without a usable CUDA device every
 *     compute entry point fails with TFS_ERR_CUDA.
 */
#ifndef TFS_GPU_H_
#define TFS_GPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TFS_ABI_VERSION 3
#define TFS_MAX_DIMS 8   /* dims carried inline per tensor; larger ranks -> TFS_ERR_UNSUPPORTED */
#define TFS_N_SLOTS 14   /* payload field slots of a TensorProto, see tfs_slot_of_kind() */

typedef int32_t tfs_status_t;
typedef struct tfs_ctx tfs_ctx_t;
typedef struct tfs_decoded tfs_decoded_t;

/* Status codes. The decode codes mirror prost 0.6.0 DecodeError descriptions so the
 * Rust shim can rebuild the same `Status::internal(..)` text tonic's ProstDecoder would produce. */
enum {
  TFS_OK = 0,
  TFS_ERR_INVALID_VARINT = 1,   /* "invalid varint" */
  TFS_ERR_BUFFER_UNDERFLOW = 2, /* "buffer underflow" */
  TFS_ERR_LENGTH_EXCEEDED = 3,  /* "delimited length exceeded" */
  TFS_ERR_WIRE_TYPE = 4,        /* "invalid wire type: X (expected Y)" */
  TFS_ERR_WIRE_TYPE_VALUE = 5,  /* "invalid wire type value: N" (6, 7) */
  TFS_ERR_INVALID_KEY = 6,      /* "invalid key value: N" (key > u32::MAX) */
  TFS_ERR_INVALID_TAG = 7,      /* "invalid tag value: 0" */
  TFS_ERR_RECURSION = 8,        /* "recursion limit reached" */
  TFS_ERR_END_GROUP = 9,        /* "unexpected end group tag" */
  TFS_ERR_UTF8 = 10,            /* "invalid string value: data is not UTF-8 encoded" */
  TFS_ERR_GRPC_FLAG = 11,       /* gRPC frame: compressed flag != 0 (tonic: Unimplemented / Internal) */
  TFS_ERR_UNSUPPORTED = 12,     /* valid wire, but uses something this library leaves to prost
                                   (rank > TFS_MAX_DIMS, dim names, a message of 4 GiB or more) */
  TFS_ERR_INVALID_ARG = 20,
  TFS_ERR_NOMEM = 21,           /* dst / pool capacity too small, or allocation failed */
  TFS_ERR_CUDA = 22,            /* CUDA runtime error; tfs_last_error_detail() has the text */
  TFS_ERR_BATCH = 23            /* at least one message of the batch failed; see per-message status */
};

/* TensorProto payload fields, numbered by their proto field number (tensor.proto:36-83). */
enum {
  TFS_KIND_NONE = 0,
  /* 1..3: the members of Feature.kind (feature.proto:79-81); only inside TFS_MSG_INPUT / TFS_MSG_CLASSIFY_REQUEST */
  TFS_KIND_BYTES_LIST = 1,     /* BytesList.value, repeated bytes: TFS_T_RAGGED block like string_val (aux = total bytes) */
  TFS_KIND_FLOAT_LIST = 2,     /* FloatList.value, packed fixed32 (src TFS_T_F32) */
  TFS_KIND_INT64_LIST = 3,     /* Int64List.value, packed varint (src TFS_T_I64) */
  TFS_KIND_TENSOR_CONTENT = 4, /* bytes */
  TFS_KIND_FLOAT_VAL = 5,      /* packed fixed32 */
  TFS_KIND_DOUBLE_VAL = 6,     /* packed fixed64 */
  TFS_KIND_INT_VAL = 7,        /* packed varint, int32 sign-extended */
  TFS_KIND_STRING_VAL = 8,     /* repeated bytes (not packed): one `42 len bytes` record per element; TFS_T_RAGGED */
  TFS_KIND_SCOMPLEX_VAL = 9,   /* packed fixed32 */
  TFS_KIND_INT64_VAL = 10,     /* packed varint */
  TFS_KIND_BOOL_VAL = 11,      /* packed varint 0/1 */
  TFS_KIND_DCOMPLEX_VAL = 12,  /* packed fixed64 */
  TFS_KIND_HALF_VAL = 13,      /* packed varint of the 16-bit pattern held in an int32 */
  /* The two message-typed fields (tensor.proto:72-76): `repeated ResourceHandleProto` / `repeated VariantTensorDataProto`.
   * Carried like string_val, as a TFS_T_RAGGED block whose element i is the ENCODED BODY of message i (what follows the
   * element's key and length on the wire).  Encode writes `72 len body` / `7a len body` per element; decode validates every
   * element exactly as prost's merge would - the nested TensorProtos of a variant, their nested variants, ..., down to
   * "recursion limit reached" at depth 100 - and returns the bodies; the host side opens them (a ResourceHandleProto is five
   * scalars; the tensors of a VariantTensorDataProto go through tfs_decode again as TFS_MSG_TENSOR_PROTO, see
   * include/tfs_codec.hpp).  Slots 12 and 13. */
  TFS_KIND_RESOURCE_HANDLE_VAL = 14,
  TFS_KIND_VARIANT_VAL = 15,
  TFS_KIND_UINT32_VAL = 16,    /* packed varint */
  TFS_KIND_UINT64_VAL = 17     /* packed varint */
};

/* In-memory element types (little-endian, dense). */
enum {
  TFS_T_NONE = 0,
  TFS_T_U8 = 1,
  TFS_T_I8 = 2,
  TFS_T_U16 = 3,
  TFS_T_I16 = 4,
  TFS_T_I32 = 5,
  TFS_T_U32 = 6,
  TFS_T_I64 = 7,
  TFS_T_U64 = 8,
  TFS_T_F16 = 9,   /* IEEE binary16 bit pattern */
  TFS_T_BF16 = 10, /* bfloat16 bit pattern */
  TFS_T_F32 = 11,
  TFS_T_F64 = 12,
  TFS_T_BOOL = 13, /* one byte per element, non-zero = true (Rust `bool`) */
  TFS_T_RAGGED = 14 /* `repeated bytes` (string_val, Rust Vec<Vec<u8>>) flattened into ONE block:
                       (n_elems + 1) little-endian u64 offsets, then the bytes of all elements back to back;
                       element i is bytes[offsets[i] .. offsets[i+1]), offsets[0] = 0. The decoder writes the
                       same block into the value pool, so a decoded string_val can be encoded again as is. */
};

/* Arithmetic fused into the u8 -> f32 widening of float_val (reference src/lib.rs:237-242:
 * `*p as f32` followed by the user's preprocessing_fn; examples/prediction.rs:43 uses `v / 255.`). */
enum {
  TFS_OP_NONE = 0,
  TFS_OP_DIV = 1, /* y = (float)x / param   (IEEE round-to-nearest division) */
  TFS_OP_MUL = 2  /* y = (float)x * param */
};

enum {
  TFS_MSG_TENSOR_PROTO = 0,     /* bare tensorflow.TensorProto */
  TFS_MSG_PREDICT_REQUEST = 1,  /* predict.proto:12-31 */
  TFS_MSG_PREDICT_RESPONSE = 2, /* predict.proto:34-40 */
  /* Encode only (what the reference's build_input returns, src/lib.rs:413-431, and the request that carries it):
   * tensorflow.serving.Input { example_list { examples: [Example { features { feature: map<string, Feature> } }] } }.
   * Described with the same arrays: every FEATURE is one tfs_tensor_desc_t (name = map key; dtype / shape / version
   * unused; 0 or 1 field of kind TFS_KIND_*_LIST, 0 = Feature { kind: None }), `version_number` = index of the Example
   * it belongs to (ascending inside the message), and tfs_msg_desc_t.n_filters = number of Examples in the list
   * (first_filter unused; an Example without features is an empty map). */
  TFS_MSG_INPUT = 3,            /* input.proto:77-82 */
  TFS_MSG_CLASSIFY_REQUEST = 4, /* classification.proto:33-40 = regression.proto:23-30: model_spec = 1, input = 2 */
  /* The metadata call of the reference (src/lib.rs:300-312), get_model_metadata.proto:16-35.
   * REQUEST (encode and decode): model_spec = 1, repeated string metadata_field = 2 - the message's `filters`; no tensors.
   * RESPONSE (decode only - the client never builds one): model_spec = 1, map<string, google.protobuf.Any> metadata = 2.
   * Every map entry comes back as ONE tensor record: key = the map key, dtype 0, no shape, and a string_val of exactly
   * two elements, [Any.type_url, Any.value].  prost keeps Any.value as opaque bytes (the SignatureDefMap inside is not
   * parsed by the reference either); type_url is a `string` and must be UTF-8; repeated occurrences merge as prost
   * merges them (the last type_url / value of an entry wins). */
  TFS_MSG_METADATA_REQUEST = 5,
  TFS_MSG_METADATA_RESPONSE = 6
};

enum { TFS_VERSION_NONE = 0, TFS_VERSION_NUMBER = 1, TFS_VERSION_LABEL = 2 }; /* model.proto:22-28 */

/* Encode / decode flags. */
enum {
  TFS_F_GRPC_PREFIX = 1u << 0,   /* encode: write tonic's 5-byte frame prefix (00 + u32 BE length) in
                                    front of every message; decode: each message starts with it */
  TFS_F_ALIGN_PAYLOAD = 1u << 1, /* encode: place each message so that its largest bulk payload starts
                                    on a 16-byte boundary of dst (messages are then not back-to-back) */
  TFS_F_HALF_TO_F32 = 1u << 2,   /* decode: half_val of a DT_HALF(19)/DT_BFLOAT16(14) tensor is widened
                                    to f32 instead of prost's Vec<i32> */
  TFS_F_CONTENT_HALF_TO_F32 = 1u << 3, /* decode: tensor_content of a DT_HALF/DT_BFLOAT16 tensor is
                                    widened to f32 */
  TFS_F_ASYNC = 1u << 4          /* tfs_encode (device destination) only: return as soon as the offsets and lengths of
                                    the messages are known, while the payload kernels are still writing dst.  dst is
                                    complete in stream order: for later work on the ctx's stream (a tfs_decode* of the
                                    same bytes, a copy queued on that stream), or after tfs_sync.  Without the flag
                                    tfs_encode returns with dst complete.  Ignored by the *_host entry points. */
};

typedef struct {
  uint64_t off; /* byte offset into the blob the span refers to */
  uint32_t len;
  uint32_t _pad;
} tfs_span_t;

typedef struct {
  const char* ptr; /* host pointer, not NUL-terminated */
  uint32_t len;
  uint32_t _pad;
} tfs_str_t;

/* ----------------------------------------------------------------------------------------------
 * Encode-side descriptors: a plain-data picture of the prost structs the reference builds at
 * src/lib.rs:249-263 (TensorProto, then the `inputs` map, then PredictRequest).
 * -------------------------------------------------------------------------------------------- */

/* One payload field of a TensorProto (`float_val: pixels` at src/lib.rs:252). */
typedef struct {
  int32_t kind;     /* TFS_KIND_* */
  int32_t src_type; /* TFS_T_* of the memory at `data` */
  int32_t dst_type; /* element type that goes on the wire; TFS_T_NONE = the field's natural type.
                       Allowed conversions: identical; U8->F32 (float_val, with op); F32->F16 / F32->BF16
                       (half_val or tensor_content, round-to-nearest-even, NaN kept quiet with sign);
                       F16/BF16/U16 -> half_val pattern; I8/U8/I16/U16 -> I32 (int_val). */
  int32_t op;       /* TFS_OP_* (only with U8->F32) */
  float op_param;
  uint32_t _pad;
  uint64_t n_elems; /* number of source elements (bytes for raw tensor_content with src_type U8) */
  const void* data; /* device pointer for tfs_encode, host pointer for tfs_encode_host.  Device arrays that are
                       varint-coded (int_val, int64_val, half_val, uint32/64_val, Int64List) must be aligned to their
                       element size, ragged blocks to 8 bytes; byte-copied fields and host arrays may sit anywhere */
  uint64_t aux;     /* TFS_KIND_STRING_VAL (src_type TFS_T_RAGGED): total bytes of all elements, i.e. offsets[n_elems]
                       (the host needs it for tfs_encode_bound and cannot read device memory); 0 otherwise */
} tfs_field_desc_t;

/* One TensorProto together with the map key it is stored under (src/lib.rs:256-257). */
typedef struct {
  tfs_str_t name;          /* map key; ignored for TFS_MSG_TENSOR_PROTO */
  int32_t dtype;           /* DataType enum value (types.proto:12-68); 0 is omitted like prost does */
  int32_t version_number;  /* tensor.proto:29 */
  int32_t has_shape;       /* Option<TensorShapeProto> is Some (always at src/lib.rs:251) */
  int32_t unknown_rank;    /* tensor_shape.proto:45 */
  uint32_t n_dims;         /* <= TFS_MAX_DIMS; Dim.name is always "" (src/lib.rs:233) */
  uint32_t first_field;    /* index into tfs_batch_t.fields */
  uint32_t n_fields;       /* payload fields of this tensor, any order.  A kind may appear more than once: the arrays are
                              then ONE field on the wire, concatenated in descriptor order (e.g. a [B,H,W,3] float_val
                              assembled from B separately stored u8 images, each with its own fused u8 -> f32 scale) */
  uint32_t _pad;
  int64_t dims[TFS_MAX_DIMS];
} tfs_tensor_desc_t;

/* One message of the batch (src/lib.rs:259-263; ModelSpec from build_model_spec, src/lib.rs:433-447). */
typedef struct {
  int32_t kind;            /* TFS_MSG_* */
  int32_t has_model_spec;  /* Option<ModelSpec> is Some */
  tfs_str_t model_name;
  tfs_str_t signature_name;
  tfs_str_t version_label;
  int32_t version_choice;  /* TFS_VERSION_* */
  uint32_t _pad;
  int64_t version;
  uint32_t first_tensor;   /* map entries in the order they are to be emitted */
  uint32_t n_tensors;
  uint32_t first_filter;   /* PredictRequest.output_filter, index into tfs_batch_t.filters */
  uint32_t n_filters;
} tfs_msg_desc_t;

typedef struct {
  const tfs_msg_desc_t* msgs;
  const tfs_tensor_desc_t* tensors;
  const tfs_field_desc_t* fields;
  const tfs_str_t* filters;
  uint32_t n_msgs;
  uint32_t n_tensors;
  uint32_t n_fields;
  uint32_t n_filters;
} tfs_batch_t;

/* ----------------------------------------------------------------------------------------------
 * Decode-side results: what prost's derived `merge` leaves in PredictResponse / PredictRequest /
 * TensorProto, as plain data.  Spans point into tfs_decoded_view_t.strings.
 * -------------------------------------------------------------------------------------------- */

typedef struct {
  tfs_span_t key;          /* map key ("" when the entry had none) */
  uint32_t msg_index;
  int32_t dtype;
  int32_t version_number;
  int32_t has_shape;
  int32_t unknown_rank;
  uint32_t n_dims;
  int64_t dims[TFS_MAX_DIMS];
  uint64_t n_elems[TFS_N_SLOTS]; /* elements per payload slot (bytes for tensor_content) */
  uint64_t dst_off[TFS_N_SLOTS]; /* byte offset of that slot's values inside the pool */
  int32_t dst_type[TFS_N_SLOTS]; /* TFS_T_* the values are written as */
} tfs_tensor_out_t;

typedef struct {
  tfs_status_t status;     /* TFS_OK or the decode error of this message */
  int32_t has_model_spec;
  uint64_t err_offset;     /* byte offset (within the message) where the error was detected */
  tfs_span_t model_name;
  tfs_span_t signature_name;
  tfs_span_t version_label;
  int32_t version_choice;
  uint32_t _pad;
  int64_t version;
  uint32_t first_tensor;   /* map entries in wire order; a repeated key means "last wins" */
  uint32_t n_tensors;
  uint32_t first_filter;
  uint32_t n_filters;
} tfs_msg_out_t;

typedef struct {
  const tfs_msg_out_t* msgs;
  const tfs_tensor_out_t* tensors;
  const tfs_span_t* filters;   /* output_filter strings of decoded PredictRequests */
  const uint8_t* strings;      /* blob the spans index */
  uint64_t strings_bytes;
  uint64_t pool_bytes;         /* size the dst pool of tfs_decode must have */
  uint32_t n_msgs;
  uint32_t n_tensors;
  uint32_t n_filters;
  uint32_t n_failed;           /* messages whose status != TFS_OK */
} tfs_decoded_view_t;

/* ----------------------------------------------------------------------------------------------
 * Context, memory and stream plumbing.  Replaces nothing in the reference: the client owns no
 * device today; a ctx is created where the channels are (src/lib.rs:117-145) and shared by clones
 * (src/lib.rs:148-156).
 * -------------------------------------------------------------------------------------------- */

int32_t tfs_abi_version(void);
const char* tfs_strerror(tfs_status_t status);

/* Create a context on CUDA device `device`.  `pinned_bytes` sizes the pinned staging slabs used by
 * the *_host entry points (0 = allocate lazily). */
tfs_status_t tfs_ctx_create(int32_t device, uint64_t pinned_bytes, tfs_ctx_t** out);
tfs_status_t tfs_ctx_destroy(tfs_ctx_t* ctx);
/* Text of the last failure on this ctx (CUDA error string, offending argument, ...). */
const char* tfs_last_error_detail(const tfs_ctx_t* ctx);
/* After a decode call returned TFS_ERR_BATCH: index of the FIRST failed message of that batch and the byte offset inside
 * it where the error was detected (what prost's DecodeError would point at; for an over-long varint inside a packed body
 * the offset is that of the 16-byte chunk holding it).  TFS_ERR_INVALID_ARG when the last call on this ctx had no failed
 * message.  Every message's own status / err_offset is in tfs_decoded_view_t.msgs[]. */
tfs_status_t tfs_last_error_location(const tfs_ctx_t* ctx, uint32_t* msg_index, uint64_t* byte_offset);
/* Run the ctx's work on `cuda_stream` (a cudaStream_t; NULL = the ctx's own non-blocking stream, which is NOT ordered
 * against the default stream).  The CUDA runtime's special handles are accepted as they are: cudaStreamLegacy
 * ((void*)1, the default stream torch and most libraries report as "stream 0") and cudaStreamPerThread ((void*)2). */
tfs_status_t tfs_ctx_set_stream(tfs_ctx_t* ctx, void* cuda_stream);
tfs_status_t tfs_sync(tfs_ctx_t* ctx);
/* Number of kernels this ctx has launched since creation (bench.py's gpu_launches). */
uint64_t tfs_launch_count(const tfs_ctx_t* ctx);

/* Per-kernel device timing (instrumentation; nothing in the reference corresponds to it).  While
 * enabled, every launch of the kernels below is bracketed by CUDA events on the ctx's stream;
 * tfs_prof_read synchronises the stream and returns the launches and summed milliseconds so far. */
enum {
  TFS_PK_COPY = 0,           /* K5 bulk copy / cast */
  TFS_PK_VARINT_SIZE = 1,    /* K1 */
  TFS_PK_VARINT_SCATTER = 2, /* K3 */
  TFS_PK_VARINT_COUNT = 3,   /* K4a */
  TFS_PK_VARINT_GATHER = 4,  /* K4c */
  TFS_PK_COPY_TMA = 5,       /* K5, TMA ring (source and destination congruent mod 16) */
  TFS_PK_WALK = 6,           /* K6d, small-batch decode plan (both structure walks) */
  TFS_PK_FRAME = 7,          /* K6a-c, small-batch encode framing */
  TFS_PK_COPY_CAST = 8       /* K5 / K7, staged widening / narrowing copies (u8 -> f32, f32 <-> f16 / bf16) */
};
tfs_status_t tfs_prof_enable(tfs_ctx_t* ctx, int32_t on);
tfs_status_t tfs_prof_read(tfs_ctx_t* ctx, int32_t kernel_id, uint64_t* launches_out, double* ms_out, int32_t reset);

/* Pinned host slabs: the bridge between tonic's BytesMut and HBM. */
tfs_status_t tfs_host_alloc(tfs_ctx_t* ctx, uint64_t bytes, void** out);
tfs_status_t tfs_host_free(tfs_ctx_t* ctx, void* p);
tfs_status_t tfs_dev_alloc(tfs_ctx_t* ctx, uint64_t bytes, void** out);
tfs_status_t tfs_dev_free(tfs_ctx_t* ctx, void* p);
tfs_status_t tfs_copy_h2d(tfs_ctx_t* ctx, void* dst_dev, const void* src_host, uint64_t bytes);
tfs_status_t tfs_copy_d2h(tfs_ctx_t* ctx, void* dst_host, const void* src_dev, uint64_t bytes);

/* ----------------------------------------------------------------------------------------------
 * Encode.  Replaces prost's `Message::encoded_len` + `Message::encode` as called by tonic's
 * ProstEncoder for the request built at src/lib.rs:259-263 and sent at src/lib.rs:265.
 * -------------------------------------------------------------------------------------------- */

/* Exact wire length of every message (prost `encoded_len`), without writing anything.
 * `msg_len_out` is a host array of n_msgs entries (the 5-byte prefix is NOT included). */
tfs_status_t tfs_encode_size(tfs_ctx_t* ctx, const tfs_batch_t* batch, uint64_t* msg_len_out);

/* Host-side upper bound of the bytes tfs_encode may need in dst for this batch and flags. */
uint64_t tfs_encode_bound(const tfs_batch_t* batch, uint32_t flags);

/* Encode the whole batch into device memory `dst_dev` (capacity dst_capacity bytes).
 * Field data pointers are device pointers.  On return (the call synchronises the stream):
 * msg_off_out[i] = byte offset in dst_dev of message i's first byte (of its gRPC prefix when
 * TFS_F_GRPC_PREFIX), msg_len_out[i] = its protobuf length (prefix excluded), *total_out = one past
 * the last byte written. */
tfs_status_t tfs_encode(tfs_ctx_t* ctx, const tfs_batch_t* batch, void* dst_dev, uint64_t dst_capacity,
                        uint32_t flags, uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out);

/* Same, end to end from host memory to host memory: field data are host pointers, dst_host is a host
 * buffer (pinned by tfs_host_alloc for full speed); host->device and device->host copies happen
 * inside the call. */
tfs_status_t tfs_encode_host(tfs_ctx_t* ctx, const tfs_batch_t* batch, void* dst_host, uint64_t dst_capacity,
                             uint32_t flags, uint64_t* msg_off_out, uint64_t* msg_len_out,
                             uint64_t* total_out);

/* ----------------------------------------------------------------------------------------------
 * Decode.  Replaces prost's `Message::decode` as called by tonic's ProstDecoder on the response
 * body of src/lib.rs:265-266.
 * -------------------------------------------------------------------------------------------- */

/* Pass 1: walk the structure of n_msgs messages of kind `msg_kind` stored in device memory
 * wire_dev[msg_off[i] .. msg_off[i]+msg_len[i]) (msg_off / msg_len are host arrays; with
 * TFS_F_GRPC_PREFIX each region starts with the 5-byte prefix and msg_len includes it), count the
 * elements of every payload field and lay out the value pool.  The result handle owns host arrays
 * described by tfs_decoded_view(). Returns TFS_ERR_BATCH when some messages are malformed (the others
 * are still usable). */
tfs_status_t tfs_decode_scan(tfs_ctx_t* ctx, const void* wire_dev, uint64_t wire_bytes, const uint64_t* msg_off,
                             const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags,
                             tfs_decoded_t** out);
tfs_status_t tfs_decoded_view(const tfs_decoded_t* dec, tfs_decoded_view_t* view);
/* Pass 2: write every payload value into the device pool (pool_bytes >= view.pool_bytes; 16-byte aligned, since it
 * holds typed arrays - every (tensor, field) region starts on a 64-byte multiple from the pool's base; an unaligned
 * device or pinned pool is TFS_ERR_INVALID_ARG, a pageable host pool may sit anywhere).  Synchronous: the values
 * are in the pool when the call returns (tfs_decode_into is the form whose value kernels run on behind the call). */
tfs_status_t tfs_decode(tfs_ctx_t* ctx, tfs_decoded_t* dec, void* pool_dev, uint64_t pool_bytes);
tfs_status_t tfs_decoded_free(tfs_ctx_t* ctx, tfs_decoded_t* dec);

/* Both passes in one call, for callers that own a value pool of known capacity (the shim's staging pool):
 * the values are written right behind the scan, without a round trip to the host in between.  When the pool
 * turns out to be too small nothing is written and TFS_ERR_NOMEM is returned; the handle is still valid and
 * its view says how many bytes are needed (then call tfs_decode with a larger pool).  Like tfs_decode the
 * value kernels may still be running when the call returns (the view is complete): tfs_sync() before the pool
 * is read from the host or from another stream. */
tfs_status_t tfs_decode_into(tfs_ctx_t* ctx, const void* wire_dev, uint64_t wire_bytes, const uint64_t* msg_off,
                             const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, void* pool_dev,
                             uint64_t pool_capacity, tfs_decoded_t** out);

/* End to end: wire bytes in host memory, values written to a host pool (both pinned for full speed).
 * The pool is allocated by the caller after looking at the view: call tfs_decode_host to get the handle and the
 * sizes, then tfs_decode_host_finish.  A pageable wire buffer is staged on the device by the first call; a pinned /
 * registered one is read IN PLACE by both calls, so it must stay alive and unmodified until tfs_decode_host_finish
 * has returned. */
tfs_status_t tfs_decode_host(tfs_ctx_t* ctx, const void* wire_host, uint64_t wire_bytes, const uint64_t* msg_off,
                             const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags,
                             tfs_decoded_t** out);
tfs_status_t tfs_decode_host_finish(tfs_ctx_t* ctx, tfs_decoded_t* dec, void* pool_host, uint64_t pool_bytes);
/* Host-memory form of tfs_decode_into.  Pinned buffers (tfs_host_alloc) are read and written in place by the
 * kernels; pageable ones are staged through device memory.  Synchronous: the values are in pool_host on return. */
tfs_status_t tfs_decode_host_into(tfs_ctx_t* ctx, const void* wire_host, uint64_t wire_bytes, const uint64_t* msg_off,
                                  const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, void* pool_host,
                                  uint64_t pool_capacity, tfs_decoded_t** out);

/* ----------------------------------------------------------------------------------------------
 * Pinned-host staging pool: the bridge between tonic's byte buffers and HBM.  Replaces nothing in the reference
 * (its buffers are tonic's BytesMut, grown by `reserve`, src/lib.rs:265 via ProstEncoder); it is created where
 * the reference creates its channels (src/lib.rs:132-138) and serves the codec at the call site (:265).
 *
 * `tfs_ctx_create(device, pinned_bytes, ..)` pins one region of that size ONCE; slabs are carved from it in
 * power-of-two classes (4 KiB up) and recycled, so acquiring one on the request path is a lock and a pop - no
 * cudaHostAlloc (which synchronises the device).  A request larger than what is left falls back to a one-off pinned
 * allocation that is freed again on release.  Thread-safe.
 *
 * `align_skew`: the returned pointer is congruent to it modulo 16.  A received message whose bulk payload starts
 * `h` bytes into it is placed at skew (16 - h % 16) % 16: the payload then sits on a 16-byte boundary and decodes
 * through the TMA ring instead of the byte-shifting copy (on encode TFS_F_ALIGN_PAYLOAD does the same inside dst).
 * -------------------------------------------------------------------------------------------- */
typedef struct {
  void* ptr;         /* pinned host memory, usable by the *_host entry points without staging */
  uint64_t capacity; /* bytes available at ptr */
  uint64_t id;       /* opaque: hand the struct back unchanged */
} tfs_slab_t;
tfs_status_t tfs_stage_acquire(tfs_ctx_t* ctx, uint64_t bytes, uint32_t align_skew, tfs_slab_t* out);
tfs_status_t tfs_stage_release(tfs_ctx_t* ctx, const tfs_slab_t* slab);
/* bytes of the pool currently handed out / its size / how many acquires had to fall back to a one-off allocation */
tfs_status_t tfs_stage_stats(tfs_ctx_t* ctx, uint64_t* in_use, uint64_t* pool_bytes, uint64_t* fallbacks);

/* ----------------------------------------------------------------------------------------------
 * Engine: worker threads over one or several GPUs, asynchronous submission, request coalescing.
 *
 * The reference's client is `Clone` and is used from concurrent tokio tasks (src/lib.rs:148-156,
 * examples/async.rs:29-46), every task sending ONE request per call (src/lib.rs:265).  A tfs_ctx_t is neither
 * re-entrant nor asynchronous; the engine puts `lanes_per_device` contexts on each listed GPU, each with its own
 * stream, staging pool and host worker thread.  A submitted job runs the ordinary host entry point
 * (tfs_encode_host / tfs_decode_host_into) on a lane's thread, so the upload of one job overlaps the download of
 * the previous one, and several GPUs work at once from ONE client process.  Requests are independent: no
 * collective, no peer access; every GPU touches only its own memory.
 *
 * All arrays handed to a submit call (descriptors, payloads, output arrays, destination) must stay valid until
 * tfs_engine_wait() has returned for its ticket.  Tickets may be waited for in any order, from any thread, once.
 * -------------------------------------------------------------------------------------------- */
typedef struct tfs_engine tfs_engine_t;
typedef uint64_t tfs_ticket_t;
tfs_status_t tfs_engine_create(const int32_t* devices, uint32_t n_devices, uint32_t lanes_per_device, uint64_t pinned_bytes_per_lane,
                               tfs_engine_t** out);
tfs_status_t tfs_engine_destroy(tfs_engine_t* eng);
uint32_t tfs_engine_lanes(const tfs_engine_t* eng);
/* The context of lane `lane` (for its staging pool: tfs_stage_acquire / release are thread-safe; do not call compute
 * entry points on it while jobs are in flight). */
tfs_ctx_t* tfs_engine_ctx(tfs_engine_t* eng, uint32_t lane);
/* Asynchronous forms of tfs_encode_host / tfs_decode_host_into.  `lane` = -1: the lane with the fewest bytes queued. */
tfs_status_t tfs_engine_encode_host(tfs_engine_t* eng, int32_t lane, const tfs_batch_t* batch, void* dst_host, uint64_t dst_capacity,
                                    uint32_t flags, uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out,
                                    tfs_ticket_t* ticket);
tfs_status_t tfs_engine_decode_host_into(tfs_engine_t* eng, int32_t lane, const void* wire_host, uint64_t wire_bytes,
                                         const uint64_t* msg_off, const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind,
                                         uint32_t flags, void* pool_host, uint64_t pool_capacity, tfs_decoded_t** out,
                                         tfs_ticket_t* ticket);
/* Blocks until the job is done and returns ITS status (what the synchronous entry point would have returned);
 * detail_out (optional, >= 256 bytes) receives the lane's error text.  A decode handle is freed with
 * tfs_engine_decoded_free (it belongs to the lane that made it). */
tfs_status_t tfs_engine_wait(tfs_engine_t* eng, tfs_ticket_t ticket, char* detail_out);
tfs_status_t tfs_engine_decoded_free(tfs_engine_t* eng, tfs_decoded_t* dec);
/* A batch of independent messages decoded on ALL lanes at once (BASELINE config 5: 8 192 responses over 8 GPUs).  The
 * messages are cut into tfs_engine_lanes() contiguous runs of about equal bytes (tfs_shard_ranges); run r is decoded by
 * lane r into pool_host[r] (capacity pool_capacity[r]) and described by out[r], whose message / tensor indices count
 * from the run's first message, first_msg_out[r].  Synchronous; returns the first non-OK status among the runs. */
tfs_status_t tfs_engine_decode_host_sharded(tfs_engine_t* eng, const void* wire_host, uint64_t wire_bytes, const uint64_t* msg_off,
                                            const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags,
                                            void* const* pool_host, const uint64_t* pool_capacity, tfs_decoded_t** out,
                                            uint32_t* first_msg_out);
/* Contiguous split of n_msgs messages into n_shards runs of about equal bytes: first_out[s] = index of run s's first
 * message (first_out[n_shards] = n_msgs).  Host only. */
tfs_status_t tfs_shard_ranges(const uint64_t* msg_bytes, uint32_t n_msgs, uint32_t n_shards, uint32_t* first_out);

/* Coalescer: callers that each hold ONE message (the reference's call site, src/lib.rs:265) are gathered into one
 * batch per window - every roofline figure of this library is per launch over a batch.  tfs_coalescer_encode blocks
 * until the batch its message joined has been encoded; the bytes are handed over IN PLACE: *wire_out points into a
 * pinned slab shared by the batch (what the transport sends from, no copy in between), valid until the caller
 * releases its reference with tfs_coalescer_release.  Thread-safe; a batch closes when `max_msgs` messages have
 * joined or `window_us` microseconds after its first one. */
typedef struct tfs_coalescer tfs_coalescer_t;
tfs_status_t tfs_coalescer_create(tfs_engine_t* eng, uint32_t window_us, uint32_t max_msgs, tfs_coalescer_t** out);
tfs_status_t tfs_coalescer_destroy(tfs_coalescer_t* co);
tfs_status_t tfs_coalescer_encode(tfs_coalescer_t* co, const tfs_batch_t* one_message, uint32_t flags, const uint8_t** wire_out,
                                  uint64_t* len_out, uint64_t* lease_out);
tfs_status_t tfs_coalescer_release(tfs_coalescer_t* co, uint64_t lease);
/* batches launched / messages encoded so far */
tfs_status_t tfs_coalescer_stats(tfs_coalescer_t* co, uint64_t* batches, uint64_t* messages);

/* ----------------------------------------------------------------------------------------------
 * Allocation-free packing of one int64 array (no counterpart in the reference: the reference never
 * touches a GPU, so it has no notion of a caller that must not allocate device memory).
 *
 * Writes the packed-varint body of an `int64_val` field (tensor.proto:64; the bytes that follow the
 * field's key and length) for src_dev[0 .. n_elems) to dst_dev, and the body's length in bytes to
 * *body_len_dev.  All three pointers are device pointers supplied by the caller; the work is queued on
 * `cuda_stream` (a cudaStream_t, NULL = the default stream) of the CURRENT device and the call returns
 * without waiting.  The bytes equal what tfs_encode produces for the same values.
 *
 * It exists for callers that cannot afford, or are not allowed, a cudaMalloc: a codec running inside
 * a request-serving poll loop (cudaMalloc synchronises the device and would stall every other request
 * in flight), or a client sharing its GPU with a TensorFlow Serving process that has already taken
 * nearly all device memory.  Its working tables live in a fixed block of device memory that belongs to
 * the library image, so:
 *   - n_elems may be at most TFS_NOALLOC_MAX_ELEMS; more returns TFS_ERR_TOO_LARGE and launches nothing;
 *   - all calls share that block: calls that may overlap on the GPU (independent streams, several host
 *     threads) race and must be ordered by the caller; calls queued on one stream are safe;
 *   - the first call made while a given device is current looks up the block's address (one synchronous
 *     runtime query): make that first call outside of any stream capture;
 *   - if the body is longer than dst_capacity nothing is written to dst_dev; *body_len_dev still
 *     receives the length that would be needed, so compare it with dst_capacity after synchronising.
 * -------------------------------------------------------------------------------------------- */
#define TFS_NOALLOC_MAX_ELEMS (1ull << 24)
enum { TFS_ERR_TOO_LARGE = 24 }; /* input larger than the fixed working block of tfs_pack_int64_noalloc */
tfs_status_t tfs_pack_int64_noalloc(const int64_t* src_dev, uint64_t n_elems, void* dst_dev, uint64_t dst_capacity,
                                    uint64_t* body_len_dev, void* cuda_stream);

/* ----------------------------------------------------------------------------------------------
 * Sharding helper (host only).  Requests are independent, so a batch is split across GPUs with no
 * collective: greedy largest-first balancing on bytes.  shard_out[i] in [0, n_shards).
 * -------------------------------------------------------------------------------------------- */
tfs_status_t tfs_shard_plan(const uint64_t* msg_bytes, uint32_t n_msgs, uint32_t n_shards, uint32_t* shard_out);

/* Slot index (0..TFS_N_SLOTS-1) of a payload kind, -1 if the kind has no slot. */
int32_t tfs_slot_of_kind(int32_t kind);
int32_t tfs_kind_of_slot(int32_t slot);
/* Size in bytes of a TFS_T_* element, 0 for TFS_T_NONE. */
uint32_t tfs_type_size(int32_t type);

#ifdef __cplusplus
}
#endif
#endif /* TFS_GPU_H_ */
