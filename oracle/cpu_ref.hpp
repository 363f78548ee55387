// cpu_ref.hpp — CPU ORACLE (test infrastructure, NOT product code).
//
// A plain C++17 restatement of what prost 0.6.0 + prost-derive 0.6.0 do for the messages the reference
// sends at src/lib.rs:265: TensorProto (protos/tensorflow/core/framework/tensor.proto:14-84),
// TensorShapeProto (tensor_shape.proto:13-46), ModelSpec (protos/tensorflow_serving/apis/model.proto:9-33),
// PredictRequest / PredictResponse (predict.proto:12-40).  prost / tonic / bytes are only PINNED in the
// reference's Cargo.lock (prost 0.6.0 Cargo.lock:831-832, prost-derive 0.6.0 :857-858, bytes 0.5.3
// :131-132, tonic 0.1.0 :1264-1265) and are not vendored, so their algorithm is restated here from the
// protobuf wire format and those crate versions' published behaviour, keeping
// prost's PASS STRUCTURE (no cached sizes: nested lengths are recomputed by every caller; one put per
// element; push-with-growth on decode) so that this file is also the timed CPU baseline.
//
// PARITY UNPINNED: the reference has no tests, golden vectors or fixtures for this path (SURVEY.md §4,
// §8c) and cannot be compiled here (no rustc/cargo/protoc).  This oracle is pinned instead against
// (1) the known-answer vectors of SURVEY.md Appendix B and (2) upb (python protobuf 6.33.6 with
// tensorboard's compiled TensorProto) as an independent implementation — tests/golden/.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use
// anything in oracle/.  The product (rust-tensorflow-serving2_b200/) never does.
#pragma once
#include <cstdint>
#include <cstring>
#include <optional>
#include <string>
#include <utility>
#include <vector>

namespace cpu_ref {

// ---- prost::DecodeError description classes; numeric values = tfs_gpu.h TFS_ERR_* ----
enum class Err : int32_t {
  Ok = 0,
  InvalidVarint = 1,
  BufferUnderflow = 2,
  LengthExceeded = 3,
  WireType = 4,
  WireTypeValue = 5,
  InvalidKey = 6,
  InvalidTag = 7,
  Recursion = 8,
  EndGroup = 9,
  Utf8 = 10,
  GrpcFlag = 11,
  Unsupported = 12,
};

struct DecodeError {
  Err code;
  uint64_t offset;  // bytes consumed from the message when the error was raised
};

// ---- bytes::Buf over a slice (bytes 0.5.3 `impl Buf for &[u8]`) ----
struct Buf {
  const uint8_t* base;
  const uint8_t* p;
  const uint8_t* end;
  Buf(const uint8_t* b, size_t n) : base(b), p(b), end(b + n) {}
  size_t remaining() const { return static_cast<size_t>(end - p); }
  bool has_remaining() const { return p != end; }
  uint64_t consumed() const { return static_cast<uint64_t>(p - base); }
  void advance(size_t n) { p += n; }
};

// ---- bytes::BufMut over a growable byte vector (BytesMut) ----
struct BufMut {
  std::vector<uint8_t>& v;
  explicit BufMut(std::vector<uint8_t>& vec) : v(vec) {}
  void put_u8(uint8_t b) { v.push_back(b); }
  void put_slice(const void* s, size_t n) {
    const uint8_t* c = static_cast<const uint8_t*>(s);
    v.insert(v.end(), c, c + n);
  }
};

enum WireType : uint32_t { Varint = 0, SixtyFourBit = 1, LengthDelimited = 2, StartGroup = 3, EndGroup = 4, ThirtyTwoBit = 5 };

// ---- generated structs (prost-build output shape for the .proto files above) ----
struct Dim {
  int64_t size = 0;
  std::string name;
  bool operator==(const Dim& o) const { return size == o.size && name == o.name; }
};
struct TensorShapeProto {
  std::vector<Dim> dim;
  bool unknown_rank = false;
  bool operator==(const TensorShapeProto& o) const { return dim == o.dim && unknown_rank == o.unknown_rank; }
};
struct TensorProto {
  int32_t dtype = 0;                                // tag 1 (enumeration stored as i32)
  std::optional<TensorShapeProto> tensor_shape;     // tag 2
  int32_t version_number = 0;                       // tag 3
  std::vector<uint8_t> tensor_content;              // tag 4
  std::vector<int32_t> half_val;                    // tag 13
  std::vector<float> float_val;                     // tag 5
  std::vector<double> double_val;                   // tag 6
  std::vector<int32_t> int_val;                     // tag 7
  std::vector<std::vector<uint8_t>> string_val;     // tag 8
  std::vector<float> scomplex_val;                  // tag 9
  std::vector<int64_t> int64_val;                   // tag 10
  std::vector<uint8_t> bool_val;                    // tag 11 (Vec<bool>)
  std::vector<double> dcomplex_val;                 // tag 12
  std::vector<uint32_t> uint32_val;                 // tag 16
  std::vector<uint64_t> uint64_val;                 // tag 17
  // tag 14 `repeated ResourceHandleProto resource_handle_val`, tag 15 `repeated VariantTensorDataProto variant_val`
  // (tensor.proto:72-76): kept as the ENCODED BODY of every element.  On decode each element has been merged into the real
  // struct first (ResourceHandleProto / VariantTensorDataProto below, recursion limit and all), so the bytes are exactly the
  // messages prost accepted; on encode they are written back as `key len body` per element.
  std::vector<std::vector<uint8_t>> resource_handle_val;
  std::vector<std::vector<uint8_t>> variant_val;
  bool is_default() const;
};
// resource_handle.proto:13-30
struct ResourceHandleProto {
  std::string device, container, name;  // tags 1, 2, 3
  uint64_t hash_code = 0;               // tag 4
  std::string maybe_type_name;          // tag 5
};
// tensor.proto:86-94
struct VariantTensorDataProto {
  std::string type_name;             // tag 1
  std::vector<uint8_t> metadata;     // tag 2
  std::vector<TensorProto> tensors;  // tag 3 (recursive)
};
void encode_raw(const ResourceHandleProto& m, BufMut& buf);
void encode_raw(const VariantTensorDataProto& m, BufMut& buf);
size_t encoded_len(const ResourceHandleProto& m);
size_t encoded_len(const VariantTensorDataProto& m);
struct ModelSpec {
  std::string name;  // tag 1
  // oneof version_choice { Int64Value version = 2 (prost maps the wrapper to i64); string version_label = 4 }
  int32_t version_choice = 0;  // 0 none, 1 version, 2 label
  int64_t version = 0;
  std::string version_label;
  std::string signature_name;  // tag 3
};
using TensorMap = std::vector<std::pair<std::string, TensorProto>>;  // HashMap in caller / wire order
struct PredictRequest {
  std::optional<ModelSpec> model_spec;     // tag 1
  TensorMap inputs;                        // tag 2
  std::vector<std::string> output_filter;  // tag 3
};
struct PredictResponse {
  TensorMap outputs;                    // tag 1
  std::optional<ModelSpec> model_spec;  // tag 2
};

// ---- the metadata call (src/lib.rs:300-312; get_model_metadata.proto:16-35, google/protobuf/any.proto) ----
struct Any {               // prost_types::Any
  std::string type_url;    // tag 1 (string: UTF-8 checked on decode)
  std::vector<uint8_t> value;  // tag 2 (bytes: opaque - the reference never unpacks the SignatureDefMap inside)
};
struct GetModelMetadataRequest {
  std::optional<ModelSpec> model_spec;      // tag 1
  std::vector<std::string> metadata_field;  // tag 2
};
struct GetModelMetadataResponse {
  std::optional<ModelSpec> model_spec;                 // tag 1
  std::vector<std::pair<std::string, Any>> metadata;   // tag 2: map<string, Any>, wire order kept
};
size_t encoded_len(const GetModelMetadataRequest& m);
void encode_raw(const GetModelMetadataRequest& m, BufMut& buf);
size_t encoded_len(const GetModelMetadataResponse& m);
void encode_raw(const GetModelMetadataResponse& m, BufMut& buf);
bool decode(const uint8_t* p, size_t n, GetModelMetadataRequest* out, DecodeError* err);
bool decode(const uint8_t* p, size_t n, GetModelMetadataResponse* out, DecodeError* err);

// ---- Example / Features / Input (feature.proto:65-87, example.proto:88-90, input.proto:15-82, classification.proto:33-40;
//      built by build_input + `impl From<Payload> for Feature`, src/lib.rs:413-431, 514-571) ----
struct Feature {
  // oneof kind { BytesList bytes_list = 1; FloatList float_list = 2; Int64List int64_list = 3 }
  int32_t kind = 0;  // 0 = None
  std::vector<std::vector<uint8_t>> bytes_list;  // BytesList.value (repeated bytes = 1)
  std::vector<float> float_list;                 // FloatList.value (packed = 1)
  std::vector<int64_t> int64_list;               // Int64List.value (packed = 1)
};
using FeatureMap = std::vector<std::pair<std::string, Feature>>;  // HashMap in caller order
struct Example {
  std::optional<FeatureMap> features;  // Option<Features>; build_input always sets Some (src/lib.rs:420)
};
struct Input {
  // oneof kind { ExampleList example_list = 1; ExampleListWithContext example_list_with_context = 2 }: only the first
  bool has_example_list = true;
  std::vector<Example> examples;
};
struct ClassificationRequest {  // RegressionRequest has the same fields (regression.proto:23-30)
  std::optional<ModelSpec> model_spec;  // tag 1
  std::optional<Input> input;           // tag 2
};
size_t encoded_len(const Input& m);
size_t encoded_len(const ClassificationRequest& m);
void encode_raw(const Input& m, BufMut& buf);
void encode_raw(const ClassificationRequest& m, BufMut& buf);

// ---- prost::encoding primitives ----
size_t encoded_len_varint(uint64_t value);
void encode_varint(uint64_t value, BufMut& buf);
bool decode_varint(Buf& buf, uint64_t* out);  // false => "invalid varint"

// ---- prost::Message surface ----
size_t encoded_len(const TensorProto& m);
size_t encoded_len(const PredictRequest& m);
size_t encoded_len(const PredictResponse& m);
void encode_raw(const TensorProto& m, BufMut& buf);
void encode_raw(const PredictRequest& m, BufMut& buf);
void encode_raw(const PredictResponse& m, BufMut& buf);

// tonic ProstEncoder::encode: len = encoded_len(); reserve(len); Message::encode (which sizes again)
template <class M>
void prost_encoder_encode(const M& m, std::vector<uint8_t>& dst) {
  size_t len = encoded_len(m);
  dst.reserve(dst.size() + len);
  size_t required = encoded_len(m);  // Message::encode's own capacity check
  (void)required;
  BufMut b(dst);
  encode_raw(m, b);
}

// Message::decode: default value, then merge until the buffer is empty.
bool decode(const uint8_t* p, size_t n, TensorProto* out, DecodeError* err);
bool decode(const uint8_t* p, size_t n, PredictRequest* out, DecodeError* err);
bool decode(const uint8_t* p, size_t n, PredictResponse* out, DecodeError* err);

// Rust core::str::from_utf8 validity.
bool is_utf8(const uint8_t* p, size_t n);

// f32 -> f16 / bf16 bit patterns, round-to-nearest-even, NaN stays a quiet NaN with its sign
// (the cast policy of DESIGN.md; the reference has no half code, tensor.proto:43-45 only fixes the carrier).
uint16_t f32_to_f16_bits(float f);
uint16_t f32_to_bf16_bits(float f);
float f16_bits_to_f32(uint16_t h);
float bf16_bits_to_f32(uint16_t h);

}  // namespace cpu_ref
