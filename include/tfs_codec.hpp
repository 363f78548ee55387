// tfs_codec.hpp — C++ host-side mirror of the seam the reference uses for its hot path, written on top of the C ABI
// (include/tfs_gpu.h) only.
//
// The reference is Rust; its toolchain is absent on the build box, so the host side is C++ (the Rust shim with the same
// shape is in rust/codec.rs).  Names follow the reference: the message structs are the prost-generated ones the client
// fills at src/lib.rs:249-263 (`TensorProto { dtype, tensor_shape, float_val, .. }`, `PredictRequest { model_spec,
// inputs, output_filter }`, `PredictResponse { outputs, model_spec }`, `ModelSpec` from build_model_spec,
// src/lib.rs:433-447), and `GpuCodec` has tonic's `Codec` surface: `encoder()` / `decoder()`, whose `encode(item, dst)`
// appends exactly the protobuf bytes of one message to a byte buffer (tonic adds the 5-byte frame prefix itself) and
// whose `decode(src)` consumes one complete message body — what `ProstCodec` does around
// `self.prediction_client.predict(request)` (src/lib.rs:265).  Errors surface like tonic's: a `Status` with code
// Internal and the text of prost's DecodeError class.
//
// Maps are ordered vectors of (key, TensorProto): entries are emitted in the order given (prost iterates a HashMap in
// unspecified order); on decode a repeated key keeps the last entry, as HashMap::insert does.
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <deque>
#include <optional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "tfs_gpu.h"

namespace tfs_codec {

struct Status : std::runtime_error {  // tonic::Status { code: Internal, message }
  int code;                           // TFS_* status that caused it
  Status(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

namespace tensor_shape_proto {
struct Dim {
  int64_t size = 0;
  std::string name;  // always "" on this path (src/lib.rs:233)
};
}  // namespace tensor_shape_proto

struct TensorShapeProto {
  std::vector<tensor_shape_proto::Dim> dim;
  bool unknown_rank = false;
};

struct TensorProto {
  int32_t dtype = 0;
  std::optional<TensorShapeProto> tensor_shape;
  int32_t version_number = 0;
  std::vector<uint8_t> tensor_content;
  std::vector<int32_t> half_val;
  std::vector<float> float_val;
  std::vector<double> double_val;
  std::vector<int32_t> int_val;
  std::vector<std::vector<uint8_t>> string_val;  // Vec<Vec<u8>>
  std::vector<float> scomplex_val;
  std::vector<int64_t> int64_val;
  std::vector<uint8_t> bool_val;  // Vec<bool>
  std::vector<double> dcomplex_val;
  std::vector<uint32_t> uint32_val;
  std::vector<uint64_t> uint64_val;
  // `repeated ResourceHandleProto` / `repeated VariantTensorDataProto` (tensor.proto:72-76): element i is the ENCODED BODY
  // of message i; the library writes / validates them, the owner of the struct opens them (a variant's `tensors` are
  // TensorProtos again: decode them with msg_kind TFS_MSG_TENSOR_PROTO)
  std::vector<std::vector<uint8_t>> resource_handle_val;
  std::vector<std::vector<uint8_t>> variant_val;
};

struct ModelSpec {
  std::string name;
  std::optional<int64_t> version;            // oneof version_choice { Int64Value version = 2;
  std::optional<std::string> version_label;  //                        string version_label = 4 }
  std::string signature_name;
};

using TensorMap = std::vector<std::pair<std::string, TensorProto>>;

struct PredictRequest {
  std::optional<ModelSpec> model_spec;
  TensorMap inputs;
  std::vector<std::string> output_filter;
};

struct PredictResponse {
  TensorMap outputs;
  std::optional<ModelSpec> model_spec;
};

// ---- the metadata call (reference src/lib.rs:300-312; get_model_metadata.proto:16-35) ----
struct Any {  // prost_types::Any: the value stays opaque bytes (the reference never unpacks the SignatureDefMap)
  std::string type_url;
  std::vector<uint8_t> value;
};
struct GetModelMetadataRequest {
  std::optional<ModelSpec> model_spec;
  std::vector<std::string> metadata_field;
};
struct GetModelMetadataResponse {
  std::optional<ModelSpec> model_spec;
  std::vector<std::pair<std::string, Any>> metadata;  // HashMap<String, Any>: a repeated key replaces the earlier entry
};

// ---- Classify / Regress inputs: `Payload`, `impl From<Payload> for Feature`, `build_input` (reference src/lib.rs:413-431,
//      497-571) ----
struct Payload {  // enum Payload { Bytes(Vec<Vec<u8>>), Ints(Vec<i64>), Floats(Vec<f32>) }
  enum Kind { None = 0, Bytes = 1, Floats = 2, Ints = 3 } kind = None;  // numbered like Feature.kind (feature.proto:79-81)
  std::vector<std::vector<uint8_t>> bytes;
  std::vector<float> floats;
  std::vector<int64_t> ints;
  static Payload Of(std::vector<std::vector<uint8_t>> v) { Payload p; p.kind = Bytes; p.bytes = std::move(v); return p; }
  static Payload Of(std::vector<float> v) { Payload p; p.kind = Floats; p.floats = std::move(v); return p; }
  static Payload Of(std::vector<int64_t> v) { Payload p; p.kind = Ints; p.ints = std::move(v); return p; }
};
using Feature = Payload;                                                // Feature { kind: Some(list) } carries exactly a Payload
using Features = std::vector<std::pair<std::string, Feature>>;          // map<string, Feature>, in the caller's order
struct Example {
  Features features;  // Option<Features> is always Some in build_input
};
struct Input {
  std::vector<Example> examples;  // kind: Some(ExampleList { examples })
};
// `fn build_input(&self, payload_map: HashMap<S, V>) -> Input`: one Example holding the whole map
inline Input build_input(Features payload_map) {
  Input in;
  in.examples.push_back(Example{std::move(payload_map)});
  return in;
}
struct ClassificationRequest {  // RegressionRequest has the same two fields
  std::optional<ModelSpec> model_spec;
  Input input;
};

// One tfs_ctx_t (one GPU); shared by the codecs made from it, like the client's channels are shared by its clones.
class GpuContext {
 public:
  explicit GpuContext(int device = 0) {
    const tfs_status_t st = tfs_ctx_create(device, 0, &ctx_);
    if (st != TFS_OK) throw Status(st, "no usable CUDA device: this codec has no CPU fallback");
  }
  ~GpuContext() { tfs_ctx_destroy(ctx_); }
  GpuContext(const GpuContext&) = delete;
  GpuContext& operator=(const GpuContext&) = delete;
  tfs_ctx_t* get() const { return ctx_; }
  void check(tfs_status_t st) const {
    if (st != TFS_OK) throw Status(st, std::string(tfs_strerror(st)) + ": " + tfs_last_error_detail(ctx_));
  }

 private:
  tfs_ctx_t* ctx_ = nullptr;
};

namespace detail {

inline tfs_str_t str(const std::string& s) { return tfs_str_t{s.data(), static_cast<uint32_t>(s.size()), 0}; }

// plain-data picture of a batch of messages: the descriptors point into the messages' vectors; only string_val
// (a vector of vectors) is flattened into a block of its own
struct Described {
  std::deque<std::vector<uint64_t>> ragged;  // flattened string_val blocks (stable addresses)
  std::vector<tfs_msg_desc_t> msgs;
  std::vector<tfs_tensor_desc_t> tensors;
  std::vector<tfs_field_desc_t> fields;
  std::vector<tfs_str_t> filters;
  tfs_batch_t batch() const {
    return tfs_batch_t{msgs.data(), tensors.data(), fields.data(), filters.data(), static_cast<uint32_t>(msgs.size()),
                       static_cast<uint32_t>(tensors.size()), static_cast<uint32_t>(fields.size()), static_cast<uint32_t>(filters.size())};
  }
  void add_tensor(const std::string& name, const TensorProto& t) {
    tfs_tensor_desc_t d;
    std::memset(&d, 0, sizeof(d));
    d.name = str(name);
    d.dtype = t.dtype;
    d.version_number = t.version_number;
    if (t.tensor_shape) {
      if (t.tensor_shape->dim.size() > TFS_MAX_DIMS) throw Status(TFS_ERR_UNSUPPORTED, "rank above TFS_MAX_DIMS: leave this message to prost");
      d.has_shape = 1;
      d.unknown_rank = t.tensor_shape->unknown_rank;
      d.n_dims = static_cast<uint32_t>(t.tensor_shape->dim.size());
      for (uint32_t i = 0; i < d.n_dims; ++i) d.dims[i] = t.tensor_shape->dim[i].size;
    }
    d.first_field = static_cast<uint32_t>(fields.size());
    auto push = [&](int32_t kind, int32_t ty, size_t n, const void* p) {
      if (n == 0) return;
      tfs_field_desc_t f;
      std::memset(&f, 0, sizeof(f));
      f.kind = kind;
      f.src_type = ty;
      f.n_elems = n;
      f.data = p;
      fields.push_back(f);
    };
    push(TFS_KIND_TENSOR_CONTENT, TFS_T_U8, t.tensor_content.size(), t.tensor_content.data());
    push(TFS_KIND_FLOAT_VAL, TFS_T_F32, t.float_val.size(), t.float_val.data());
    push(TFS_KIND_DOUBLE_VAL, TFS_T_F64, t.double_val.size(), t.double_val.data());
    push(TFS_KIND_INT_VAL, TFS_T_I32, t.int_val.size(), t.int_val.data());
    auto push_ragged = [&](int32_t kind, const std::vector<std::vector<uint8_t>>& elems) {
      if (elems.empty()) return;  // Vec<Vec<u8>> flattened into one TFS_T_RAGGED block: (n + 1) u64 offsets, then the bytes
      const size_t n = elems.size();
      size_t total = 0;
      for (const auto& e : elems) total += e.size();
      ragged.emplace_back((8 * (n + 1) + total + 7) / 8);  // uint64_t storage keeps the block 8-byte aligned
      uint64_t* offs = ragged.back().data();
      uint8_t* bytes = reinterpret_cast<uint8_t*>(offs + n + 1);
      uint64_t at = 0;
      for (size_t i = 0; i < n; ++i) {
        offs[i] = at;
        if (!elems[i].empty()) std::memcpy(bytes + at, elems[i].data(), elems[i].size());
        at += elems[i].size();
      }
      offs[n] = at;
      push(kind, TFS_T_RAGGED, n, offs);
      fields.back().aux = total;
    };
    push_ragged(TFS_KIND_STRING_VAL, t.string_val);
    push_ragged(TFS_KIND_RESOURCE_HANDLE_VAL, t.resource_handle_val);
    push_ragged(TFS_KIND_VARIANT_VAL, t.variant_val);
    push(TFS_KIND_SCOMPLEX_VAL, TFS_T_F32, t.scomplex_val.size(), t.scomplex_val.data());
    push(TFS_KIND_INT64_VAL, TFS_T_I64, t.int64_val.size(), t.int64_val.data());
    push(TFS_KIND_BOOL_VAL, TFS_T_BOOL, t.bool_val.size(), t.bool_val.data());
    push(TFS_KIND_DCOMPLEX_VAL, TFS_T_F64, t.dcomplex_val.size(), t.dcomplex_val.data());
    push(TFS_KIND_HALF_VAL, TFS_T_I32, t.half_val.size(), t.half_val.data());
    push(TFS_KIND_UINT32_VAL, TFS_T_U32, t.uint32_val.size(), t.uint32_val.data());
    push(TFS_KIND_UINT64_VAL, TFS_T_U64, t.uint64_val.size(), t.uint64_val.data());
    d.n_fields = static_cast<uint32_t>(fields.size()) - d.first_field;
    tensors.push_back(d);
  }
  std::vector<uint64_t>* flatten(const std::vector<std::vector<uint8_t>>& v, uint64_t* total_out) {
    const size_t n = v.size();
    size_t total = 0;
    for (const auto& e : v) total += e.size();
    ragged.emplace_back((8 * (n + 1) + total + 7) / 8);
    uint64_t* offs = ragged.back().data();
    uint8_t* bytes = reinterpret_cast<uint8_t*>(offs + n + 1);
    uint64_t at = 0;
    for (size_t i = 0; i < n; ++i) {
      offs[i] = at;
      if (!v[i].empty()) std::memcpy(bytes + at, v[i].data(), v[i].size());
      at += v[i].size();
    }
    offs[n] = at;
    *total_out = total;
    return &ragged.back();
  }
  // TFS_MSG_INPUT / TFS_MSG_CLASSIFY_REQUEST: one descriptor per feature, version_number = index of its Example
  void add_input(int32_t kind, const std::optional<ModelSpec>& ms, const Input& in) {
    tfs_msg_desc_t m;
    std::memset(&m, 0, sizeof(m));
    m.kind = kind;
    m.first_tensor = static_cast<uint32_t>(tensors.size());
    for (size_t e = 0; e < in.examples.size(); ++e) {
      for (const auto& kv : in.examples[e].features) {
        tfs_tensor_desc_t d;
        std::memset(&d, 0, sizeof(d));
        d.name = str(kv.first);
        d.version_number = static_cast<int32_t>(e);
        d.first_field = static_cast<uint32_t>(fields.size());
        const Feature& ft = kv.second;
        if (ft.kind != Payload::None) {
          tfs_field_desc_t f;
          std::memset(&f, 0, sizeof(f));
          f.kind = static_cast<int32_t>(ft.kind);
          if (ft.kind == Payload::Bytes) {
            f.src_type = TFS_T_RAGGED;
            f.n_elems = ft.bytes.size();
            f.data = flatten(ft.bytes, &f.aux)->data();
          } else if (ft.kind == Payload::Floats) {
            f.src_type = TFS_T_F32;
            f.n_elems = ft.floats.size();
            f.data = ft.floats.data();
          } else {
            f.src_type = TFS_T_I64;
            f.n_elems = ft.ints.size();
            f.data = ft.ints.data();
          }
          fields.push_back(f);
          d.n_fields = 1;
        }
        tensors.push_back(d);
      }
    }
    if (ms && kind == TFS_MSG_CLASSIFY_REQUEST) {
      m.has_model_spec = 1;
      m.model_name = str(ms->name);
      m.signature_name = str(ms->signature_name);
      if (ms->version) {
        m.version_choice = TFS_VERSION_NUMBER;
        m.version = *ms->version;
      } else if (ms->version_label) {
        m.version_choice = TFS_VERSION_LABEL;
        m.version_label = str(*ms->version_label);
      }
    }
    m.n_tensors = static_cast<uint32_t>(tensors.size()) - m.first_tensor;
    m.n_filters = static_cast<uint32_t>(in.examples.size());  // the number of Examples (see TFS_MSG_INPUT)
    msgs.push_back(m);
  }
  void add_message(int32_t kind, const std::optional<ModelSpec>& ms, const TensorMap& map, const std::vector<std::string>* filter) {
    tfs_msg_desc_t m;
    std::memset(&m, 0, sizeof(m));
    m.kind = kind;
    m.first_tensor = static_cast<uint32_t>(tensors.size());
    m.first_filter = static_cast<uint32_t>(filters.size());
    for (const auto& kv : map) add_tensor(kv.first, kv.second);
    if (ms) {
      m.has_model_spec = 1;
      m.model_name = str(ms->name);
      m.signature_name = str(ms->signature_name);
      if (ms->version) {
        m.version_choice = TFS_VERSION_NUMBER;
        m.version = *ms->version;
      } else if (ms->version_label) {
        m.version_choice = TFS_VERSION_LABEL;
        m.version_label = str(*ms->version_label);
      }
    }
    if (filter)
      for (const auto& f : *filter) filters.push_back(str(f));
    m.n_tensors = static_cast<uint32_t>(tensors.size()) - m.first_tensor;
    m.n_filters = static_cast<uint32_t>(filters.size()) - m.first_filter;
    msgs.push_back(m);
  }
};

template <class T>
void take(std::vector<T>& dst, const tfs_tensor_out_t& t, int slot, const uint8_t* pool) {
  const uint64_t n = t.n_elems[slot];
  if (n == 0) return;
  const T* p = reinterpret_cast<const T*>(pool + t.dst_off[slot]);
  dst.assign(p, p + n);  // one exact-capacity copy (prost grows its Vecs by doubling instead)
}

inline TensorProto tensor_from(const tfs_tensor_out_t& t, const uint8_t* pool) {
  TensorProto o;
  o.dtype = t.dtype;
  o.version_number = t.version_number;
  if (t.has_shape) {
    TensorShapeProto s;
    s.unknown_rank = t.unknown_rank != 0;
    for (uint32_t i = 0; i < t.n_dims; ++i) s.dim.push_back(tensor_shape_proto::Dim{t.dims[i], ""});
    o.tensor_shape = std::move(s);
  }
  take(o.tensor_content, t, 0, pool);
  take(o.float_val, t, 1, pool);
  take(o.double_val, t, 2, pool);
  take(o.int_val, t, 3, pool);
  auto take_ragged = [&](std::vector<std::vector<uint8_t>>& dst, int slot) {  // comes back as the same flattened block
    const uint64_t n = t.n_elems[slot];
    if (!n) return;
    const uint8_t* block = pool + t.dst_off[slot];
    const uint8_t* bytes = block + 8 * (n + 1);
    dst.resize(n);
    for (uint64_t i = 0; i < n; ++i) {
      uint64_t a, b;
      std::memcpy(&a, block + 8 * i, 8);
      std::memcpy(&b, block + 8 * (i + 1), 8);
      dst[i].assign(bytes + a, bytes + b);
    }
  };
  take_ragged(o.string_val, 4);
  take_ragged(o.resource_handle_val, 12);
  take_ragged(o.variant_val, 13);
  take(o.scomplex_val, t, 5, pool);
  take(o.int64_val, t, 6, pool);
  take(o.bool_val, t, 7, pool);
  take(o.dcomplex_val, t, 8, pool);
  take(o.half_val, t, 9, pool);
  take(o.uint32_val, t, 10, pool);
  take(o.uint64_val, t, 11, pool);
  return o;
}

inline std::string text(const tfs_decoded_view_t& v, const tfs_span_t& s) {
  return std::string(reinterpret_cast<const char*>(v.strings) + s.off, s.len);
}

inline std::optional<ModelSpec> spec_from(const tfs_decoded_view_t& v, const tfs_msg_out_t& m) {
  if (!m.has_model_spec) return std::nullopt;
  ModelSpec ms;
  ms.name = text(v, m.model_name);
  ms.signature_name = text(v, m.signature_name);
  if (m.version_choice == TFS_VERSION_NUMBER) ms.version = m.version;
  if (m.version_choice == TFS_VERSION_LABEL) ms.version_label = text(v, m.version_label);
  return ms;
}

// HashMap::insert semantics on an ordered vector: a repeated key keeps its first position and the LAST value
inline void insert_last_wins(TensorMap& map, std::string key, TensorProto t) {
  for (auto& kv : map)
    if (kv.first == key) {
      kv.second = std::move(t);
      return;
    }
  map.emplace_back(std::move(key), std::move(t));
}

}  // namespace detail

// tonic `Encoder<Item = PredictRequest>`: appends the protobuf bytes of the message(s) to dst.
class GpuEncoder {
 public:
  explicit GpuEncoder(std::shared_ptr<GpuContext> ctx) : ctx_(std::move(ctx)) {}
  void encode(const PredictRequest& item, std::vector<uint8_t>& dst) const { encode_batch(&item, 1, dst, nullptr); }
  void encode(const PredictResponse& item, std::vector<uint8_t>& dst) const {
    detail::Described d;
    d.add_message(TFS_MSG_PREDICT_RESPONSE, item.model_spec, item.outputs, nullptr);
    run(d, dst, nullptr);
  }
  void encode(const GetModelMetadataRequest& item, std::vector<uint8_t>& dst) const {
    detail::Described d;
    d.add_message(TFS_MSG_METADATA_REQUEST, item.model_spec, TensorMap{}, &item.metadata_field);
    run(d, dst, nullptr);
  }
  // Classify / Regress: what `build_input` returns, and the request around it
  void encode(const Input& item, std::vector<uint8_t>& dst) const {
    detail::Described d;
    d.add_input(TFS_MSG_INPUT, std::nullopt, item);
    run(d, dst, nullptr);
  }
  void encode(const ClassificationRequest& item, std::vector<uint8_t>& dst) const {
    detail::Described d;
    d.add_input(TFS_MSG_CLASSIFY_REQUEST, item.model_spec, item.input);
    run(d, dst, nullptr);
  }
  // the batched form the shim uses for the requests of one window: message i occupies
  // dst[base + offsets[i] .. + lengths[i])
  void encode_batch(const PredictRequest* items, size_t n, std::vector<uint8_t>& dst,
                    std::vector<std::pair<uint64_t, uint64_t>>* spans) const {
    detail::Described d;
    for (size_t i = 0; i < n; ++i) d.add_message(TFS_MSG_PREDICT_REQUEST, items[i].model_spec, items[i].inputs, &items[i].output_filter);
    run(d, dst, spans);
  }
  // prost Message::encoded_len
  uint64_t encoded_len(const PredictRequest& item) const {
    detail::Described d;
    d.add_message(TFS_MSG_PREDICT_REQUEST, item.model_spec, item.inputs, &item.output_filter);
    // the size pass works on device data: stage through the host entry point's bound instead
    std::vector<uint8_t> tmp;
    std::vector<std::pair<uint64_t, uint64_t>> spans;
    run(d, tmp, &spans);
    return spans[0].second;
  }

 private:
  void run(const detail::Described& d, std::vector<uint8_t>& dst, std::vector<std::pair<uint64_t, uint64_t>>* spans) const {
    const tfs_batch_t b = d.batch();
    const uint64_t bound = tfs_encode_bound(&b, 0);
    if (bound == 0) throw Status(TFS_ERR_INVALID_ARG, "message cannot be described to the codec");
    const size_t base = dst.size();
    dst.resize(base + bound);  // BytesMut::reserve(encoded_len) in tonic's ProstEncoder
    std::vector<uint64_t> off(b.n_msgs), len(b.n_msgs);
    uint64_t total = 0;
    ctx_->check(tfs_encode_host(ctx_->get(), &b, dst.data() + base, bound, 0, off.data(), len.data(), &total));
    dst.resize(base + total);
    if (spans) {
      spans->clear();
      for (uint32_t i = 0; i < b.n_msgs; ++i) spans->emplace_back(off[i], len[i]);
    }
  }
  std::shared_ptr<GpuContext> ctx_;
};

// tonic `Decoder<Item = PredictResponse>`: consumes one complete message body.
class GpuDecoder {
 public:
  explicit GpuDecoder(std::shared_ptr<GpuContext> ctx) : ctx_(std::move(ctx)) {}

  std::optional<PredictResponse> decode(std::vector<uint8_t>& src) const {
    PredictResponse r;
    decode_one(src, TFS_MSG_PREDICT_RESPONSE, [&](const tfs_decoded_view_t& v, const tfs_msg_out_t& m, const uint8_t* pool) {
      r.model_spec = detail::spec_from(v, m);
      for (uint32_t i = 0; i < m.n_tensors; ++i) {
        const tfs_tensor_out_t& t = v.tensors[m.first_tensor + i];
        detail::insert_last_wins(r.outputs, detail::text(v, t.key), detail::tensor_from(t, pool));
      }
    });
    src.clear();
    return r;
  }
  // GetModelMetadataResponse: every map entry comes back as a record with the key and string_val = [type_url, value]
  std::optional<GetModelMetadataResponse> decode_metadata(std::vector<uint8_t>& src) const {
    GetModelMetadataResponse r;
    decode_one(src, TFS_MSG_METADATA_RESPONSE, [&](const tfs_decoded_view_t& v, const tfs_msg_out_t& m, const uint8_t* pool) {
      r.model_spec = detail::spec_from(v, m);
      for (uint32_t i = 0; i < m.n_tensors; ++i) {
        const tfs_tensor_out_t& t = v.tensors[m.first_tensor + i];
        TensorProto parts = detail::tensor_from(t, pool);
        Any a;
        if (parts.string_val.size() == 2) {
          a.type_url.assign(parts.string_val[0].begin(), parts.string_val[0].end());
          a.value = std::move(parts.string_val[1]);
        }
        std::string key = detail::text(v, t.key);
        bool replaced = false;
        for (auto& kv : r.metadata)
          if (kv.first == key) {
            kv.second = std::move(a);
            replaced = true;
            break;
          }
        if (!replaced) r.metadata.emplace_back(std::move(key), std::move(a));
      }
    });
    src.clear();
    return r;
  }
  std::optional<PredictRequest> decode_request(std::vector<uint8_t>& src) const {
    PredictRequest r;
    decode_one(src, TFS_MSG_PREDICT_REQUEST, [&](const tfs_decoded_view_t& v, const tfs_msg_out_t& m, const uint8_t* pool) {
      r.model_spec = detail::spec_from(v, m);
      for (uint32_t i = 0; i < m.n_tensors; ++i) {
        const tfs_tensor_out_t& t = v.tensors[m.first_tensor + i];
        detail::insert_last_wins(r.inputs, detail::text(v, t.key), detail::tensor_from(t, pool));
      }
      for (uint32_t i = 0; i < m.n_filters; ++i) r.output_filter.push_back(detail::text(v, v.filters[m.first_filter + i]));
    });
    src.clear();
    return r;
  }

 private:
  template <class F>
  void decode_one(const std::vector<uint8_t>& src, int32_t kind, F fill) const {
    const uint64_t off = 0, len = src.size();
    tfs_decoded_t* dec = nullptr;
    const tfs_status_t st = tfs_decode_host(ctx_->get(), src.data(), len, &off, &len, 1, kind, 0, &dec);
    struct Guard {
      tfs_ctx_t* c;
      tfs_decoded_t* d;
      ~Guard() { if (d) tfs_decoded_free(c, d); }
    } guard{ctx_->get(), dec};
    if (st != TFS_OK && st != TFS_ERR_BATCH) ctx_->check(st);
    tfs_decoded_view_t v;
    ctx_->check(tfs_decoded_view(dec, &v));
    const tfs_msg_out_t& m = v.msgs[0];
    if (m.status != TFS_OK) throw Status(m.status, tfs_strerror(m.status));  // Status::internal(DecodeError text)
    std::vector<uint8_t> pool(v.pool_bytes ? v.pool_bytes : 1);
    ctx_->check(tfs_decode_host_finish(ctx_->get(), dec, pool.data(), pool.size()));
    fill(v, m, pool.data());
  }
  std::shared_ptr<GpuContext> ctx_;
};

// `impl tonic::codec::Codec for GpuCodec { type Encode = PredictRequest; type Decode = PredictResponse; .. }`
class GpuCodec {
 public:
  explicit GpuCodec(std::shared_ptr<GpuContext> ctx = std::make_shared<GpuContext>(0)) : ctx_(std::move(ctx)) {}
  GpuEncoder encoder() const { return GpuEncoder(ctx_); }
  GpuDecoder decoder() const { return GpuDecoder(ctx_); }

 private:
  std::shared_ptr<GpuContext> ctx_;
};

}  // namespace tfs_codec
