// cpu_ref_capi.cpp — CPU ORACLE (test infrastructure, NOT product code): a C surface over cpu_ref.cpp
// that takes the SAME plain-data descriptors as include/tfs_gpu.h, so tests and bench.py can feed one
// batch to both the CUDA library and this restatement and compare byte for byte.
//
// What the reference does on the CPU for this path and what is restated here:
//   - message construction + the u8 -> f32 (+ preprocessing fn) cast loop  src/lib.rs:229-263
//   - tonic ProstEncoder::encode -> prost Message::encode                  reached from src/lib.rs:265
//   - tonic ProstDecoder::decode -> prost Message::decode                  src/lib.rs:265-266
// "parity unpinned": see cpu_ref.hpp.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <memory>
#include <thread>

#include "../include/tfs_gpu.h"
#include "cpu_ref.hpp"

using namespace cpu_ref;

namespace {

uint32_t type_size(int32_t t) {
  switch (t) {
    case TFS_T_U8: case TFS_T_I8: case TFS_T_BOOL: return 1;
    case TFS_T_U16: case TFS_T_I16: case TFS_T_F16: case TFS_T_BF16: return 2;
    case TFS_T_I32: case TFS_T_U32: case TFS_T_F32: return 4;
    case TFS_T_I64: case TFS_T_U64: case TFS_T_F64: return 8;
    default: return 0;
  }
}
int slot_of_kind(int32_t kind) {
  static const int kinds[TFS_N_SLOTS] = {4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 16, 17, 14, 15};
  for (int i = 0; i < TFS_N_SLOTS; ++i)
    if (kinds[i] == kind) return i;
  return -1;
}

template <class T>
const T* as(const tfs_field_desc_t& f) { return static_cast<const T*>(f.data); }

// Build one payload field of a prost TensorProto the way the reference's CPU code would hold it
// (e.g. `pixels: Vec<f32>` from `raw_pixels().iter().map(|p| *p as f32).map(preprocessing_fn)`,
// src/lib.rs:237-242).  Returns false on a descriptor the ABI does not allow.
bool build_field(const tfs_field_desc_t& f, TensorProto* t) {
  const uint64_t n = f.n_elems;
  const int32_t dst = f.dst_type;
  if (n != 0 && f.data == nullptr) return false;
  if (f.op != TFS_OP_NONE && !(f.src_type == TFS_T_U8 && (f.kind == TFS_KIND_FLOAT_VAL || f.kind == TFS_KIND_SCOMPLEX_VAL)))
    return false;
  switch (f.kind) {
    case TFS_KIND_FLOAT_VAL:
    case TFS_KIND_SCOMPLEX_VAL: {
      std::vector<float>& v = f.kind == TFS_KIND_FLOAT_VAL ? t->float_val : t->scomplex_val;
      if (dst != TFS_T_NONE && dst != TFS_T_F32) return false;
      if (f.src_type == TFS_T_F32) {
        v.assign(as<float>(f), as<float>(f) + n);
      } else if (f.src_type == TFS_T_U8) {
        v.resize(n);
        const uint8_t* s = as<uint8_t>(f);
        for (uint64_t i = 0; i < n; ++i) {
          float x = static_cast<float>(s[i]);  // `*p as f32`
          if (f.op == TFS_OP_DIV) x = x / f.op_param;       // e.g. |v| v / 255.  (examples/prediction.rs:43)
          else if (f.op == TFS_OP_MUL) x = x * f.op_param;
          v[i] = x;
        }
      } else {
        return false;
      }
      return true;
    }
    case TFS_KIND_DOUBLE_VAL:
    case TFS_KIND_DCOMPLEX_VAL: {
      std::vector<double>& v = f.kind == TFS_KIND_DOUBLE_VAL ? t->double_val : t->dcomplex_val;
      if (f.src_type != TFS_T_F64 || (dst != TFS_T_NONE && dst != TFS_T_F64)) return false;
      v.assign(as<double>(f), as<double>(f) + n);
      return true;
    }
    case TFS_KIND_INT_VAL: {
      if (dst != TFS_T_NONE && dst != TFS_T_I32) return false;
      std::vector<int32_t>& v = t->int_val;
      v.resize(n);
      switch (f.src_type) {
        case TFS_T_I32: std::copy(as<int32_t>(f), as<int32_t>(f) + n, v.begin()); break;
        case TFS_T_I8: for (uint64_t i = 0; i < n; ++i) v[i] = as<int8_t>(f)[i]; break;
        case TFS_T_U8: for (uint64_t i = 0; i < n; ++i) v[i] = as<uint8_t>(f)[i]; break;
        case TFS_T_I16: for (uint64_t i = 0; i < n; ++i) v[i] = as<int16_t>(f)[i]; break;
        case TFS_T_U16: for (uint64_t i = 0; i < n; ++i) v[i] = as<uint16_t>(f)[i]; break;
        default: return false;
      }
      return true;
    }
    case TFS_KIND_INT64_VAL:
      if (f.src_type != TFS_T_I64 || (dst != TFS_T_NONE && dst != TFS_T_I64)) return false;
      t->int64_val.assign(as<int64_t>(f), as<int64_t>(f) + n);
      return true;
    case TFS_KIND_UINT32_VAL:
      if (f.src_type != TFS_T_U32 || (dst != TFS_T_NONE && dst != TFS_T_U32)) return false;
      t->uint32_val.assign(as<uint32_t>(f), as<uint32_t>(f) + n);
      return true;
    case TFS_KIND_UINT64_VAL:
      if (f.src_type != TFS_T_U64 || (dst != TFS_T_NONE && dst != TFS_T_U64)) return false;
      t->uint64_val.assign(as<uint64_t>(f), as<uint64_t>(f) + n);
      return true;
    case TFS_KIND_BOOL_VAL:
      if ((f.src_type != TFS_T_BOOL && f.src_type != TFS_T_U8) || (dst != TFS_T_NONE && dst != TFS_T_BOOL)) return false;
      t->bool_val.assign(as<uint8_t>(f), as<uint8_t>(f) + n);
      return true;
    case TFS_KIND_HALF_VAL: {
      std::vector<int32_t>& v = t->half_val;
      v.resize(n);
      if (f.src_type == TFS_T_I32 && (dst == TFS_T_NONE || dst == TFS_T_I32)) {
        std::copy(as<int32_t>(f), as<int32_t>(f) + n, v.begin());
      } else if ((f.src_type == TFS_T_U16 || f.src_type == TFS_T_F16 || f.src_type == TFS_T_BF16) &&
                 (dst == TFS_T_NONE || dst == f.src_type || dst == TFS_T_I32)) {
        for (uint64_t i = 0; i < n; ++i) v[i] = as<uint16_t>(f)[i];  // 16-bit pattern, upper bits zero (tensor.proto:43-45)
      } else if (f.src_type == TFS_T_F32 && dst == TFS_T_F16) {
        for (uint64_t i = 0; i < n; ++i) v[i] = f32_to_f16_bits(as<float>(f)[i]);
      } else if (f.src_type == TFS_T_F32 && dst == TFS_T_BF16) {
        for (uint64_t i = 0; i < n; ++i) v[i] = f32_to_bf16_bits(as<float>(f)[i]);
      } else {
        return false;
      }
      return true;
    }
    case TFS_KIND_TENSOR_CONTENT: {
      std::vector<uint8_t>& v = t->tensor_content;
      uint32_t w = type_size(f.src_type);
      if (w == 0) return false;
      if (dst == TFS_T_NONE || dst == f.src_type) {
        v.assign(as<uint8_t>(f), as<uint8_t>(f) + n * w);  // raw little-endian tensor memory (tensor.proto:31-36)
      } else if (f.src_type == TFS_T_F32 && (dst == TFS_T_F16 || dst == TFS_T_BF16)) {
        v.resize(n * 2);
        for (uint64_t i = 0; i < n; ++i) {
          uint16_t h = dst == TFS_T_F16 ? f32_to_f16_bits(as<float>(f)[i]) : f32_to_bf16_bits(as<float>(f)[i]);
          std::memcpy(&v[2 * i], &h, 2);
        }
      } else {
        return false;
      }
      return true;
    }
    case TFS_KIND_RESOURCE_HANDLE_VAL:  // the same block; every element is the encoded body of one message (tensor.proto:72-76)
    case TFS_KIND_VARIANT_VAL:
    case TFS_KIND_STRING_VAL: {  // TFS_T_RAGGED block: (n + 1) u64 offsets, then the bytes (tensor.proto:57)
      if (f.src_type != TFS_T_RAGGED || (dst != TFS_T_NONE && dst != TFS_T_RAGGED)) return false;
      std::vector<std::vector<uint8_t>>& elems = f.kind == TFS_KIND_STRING_VAL ? t->string_val : f.kind == TFS_KIND_VARIANT_VAL ? t->variant_val : t->resource_handle_val;
      const uint8_t* blk = as<uint8_t>(f);
      std::vector<uint64_t> off(n + 1);
      std::memcpy(off.data(), blk, 8 * (n + 1));
      if (off[0] != 0 || off[n] != f.aux) return false;
      const uint8_t* bytes = blk + 8 * (n + 1);
      elems.resize(n);
      for (uint64_t i = 0; i < n; ++i) {
        if (off[i + 1] < off[i]) return false;
        elems[i].assign(bytes + off[i], bytes + off[i + 1]);
      }
      return true;
    }
    default: return false;
  }
}

bool build_tensor(const tfs_batch_t& b, const tfs_tensor_desc_t& td, TensorProto* t) {
  t->dtype = td.dtype;
  t->version_number = td.version_number;
  if (td.n_dims > TFS_MAX_DIMS) return false;
  if (td.has_shape) {
    TensorShapeProto s;
    for (uint32_t i = 0; i < td.n_dims; ++i) s.dim.push_back(Dim{td.dims[i], ""});  // name "" (src/lib.rs:233)
    s.unknown_rank = td.unknown_rank != 0;
    t->tensor_shape = std::move(s);
  } else if (td.n_dims != 0 || td.unknown_rank) {
    return false;
  }
  if (static_cast<uint64_t>(td.first_field) + td.n_fields > b.n_fields) return false;
  uint32_t seen = 0;
  for (uint32_t i = 0; i < td.n_fields; ++i) {
    const tfs_field_desc_t& f = b.fields[td.first_field + i];
    int slot = slot_of_kind(f.kind);
    if (slot < 0) return false;
    if ((seen >> slot) & 1u) {  // the same kind again: concatenated in descriptor order (what the caller's Vec would hold)
      TensorProto more;
      if (!build_field(f, &more)) return false;
      auto app = [](auto& dst, auto& src) { dst.insert(dst.end(), src.begin(), src.end()); };
      app(t->tensor_content, more.tensor_content);
      app(t->half_val, more.half_val);
      app(t->float_val, more.float_val);
      app(t->double_val, more.double_val);
      app(t->int_val, more.int_val);
      app(t->string_val, more.string_val);
      app(t->scomplex_val, more.scomplex_val);
      app(t->int64_val, more.int64_val);
      app(t->bool_val, more.bool_val);
      app(t->dcomplex_val, more.dcomplex_val);
      app(t->uint32_val, more.uint32_val);
      app(t->uint64_val, more.uint64_val);
      app(t->resource_handle_val, more.resource_handle_val);
      app(t->variant_val, more.variant_val);
      continue;
    }
    seen |= 1u << slot;
    if (!build_field(f, t)) return false;
  }
  return true;
}

std::string str(const tfs_str_t& s) { return s.len ? std::string(s.ptr, s.len) : std::string(); }

ModelSpec build_model_spec(const tfs_msg_desc_t& m) {  // build_model_spec, src/lib.rs:433-447
  ModelSpec ms;
  ms.name = str(m.model_name);
  ms.signature_name = str(m.signature_name);
  ms.version_choice = m.version_choice;
  if (m.version_choice == TFS_VERSION_NUMBER) ms.version = m.version;
  if (m.version_choice == TFS_VERSION_LABEL) ms.version_label = str(m.version_label);
  return ms;
}

struct AnyMsg {
  int32_t kind = 0;
  TensorProto tensor;
  PredictRequest req;
  PredictResponse resp;
  Input input;
  ClassificationRequest classify;
  GetModelMetadataRequest meta_req;
  size_t encoded_len() const {
    switch (kind) {
      case TFS_MSG_METADATA_REQUEST: return cpu_ref::encoded_len(meta_req);
      case TFS_MSG_TENSOR_PROTO: return cpu_ref::encoded_len(tensor);
      case TFS_MSG_PREDICT_REQUEST: return cpu_ref::encoded_len(req);
      case TFS_MSG_INPUT: return cpu_ref::encoded_len(input);
      case TFS_MSG_CLASSIFY_REQUEST: return cpu_ref::encoded_len(classify);
      default: return cpu_ref::encoded_len(resp);
    }
  }
  void prost_encode(std::vector<uint8_t>& dst) const {
    switch (kind) {
      case TFS_MSG_METADATA_REQUEST: prost_encoder_encode(meta_req, dst); break;
      case TFS_MSG_TENSOR_PROTO: prost_encoder_encode(tensor, dst); break;
      case TFS_MSG_PREDICT_REQUEST: prost_encoder_encode(req, dst); break;
      case TFS_MSG_INPUT: prost_encoder_encode(input, dst); break;
      case TFS_MSG_CLASSIFY_REQUEST: prost_encoder_encode(classify, dst); break;
      default: prost_encoder_encode(resp, dst);
    }
  }
};

// Feature from its descriptor (`impl From<Payload> for Feature`, src/lib.rs:514-540)
bool build_feature(const tfs_batch_t& b, const tfs_tensor_desc_t& td, Feature* f) {
  if (td.dtype != 0 || td.has_shape || td.n_dims || td.n_fields > 1) return false;
  if (static_cast<uint64_t>(td.first_field) + td.n_fields > b.n_fields) return false;
  if (td.n_fields == 0) return true;  // kind: None
  const tfs_field_desc_t& d = b.fields[td.first_field];
  if (d.op != TFS_OP_NONE || (d.n_elems && !d.data)) return false;
  const uint64_t n = d.n_elems;
  switch (d.kind) {
    case TFS_KIND_BYTES_LIST: {
      if (d.src_type != TFS_T_RAGGED) return false;
      f->kind = 1;
      if (n == 0) return true;
      const uint8_t* blk = static_cast<const uint8_t*>(d.data);
      std::vector<uint64_t> off(n + 1);
      std::memcpy(off.data(), blk, 8 * (n + 1));
      if (off[0] != 0 || off[n] != d.aux) return false;
      const uint8_t* bytes = blk + 8 * (n + 1);
      f->bytes_list.resize(n);
      for (uint64_t i = 0; i < n; ++i) {
        if (off[i + 1] < off[i]) return false;
        f->bytes_list[i].assign(bytes + off[i], bytes + off[i + 1]);
      }
      return true;
    }
    case TFS_KIND_FLOAT_LIST:
      if (d.src_type != TFS_T_F32) return false;
      f->kind = 2;
      f->float_list.assign(static_cast<const float*>(d.data), static_cast<const float*>(d.data) + n);
      return true;
    case TFS_KIND_INT64_LIST:
      if (d.src_type != TFS_T_I64) return false;
      f->kind = 3;
      f->int64_list.assign(static_cast<const int64_t*>(d.data), static_cast<const int64_t*>(d.data) + n);
      return true;
    default: return false;
  }
}

bool build_input(const tfs_batch_t& b, const tfs_msg_desc_t& m, Input* in) {
  in->examples.assign(m.n_filters, Example{});
  for (auto& e : in->examples) e.features = FeatureMap{};  // Some(Features), as build_input does (src/lib.rs:420)
  int64_t last = 0;
  for (uint32_t i = 0; i < m.n_tensors; ++i) {
    const tfs_tensor_desc_t& td = b.tensors[m.first_tensor + i];
    if (td.version_number < last || static_cast<uint32_t>(td.version_number) >= m.n_filters) return false;
    last = td.version_number;
    Feature f;
    if (!build_feature(b, td, &f)) return false;
    in->examples[td.version_number].features->emplace_back(str(td.name), std::move(f));
  }
  return true;
}

bool build_msg(const tfs_batch_t& b, const tfs_msg_desc_t& m, AnyMsg* out) {
  out->kind = m.kind;
  if (static_cast<uint64_t>(m.first_tensor) + m.n_tensors > b.n_tensors) return false;
  const bool input_kind = m.kind == TFS_MSG_INPUT || m.kind == TFS_MSG_CLASSIFY_REQUEST;  // n_filters counts Examples there
  if (!input_kind && static_cast<uint64_t>(m.first_filter) + m.n_filters > b.n_filters) return false;
  if (m.version_choice < 0 || m.version_choice > 2) return false;
  if (m.kind == TFS_MSG_TENSOR_PROTO) {
    if (m.n_tensors != 1 || m.has_model_spec || m.n_filters) return false;
    return build_tensor(b, b.tensors[m.first_tensor], &out->tensor);
  }
  if (m.kind == TFS_MSG_INPUT) {
    if (m.has_model_spec) return false;
    return build_input(b, m, &out->input);
  }
  if (m.kind == TFS_MSG_CLASSIFY_REQUEST) {
    if (m.has_model_spec) out->classify.model_spec = build_model_spec(m);
    out->classify.input = Input{};
    return build_input(b, m, &*out->classify.input);
  }
  if (m.kind == TFS_MSG_METADATA_REQUEST) {  // model_spec + metadata_field (the message's filters); decode-only kinds are refused
    if (m.n_tensors) return false;
    if (m.has_model_spec) out->meta_req.model_spec = build_model_spec(m);
    for (uint32_t i = 0; i < m.n_filters; ++i) out->meta_req.metadata_field.push_back(str(b.filters[m.first_filter + i]));
    return true;
  }
  if (m.kind != TFS_MSG_PREDICT_REQUEST && m.kind != TFS_MSG_PREDICT_RESPONSE) return false;
  TensorMap map;
  for (uint32_t i = 0; i < m.n_tensors; ++i) {
    const tfs_tensor_desc_t& td = b.tensors[m.first_tensor + i];
    TensorProto t;
    if (!build_tensor(b, td, &t)) return false;
    map.emplace_back(str(td.name), std::move(t));
  }
  if (m.kind == TFS_MSG_PREDICT_REQUEST) {
    if (m.has_model_spec) out->req.model_spec = build_model_spec(m);
    out->req.inputs = std::move(map);
    for (uint32_t i = 0; i < m.n_filters; ++i) out->req.output_filter.push_back(str(b.filters[m.first_filter + i]));
  } else {
    if (m.n_filters) return false;
    if (m.has_model_spec) out->resp.model_spec = build_model_spec(m);
    out->resp.outputs = std::move(map);
  }
  return true;
}

// ------------------------------------------------------------------------------------------
// decode results as tfs_gpu.h plain data
// ------------------------------------------------------------------------------------------
struct Decoded {
  std::vector<tfs_msg_out_t> msgs;
  std::vector<tfs_tensor_out_t> tensors;
  std::vector<tfs_span_t> filters;
  std::vector<uint8_t> strings;
  std::vector<uint8_t> pool;
  uint32_t n_failed = 0;
  // kept for the timed baseline / value checks
};

tfs_span_t add_string(Decoded& d, const std::string& s) {
  tfs_span_t sp{d.strings.size(), static_cast<uint32_t>(s.size()), 0};
  d.strings.insert(d.strings.end(), s.begin(), s.end());
  return sp;
}
uint64_t pool_put(Decoded& d, const void* p, size_t bytes) {
  size_t off = (d.pool.size() + 15) & ~static_cast<size_t>(15);
  d.pool.resize(off + bytes);
  if (bytes) std::memcpy(d.pool.data() + off, p, bytes);
  return off;
}

bool emit_tensor(Decoded& d, uint32_t msg_index, const std::string& key, const TensorProto& t, uint32_t flags) {
  tfs_tensor_out_t o;
  std::memset(&o, 0, sizeof(o));
  o.key = add_string(d, key);
  o.msg_index = msg_index;
  o.dtype = t.dtype;
  o.version_number = t.version_number;
  o.has_shape = t.tensor_shape ? 1 : 0;
  if (t.tensor_shape) {
    o.unknown_rank = t.tensor_shape->unknown_rank;
    o.n_dims = static_cast<uint32_t>(t.tensor_shape->dim.size());
    if (o.n_dims > TFS_MAX_DIMS) return false;
    for (uint32_t i = 0; i < o.n_dims; ++i) {
      if (!t.tensor_shape->dim[i].name.empty()) return false;  // dim names -> UNSUPPORTED
      o.dims[i] = t.tensor_shape->dim[i].size;
    }
  }
  const bool is_half = t.dtype == 19, is_bf16 = t.dtype == 14;
  auto put = [&](int kind, int32_t type, const void* p, uint64_t n, size_t bytes) {
    int s = slot_of_kind(kind);
    o.n_elems[s] = n;
    o.dst_type[s] = type;
    o.dst_off[s] = n ? pool_put(d, p, bytes) : 0;
  };
  // tensor_content
  if ((flags & TFS_F_CONTENT_HALF_TO_F32) && (is_half || is_bf16) && t.tensor_content.size() % 2 == 0 && !t.tensor_content.empty()) {
    std::vector<float> w(t.tensor_content.size() / 2);
    for (size_t i = 0; i < w.size(); ++i) {
      uint16_t h;
      std::memcpy(&h, &t.tensor_content[2 * i], 2);
      w[i] = is_half ? f16_bits_to_f32(h) : bf16_bits_to_f32(h);
    }
    put(4, TFS_T_F32, w.data(), t.tensor_content.size(), w.size() * 4);
  } else {
    put(4, TFS_T_U8, t.tensor_content.data(), t.tensor_content.size(), t.tensor_content.size());
  }
  put(5, TFS_T_F32, t.float_val.data(), t.float_val.size(), t.float_val.size() * 4);
  put(6, TFS_T_F64, t.double_val.data(), t.double_val.size(), t.double_val.size() * 8);
  put(7, TFS_T_I32, t.int_val.data(), t.int_val.size(), t.int_val.size() * 4);
  auto put_ragged = [&](int kind, const std::vector<std::vector<uint8_t>>& elems) {  // (n+1) u64 offsets then the bytes
    size_t n = elems.size();
    std::vector<uint8_t> blob((n + 1) * 8);
    uint64_t acc = 0;
    for (size_t i = 0; i < n; ++i) {
      std::memcpy(&blob[i * 8], &acc, 8);
      acc += elems[i].size();
    }
    std::memcpy(&blob[n * 8], &acc, 8);
    for (const auto& s : elems) blob.insert(blob.end(), s.begin(), s.end());
    put(kind, TFS_T_RAGGED, blob.data(), n, n ? blob.size() : 0);
  };
  put_ragged(8, t.string_val);
  put_ragged(14, t.resource_handle_val);
  put_ragged(15, t.variant_val);
  put(9, TFS_T_F32, t.scomplex_val.data(), t.scomplex_val.size(), t.scomplex_val.size() * 4);
  put(10, TFS_T_I64, t.int64_val.data(), t.int64_val.size(), t.int64_val.size() * 8);
  put(11, TFS_T_BOOL, t.bool_val.data(), t.bool_val.size(), t.bool_val.size());
  put(12, TFS_T_F64, t.dcomplex_val.data(), t.dcomplex_val.size(), t.dcomplex_val.size() * 8);
  if ((flags & TFS_F_HALF_TO_F32) && (is_half || is_bf16)) {
    std::vector<float> w(t.half_val.size());
    for (size_t i = 0; i < w.size(); ++i) {
      uint16_t h = static_cast<uint16_t>(static_cast<uint32_t>(t.half_val[i]) & 0xFFFFu);
      w[i] = is_half ? f16_bits_to_f32(h) : bf16_bits_to_f32(h);
    }
    put(13, TFS_T_F32, w.data(), w.size(), w.size() * 4);
  } else {
    put(13, TFS_T_I32, t.half_val.data(), t.half_val.size(), t.half_val.size() * 4);
  }
  put(16, TFS_T_U32, t.uint32_val.data(), t.uint32_val.size(), t.uint32_val.size() * 4);
  put(17, TFS_T_U64, t.uint64_val.data(), t.uint64_val.size(), t.uint64_val.size() * 8);
  d.tensors.push_back(o);
  return true;
}

void emit_model_spec(Decoded& d, tfs_msg_out_t& mo, const std::optional<ModelSpec>& ms) {
  mo.has_model_spec = ms ? 1 : 0;
  if (!ms) return;
  mo.model_name = add_string(d, ms->name);
  mo.signature_name = add_string(d, ms->signature_name);
  mo.version_label = add_string(d, ms->version_label);
  mo.version_choice = ms->version_choice;
  mo.version = ms->version;
}

void decode_one(Decoded& d, uint32_t index, const uint8_t* p, uint64_t n, int32_t kind, uint32_t flags) {
  tfs_msg_out_t mo;
  std::memset(&mo, 0, sizeof(mo));
  size_t tensors0 = d.tensors.size(), filters0 = d.filters.size(), strings0 = d.strings.size(), pool0 = d.pool.size();
  DecodeError err{Err::Ok, 0};
  uint64_t skew = 0;
  bool ok = true;
  if (flags & TFS_F_GRPC_PREFIX) {  // tonic Streaming: flag byte must be 0, then u32 BE length (A.8)
    if (n < 5) {
      ok = false;
      err = {Err::BufferUnderflow, 0};
    } else if (p[0] != 0) {
      ok = false;
      err = {Err::GrpcFlag, 0};
    } else {
      uint64_t l = (uint64_t(p[1]) << 24) | (uint64_t(p[2]) << 16) | (uint64_t(p[3]) << 8) | p[4];
      if (l != n - 5) {
        ok = false;
        err = {l > n - 5 ? Err::BufferUnderflow : Err::LengthExceeded, 1};
      }
      p += 5;
      n -= 5;
      skew = 5;
    }
  }
  mo.first_tensor = static_cast<uint32_t>(tensors0);
  mo.first_filter = static_cast<uint32_t>(filters0);
  if (ok) {
    if (kind == TFS_MSG_TENSOR_PROTO) {
      TensorProto t;
      ok = decode(p, n, &t, &err);
      if (ok && !emit_tensor(d, index, "", t, flags)) { ok = false; err = {Err::Unsupported, 0}; }
    } else if (kind == TFS_MSG_PREDICT_REQUEST) {
      PredictRequest r;
      ok = decode(p, n, &r, &err);
      if (ok) {
        emit_model_spec(d, mo, r.model_spec);
        for (auto& kv : r.inputs)
          if (ok && !emit_tensor(d, index, kv.first, kv.second, flags)) { ok = false; err = {Err::Unsupported, 0}; }
        for (auto& s : r.output_filter) d.filters.push_back(add_string(d, s));
      }
    } else if (kind == TFS_MSG_METADATA_REQUEST) {
      GetModelMetadataRequest r;
      ok = decode(p, n, &r, &err);
      if (ok) {
        emit_model_spec(d, mo, r.model_spec);
        for (auto& s : r.metadata_field) d.filters.push_back(add_string(d, s));
      }
    } else if (kind == TFS_MSG_METADATA_RESPONSE) {
      // every metadata entry as the library hands it back: a record with the map key and string_val = [type_url, value]
      GetModelMetadataResponse r;
      ok = decode(p, n, &r, &err);
      if (ok) {
        emit_model_spec(d, mo, r.model_spec);
        for (auto& kv : r.metadata) {
          TensorProto t;
          t.string_val.emplace_back(kv.second.type_url.begin(), kv.second.type_url.end());
          t.string_val.push_back(kv.second.value);
          if (ok && !emit_tensor(d, index, kv.first, t, flags)) { ok = false; err = {Err::Unsupported, 0}; }
        }
      }
    } else {
      PredictResponse r;
      ok = decode(p, n, &r, &err);
      if (ok) {
        emit_model_spec(d, mo, r.model_spec);
        for (auto& kv : r.outputs)
          if (ok && !emit_tensor(d, index, kv.first, kv.second, flags)) { ok = false; err = {Err::Unsupported, 0}; }
      }
    }
  }
  if (!ok) {  // a failed message exposes nothing
    d.tensors.resize(tensors0);
    d.filters.resize(filters0);
    d.strings.resize(strings0);
    d.pool.resize(pool0);
    std::memset(&mo, 0, sizeof(mo));
    mo.first_tensor = static_cast<uint32_t>(tensors0);
    mo.first_filter = static_cast<uint32_t>(filters0);
    mo.status = static_cast<int32_t>(err.code);
    mo.err_offset = err.offset + skew;
    d.n_failed++;
  }
  mo.n_tensors = static_cast<uint32_t>(d.tensors.size() - tensors0);
  mo.n_filters = static_cast<uint32_t>(d.filters.size() - filters0);
  d.msgs.push_back(mo);
}

// ------------------------------------------------------------------------------------------
// timed baseline state
// ------------------------------------------------------------------------------------------
struct Prepared {
  std::vector<AnyMsg> msgs;                    // the prost structs the Rust client would hold
  std::vector<std::vector<uint8_t>> wire;      // one BytesMut per message (filled by an encode run)
  int32_t kind = 0;
};

template <class F>
void parallel_for(uint32_t n, int threads, F f) {
  if (threads <= 1) {
    for (uint32_t i = 0; i < n; ++i) f(i);
    return;
  }
  std::atomic<uint32_t> next{0};
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&] {
      for (;;) {
        uint32_t i = next.fetch_add(1);
        if (i >= n) break;
        f(i);
      }
    });
  for (auto& th : pool) th.join();
}

}  // namespace

extern "C" {

typedef struct oracle_decoded oracle_decoded_t;
typedef struct oracle_prepared oracle_prepared_t;

// prost `encoded_len` of every message of the batch.
int32_t oracle_encode_size(const tfs_batch_t* b, uint64_t* msg_len_out) {
  if (!b || !msg_len_out) return TFS_ERR_INVALID_ARG;
  for (uint32_t i = 0; i < b->n_msgs; ++i) {
    AnyMsg m;
    if (!build_msg(*b, b->msgs[i], &m)) return TFS_ERR_INVALID_ARG;
    msg_len_out[i] = m.encoded_len();
  }
  return TFS_OK;
}

// Encode every message (tonic ProstEncoder::encode), back to back in dst; with TFS_F_GRPC_PREFIX
// each is preceded by tonic's 5-byte frame header (00, u32 big-endian length).
int32_t oracle_encode(const tfs_batch_t* b, uint8_t* dst, uint64_t cap, uint32_t flags, uint64_t* off_out,
                      uint64_t* len_out, uint64_t* total_out) {
  if (!b || (!dst && cap)) return TFS_ERR_INVALID_ARG;
  uint64_t cursor = 0;
  for (uint32_t i = 0; i < b->n_msgs; ++i) {
    AnyMsg m;
    if (!build_msg(*b, b->msgs[i], &m)) return TFS_ERR_INVALID_ARG;
    std::vector<uint8_t> buf;
    m.prost_encode(buf);
    uint64_t need = buf.size() + ((flags & TFS_F_GRPC_PREFIX) ? 5 : 0);
    if (cursor + need > cap) return TFS_ERR_NOMEM;
    if (off_out) off_out[i] = cursor;
    if (len_out) len_out[i] = buf.size();
    if (flags & TFS_F_GRPC_PREFIX) {
      uint32_t l = static_cast<uint32_t>(buf.size());
      dst[cursor + 0] = 0;
      dst[cursor + 1] = static_cast<uint8_t>(l >> 24);
      dst[cursor + 2] = static_cast<uint8_t>(l >> 16);
      dst[cursor + 3] = static_cast<uint8_t>(l >> 8);
      dst[cursor + 4] = static_cast<uint8_t>(l);
      cursor += 5;
    }
    if (!buf.empty()) std::memcpy(dst + cursor, buf.data(), buf.size());
    cursor += buf.size();
  }
  if (total_out) *total_out = cursor;
  return TFS_OK;
}

// Decode n messages (tonic ProstDecoder::decode -> Message::decode) into tfs_gpu.h's result layout.
int32_t oracle_decode(const uint8_t* wire, const uint64_t* msg_off, const uint64_t* msg_len, uint32_t n_msgs,
                      int32_t msg_kind, uint32_t flags, oracle_decoded_t** out) {
  if (!out || (n_msgs && (!msg_off || !msg_len))) return TFS_ERR_INVALID_ARG;
  static const uint8_t kEmpty[1] = {0};
  if (!wire) wire = kEmpty;  // a batch of zero-length messages
  if ((msg_kind < 0 || msg_kind > 2) && msg_kind != TFS_MSG_METADATA_REQUEST && msg_kind != TFS_MSG_METADATA_RESPONSE) return TFS_ERR_INVALID_ARG;
  Decoded* d = new Decoded();
  for (uint32_t i = 0; i < n_msgs; ++i) decode_one(*d, i, wire + msg_off[i], msg_len[i], msg_kind, flags);
  *out = reinterpret_cast<oracle_decoded_t*>(d);
  return d->n_failed ? TFS_ERR_BATCH : TFS_OK;
}
int32_t oracle_decoded_view(const oracle_decoded_t* h, tfs_decoded_view_t* v, const uint8_t** pool) {
  const Decoded* d = reinterpret_cast<const Decoded*>(h);
  if (!d || !v) return TFS_ERR_INVALID_ARG;
  v->msgs = d->msgs.data();
  v->tensors = d->tensors.data();
  v->filters = d->filters.data();
  v->strings = d->strings.data();
  v->strings_bytes = d->strings.size();
  v->pool_bytes = d->pool.size();
  v->n_msgs = static_cast<uint32_t>(d->msgs.size());
  v->n_tensors = static_cast<uint32_t>(d->tensors.size());
  v->n_filters = static_cast<uint32_t>(d->filters.size());
  v->n_failed = d->n_failed;
  if (pool) *pool = d->pool.data();
  return TFS_OK;
}
void oracle_decoded_free(oracle_decoded_t* h) { delete reinterpret_cast<Decoded*>(h); }

// ---- timed CPU baseline (bench.py cpu_baseline / --impl reference) --------------------------------
// Build the prost structs once (the client holds them before it calls predict; untimed).
int32_t oracle_prepare(const tfs_batch_t* b, oracle_prepared_t** out) {
  if (!b || !out) return TFS_ERR_INVALID_ARG;
  auto p = std::make_unique<Prepared>();
  p->msgs.resize(b->n_msgs);
  for (uint32_t i = 0; i < b->n_msgs; ++i)
    if (!build_msg(*b, b->msgs[i], &p->msgs[i])) return TFS_ERR_INVALID_ARG;
  p->kind = b->n_msgs ? b->msgs[0].kind : 0;
  p->wire.resize(b->n_msgs);
  *out = reinterpret_cast<oracle_prepared_t*>(p.release());
  return TFS_OK;
}
void oracle_prepared_free(oracle_prepared_t* h) { delete reinterpret_cast<Prepared*>(h); }

// One pass of ProstEncoder::encode over every message, `threads` host threads (one message at a
// time per thread, each into a fresh buffer).  Returns seconds; *wire_bytes = bytes produced.
double oracle_run_encode(oracle_prepared_t* h, int32_t threads, uint64_t* wire_bytes) {
  Prepared* p = reinterpret_cast<Prepared*>(h);
  auto t0 = std::chrono::steady_clock::now();
  parallel_for(static_cast<uint32_t>(p->msgs.size()), threads, [&](uint32_t i) {
    std::vector<uint8_t> buf;  // tonic: a BytesMut that is reserve()d to encoded_len
    p->msgs[i].prost_encode(buf);
    p->wire[i].swap(buf);
  });
  auto t1 = std::chrono::steady_clock::now();
  uint64_t total = 0;
  for (auto& w : p->wire) total += w.size();
  if (wire_bytes) *wire_bytes = total;
  return std::chrono::duration<double>(t1 - t0).count();
}
// One pass of ProstDecoder::decode over the wire produced by the last oracle_run_encode.
// Returns seconds; *checksum folds a few decoded values so the work cannot be elided.
double oracle_run_decode(oracle_prepared_t* h, int32_t threads, uint64_t* wire_bytes, uint64_t* checksum) {
  Prepared* p = reinterpret_cast<Prepared*>(h);
  std::atomic<uint64_t> sum{0};
  auto t0 = std::chrono::steady_clock::now();
  parallel_for(static_cast<uint32_t>(p->wire.size()), threads, [&](uint32_t i) {
    const auto& w = p->wire[i];
    uint64_t s = 0;
    auto fold = [&](const TensorProto& t) {
      s += t.float_val.size() + t.int64_val.size() + t.half_val.size() + t.tensor_content.size() + t.double_val.size() +
           t.int_val.size() + t.bool_val.size() + t.uint32_val.size() + t.uint64_val.size();
      if (!t.float_val.empty()) { uint32_t x; std::memcpy(&x, &t.float_val.back(), 4); s += x; }
      if (!t.int64_val.empty()) s += static_cast<uint64_t>(t.int64_val.back());
      if (!t.half_val.empty()) s += static_cast<uint64_t>(t.half_val.back());
      if (!t.tensor_content.empty()) s += t.tensor_content.back();
    };
    DecodeError e;
    if (p->kind == TFS_MSG_TENSOR_PROTO) {
      TensorProto t;
      if (decode(w.data(), w.size(), &t, &e)) fold(t);
    } else if (p->kind == TFS_MSG_PREDICT_REQUEST) {
      PredictRequest r;
      if (decode(w.data(), w.size(), &r, &e))
        for (auto& kv : r.inputs) fold(kv.second);
    } else {
      PredictResponse r;
      if (decode(w.data(), w.size(), &r, &e))
        for (auto& kv : r.outputs) fold(kv.second);
    }
    sum.fetch_add(s);
  });
  auto t1 = std::chrono::steady_clock::now();
  uint64_t total = 0;
  for (auto& w : p->wire) total += w.size();
  if (wire_bytes) *wire_bytes = total;
  if (checksum) *checksum = sum.load();
  return std::chrono::duration<double>(t1 - t0).count();
}

// ---- small primitives exposed for known-answer tests ----
uint64_t oracle_encoded_len_varint(uint64_t v) { return encoded_len_varint(v); }
uint32_t oracle_f32_to_f16(float f) { return f32_to_f16_bits(f); }
uint32_t oracle_f32_to_bf16(float f) { return f32_to_bf16_bits(f); }
float oracle_f16_to_f32(uint32_t h) { return f16_bits_to_f32(static_cast<uint16_t>(h)); }
float oracle_bf16_to_f32(uint32_t h) { return bf16_bits_to_f32(static_cast<uint16_t>(h)); }
int32_t oracle_is_utf8(const uint8_t* p, uint64_t n) { return is_utf8(p, n) ? 1 : 0; }

}  // extern "C"
