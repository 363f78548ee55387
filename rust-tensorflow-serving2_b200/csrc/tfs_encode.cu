// tfs_encode.cu — batched TensorProto / PredictRequest / PredictResponse ENCODE on sm_100a.
//
// Replaces, on the reference path `self.prediction_client.predict(request)` (src/lib.rs:265), what tonic's
// ProstEncoder does with the message built at src/lib.rs:249-263:
//   prost Message::encoded_len  -> K1 k_varint_tiles<SIZE> + K2 k_field_scan + K6a k_msg_layout   (sized ONCE;
//                                  prost re-walks a nested packed varint field 5 times)
//   prost Message::encode_raw   -> K6b k_msg_scan + k_write_headers (framing bytes, device side)
//                                  K3 k_varint_tiles<SCATTER>       (int32/int64/uint32/uint64/half_val bodies)
//                                  K5 k_copy (tfs_copy.cu)          (float/double/complex/bool/tensor_content)
// Nothing here is a contraction: integer / byte work bound by HBM bandwidth, no tensor cores.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <system_error>
#include <thread>

#include "tfs_device.cuh"
#include "tfs_ragged.cuh"
#include "tfs_varint_enc.cuh"

namespace tfs {

namespace {

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
  return v;
}

// =====================================================================================================
// K2: per varint field, exclusive scan of its tile byte counts -> tile offsets + body length.
// One warp per field; each lane takes 4 consecutive tiles per step.
// =====================================================================================================
__global__ void __launch_bounds__(256) k_field_scan(EncTables T) {
  const uint32_t vf = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (vf >= T.n_vfields) return;
  const uint32_t fi = T.vfield_index[vf];
  const uint32_t t0 = T.vfield_first_tile[vf];
  const uint32_t nt = T.vfield_first_tile[vf + 1] - t0;
  uint64_t carry = 0;
  for (uint32_t base = 0; base < nt; base += 128) {
    const uint32_t i0 = base + lane_id() * 4u;
    uint32_t x[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) x[j] = (i0 + j < nt) ? T.tile_bytes[t0 + i0 + j] : 0u;
    const uint32_t s = x[0] + x[1] + x[2] + x[3];
    const uint32_t incl = warp_inclusive_scan(s);
    uint64_t o = carry + (incl - s);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (i0 + j < nt) T.tile_off[t0 + i0 + j] = o;
      o += x[j];
    }
    carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
  }
  if (lane_id() == 0) T.fields[fi].body_len = carry;
}

// =====================================================================================================
// K6: message framing.  One traversal template emits (or just counts) every byte of a message that
// is not a bulk payload body, in prost's order: ascending field number, proto3 defaults skipped,
// Option<Message> iff Some, oneof at its lowest tag, map entries in caller order.
// =====================================================================================================
__device__ __forceinline__ uint32_t dim_body_len(int64_t size) { return size != 0 ? 1u + varint_len(static_cast<uint64_t>(size)) : 0u; }

__device__ __forceinline__ uint32_t shape_body_len(const EncTensor& t) {
  uint32_t n = 0;
  for (uint32_t i = 0; i < t.n_dims; ++i) n += 2u + dim_body_len(t.dims[i]);  // key + 1-byte length + Dim
  if (t.unknown_rank) n += 2u;
  return n;
}

__device__ __forceinline__ uint64_t str_field_len(uint32_t len) { return 1u + varint_len(len) + static_cast<uint64_t>(len); }

__device__ __forceinline__ uint64_t spec_body_len(const EncMsg& m) {
  uint64_t n = 0;
  if (m.model_name.len) n += str_field_len(m.model_name.len);
  if (m.version_choice == TFS_VERSION_NUMBER) n += 2u + (m.version != 0 ? 1u + varint_len(static_cast<uint64_t>(m.version)) : 0u);
  else if (m.version_choice == TFS_VERSION_LABEL) n += str_field_len(m.version_label.len);
  if (m.signature_name.len) n += str_field_len(m.signature_name.len);
  return n;
}

// Several descriptors of the same kind in one tensor are ONE field on the wire: their bodies concatenated in descriptor
// order under a single key + length (a [B,H,W,3] tensor assembled from B separately stored images).
// Returns the summed body length of the group that starts at field i and the index one past its last member.
__device__ __forceinline__ uint64_t group_body_len(const EncTables& T, const EncTensor& t, uint32_t i, uint32_t* end) {
  const int32_t kind = T.fields[t.first_field + i].kind;
  uint64_t n = 0;
  uint32_t j = i;
  for (; j < t.n_fields && T.fields[t.first_field + j].kind == kind; ++j)
    if (T.fields[t.first_field + j].n_elems != 0) n += T.fields[t.first_field + j].body_len;
  *end = j;
  return n;
}

__device__ __forceinline__ uint64_t tensor_body_len(const EncTables& T, const EncTensor& t) {
  uint64_t n = 0;
  if (t.dtype != 0) n += 1u + varint_len(static_cast<uint64_t>(static_cast<int64_t>(t.dtype)));
  if (t.has_shape) {
    uint32_t s = shape_body_len(t);
    n += 1u + varint_len(s) + s;
  }
  if (t.version_number != 0) n += 1u + varint_len(static_cast<uint64_t>(static_cast<int64_t>(t.version_number)));
  for (uint32_t i = 0; i < t.n_fields;) {
    const EncField& f = T.fields[t.first_field + i];
    uint32_t end;
    const uint64_t body = group_body_len(T, t, i, &end);
    i = end;
    if (body == 0) continue;  // (a packed field or a record area is never empty when it has elements)
    // repeated bytes: the record area carries a key and a length per element, nothing in front of it
    n += f.is_varint == FC_RAGGED ? body : key_len(f.kind) + varint_len(body) + body;
  }
  return n;
}

template <class Sink>
__device__ void emit_string(Sink& s, uint32_t field, const uint8_t* strings, StrRef r) {
  put_key(s, field, 2);
  put_varint(s, r.len);
  s.put_bytes(strings + r.off, r.len);
}

template <class Sink>
__device__ void emit_model_spec(Sink& s, const EncTables& T, const EncMsg& m, uint32_t field, uint64_t spec_len) {
  put_key(s, field, 2);
  put_varint(s, spec_len);
  if (m.model_name.len) emit_string(s, 1, T.strings, m.model_name);
  if (m.version_choice == TFS_VERSION_NUMBER) {
    put_key(s, 2, 2);
    if (m.version != 0) {
      s.put(static_cast<uint8_t>(1u + varint_len(static_cast<uint64_t>(m.version))));
      put_key(s, 1, 0);
      put_varint(s, static_cast<uint64_t>(m.version));
    } else {
      s.put(0);
    }
  } else if (m.version_choice == TFS_VERSION_LABEL) {
    emit_string(s, 4, T.strings, m.version_label);
  }
  if (m.signature_name.len) emit_string(s, 3, T.strings, m.signature_name);
}

// Emits the TensorProto body; for every payload field the body bytes are skipped (written by K3 / K5)
// and `on_body(field_index, offset_of_body)` is called.
template <class Sink, class OnBody>
__device__ void emit_tensor_body(Sink& s, const EncTables& T, const EncTensor& t, OnBody on_body) {
  if (t.dtype != 0) {
    put_key(s, 1, 0);
    put_varint(s, static_cast<uint64_t>(static_cast<int64_t>(t.dtype)));
  }
  if (t.has_shape) {
    put_key(s, 2, 2);
    put_varint(s, shape_body_len(t));
    for (uint32_t i = 0; i < t.n_dims; ++i) {
      put_key(s, 2, 2);
      s.put(static_cast<uint8_t>(dim_body_len(t.dims[i])));
      if (t.dims[i] != 0) {
        put_key(s, 1, 0);
        put_varint(s, static_cast<uint64_t>(t.dims[i]));
      }
    }
    if (t.unknown_rank) {
      put_key(s, 3, 0);
      s.put(1);
    }
  }
  if (t.version_number != 0) {
    put_key(s, 3, 0);
    put_varint(s, static_cast<uint64_t>(static_cast<int64_t>(t.version_number)));
  }
  uint32_t s_group_open = 0;  // a key + length has been written for the current run of same-kind descriptors
  for (uint32_t i = 0; i < t.n_fields; ++i) {
    const uint32_t fi = t.first_field + i;
    const EncField& f = T.fields[fi];
    if (f.n_elems == 0) {
      if (i + 1 >= t.n_fields || T.fields[fi + 1].kind != f.kind) s_group_open = 0;
      continue;
    }
    if (s_group_open == 0) {  // first non-empty descriptor of its kind: one key + length for the whole run
      uint32_t end;
      const uint64_t body = group_body_len(T, t, i, &end);
      if (f.is_varint != FC_RAGGED) {
        put_key(s, static_cast<uint32_t>(f.kind), 2);
        put_varint(s, body);
      }
      s_group_open = 1;
    }
    on_body(fi, s.pos);
    s.skip(f.body_len);
    if (i + 1 >= t.n_fields || T.fields[fi + 1].kind != f.kind) s_group_open = 0;
  }
}

// ---- Input { example_list { examples: [Example { features { feature: map<string, Feature> } }] } } (TFS_MSG_INPUT) and
//      the request around it (TFS_MSG_CLASSIFY_REQUEST): build_input, src/lib.rs:413-431.  Every feature is one EncTensor
//      (version_number = index of its Example, ascending), EncMsg::n_filters = number of Examples.
__device__ __forceinline__ bool is_input_kind(int32_t k) { return k == TFS_MSG_INPUT || k == TFS_MSG_CLASSIFY_REQUEST; }

// body of the BytesList / FloatList / Int64List message of a feature
__device__ __forceinline__ uint64_t feature_list_len(const EncTables& T, const EncTensor& t) {
  const EncField& f = T.fields[t.first_field];
  if (f.n_elems == 0) return 0;
  return f.is_varint == FC_RAGGED ? f.body_len : 1u + varint_len(f.body_len) + f.body_len;
}
// body of the Feature message: the oneof member is written even when its list is empty; kind None -> nothing
__device__ __forceinline__ uint64_t feature_body_len(const EncTables& T, const EncTensor& t) {
  if (t.n_fields == 0) return 0;
  const uint64_t l = feature_list_len(T, t);
  return 1u + varint_len(l) + l;
}
// length of the Features map of the Example that starts at feature `ti`; *next = first feature of the next Example
__device__ __forceinline__ uint64_t features_len_from(const EncTables& T, uint32_t ti, uint32_t te, int32_t example, uint32_t* next) {
  uint64_t n = 0;
  while (ti < te && T.feat_ex[ti] == example) {  // compact arrays: one cache line serves 16 features of this serial loop
    n += T.feat_rec[ti];
    ++ti;
  }
  *next = ti;
  return n;
}
__device__ uint64_t example_list_len(const EncTables& T, const EncMsg& m) {
  const uint32_t te = m.first_tensor + m.n_tensors;
  uint32_t ti = m.first_tensor;
  uint64_t n = 0;
  for (uint32_t e = 0; e < m.n_filters; ++e) {
    const uint64_t fl = features_len_from(T, ti, te, static_cast<int32_t>(e), &ti);
    const uint64_t xl = 1u + varint_len(fl) + fl;  // Example { features: Some(..) }
    n += 1u + varint_len(xl) + xl;
  }
  return n;
}
template <class Sink, class OnBody>
__device__ void emit_input(Sink& s, const EncTables& T, uint32_t mi, OnBody on_body) {
  const EncMsg& m = T.msgs[mi];
  const uint64_t list_len = example_list_len(T, m);
  if (m.kind == TFS_MSG_CLASSIFY_REQUEST) {
    if (m.has_model_spec) emit_model_spec(s, T, m, 1, T.spec_len[mi]);
    put_key(s, 2, 2);  // input = 2
    put_varint(s, 1u + varint_len(list_len) + list_len);
  }
  put_key(s, 1, 2);  // oneof kind: example_list = 1
  put_varint(s, list_len);
  const uint32_t te = m.first_tensor + m.n_tensors;
  uint32_t ti = m.first_tensor;
  for (uint32_t e = 0; e < m.n_filters; ++e) {
    uint32_t next;
    const uint64_t fl = features_len_from(T, ti, te, static_cast<int32_t>(e), &next);
    put_key(s, 1, 2);  // examples = 1
    put_varint(s, 1u + varint_len(fl) + fl);
    put_key(s, 1, 2);  // features = 1
    put_varint(s, fl);
    for (; ti < next; ++ti) {  // the entries themselves are written by k_feature_write, one thread per feature
      s.note(&T.entry_off[ti]);
      s.skip(T.feat_rec[ti]);
    }
  }
  (void)on_body;
}

// K6f: nested lengths of every feature of an Input (one thread per feature; messages of other kinds are skipped)
__global__ void __launch_bounds__(256) k_feature_layout(EncTables T) {
  const uint32_t ti = blockIdx.x * blockDim.x + threadIdx.x;
  if (ti >= T.n_tensors) return;
  const EncTensor& t = T.tensors[ti];
  if (!is_input_kind(T.msgs[t._pad].kind)) return;  // _pad = index of the message the descriptor belongs to
  const uint64_t fl = feature_body_len(T, t);
  T.tensor_len[ti] = fl;
  const uint64_t el = (t.name.len ? str_field_len(t.name.len) : 0) + (t.n_fields ? 1u + varint_len(fl) + fl : 0);
  T.entry_len[ti] = el;
  T.feat_rec[ti] = 1u + varint_len(el) + el;
  T.feat_ex[ti] = t.version_number;
}

// K6g: the map entry of every feature (key, Feature, list header) at the place the per-message pass reserved for it;
// records where the list body goes for K3 / K5 / K8
__global__ void __launch_bounds__(256) k_feature_write(EncTables T, uint8_t* __restrict__ dst) {
  const uint32_t ti = blockIdx.x * blockDim.x + threadIdx.x;
  if (ti >= T.n_tensors || T.result[1] != 0) return;
  const EncTensor& t = T.tensors[ti];
  if (!is_input_kind(T.msgs[t._pad].kind)) return;
  ByteSink s{dst, T.entry_off[ti]};
  put_key(s, 1, 2);  // map entry of `feature`
  put_varint(s, T.entry_len[ti]);
  if (t.name.len) emit_string(s, 1, T.strings, t.name);
  if (t.n_fields == 0) return;  // Feature::default(): the value is left out of the entry
  put_key(s, 2, 2);
  put_varint(s, T.tensor_len[ti]);
  EncField& f = T.fields[t.first_field];
  put_key(s, static_cast<uint32_t>(f.kind), 2);  // bytes_list = 1 / float_list = 2 / int64_list = 3
  put_varint(s, feature_list_len(T, t));
  if (f.n_elems == 0) return;
  if (f.is_varint != FC_RAGGED) {  // packed `value = 1`
    put_key(s, 1, 2);
    put_varint(s, f.body_len);
  }
  f.dst_off = s.pos;
  if (!f.is_varint) T.segs[f.first_tile].dst = dst + s.pos;
}

template <class Sink, class OnBody>
__device__ void emit_message(Sink& s, const EncTables& T, uint32_t mi, OnBody on_body) {
  const EncMsg& m = T.msgs[mi];
  if (is_input_kind(m.kind)) {
    emit_input(s, T, mi, on_body);
    return;
  }
  if (m.kind == TFS_MSG_TENSOR_PROTO) {
    emit_tensor_body(s, T, T.tensors[m.first_tensor], on_body);
    return;
  }
  if (m.kind == TFS_MSG_METADATA_REQUEST) {  // get_model_metadata.proto:16-23: model_spec = 1, repeated string metadata_field = 2
    if (m.has_model_spec) emit_model_spec(s, T, m, 1, T.spec_len[mi]);
    for (uint32_t i = 0; i < m.n_filters; ++i) emit_string(s, 2, T.strings, T.filters[m.first_filter + i]);
    return;
  }
  const uint32_t map_field = m.kind == TFS_MSG_PREDICT_REQUEST ? 2u : 1u;
  const uint32_t spec_field = m.kind == TFS_MSG_PREDICT_REQUEST ? 1u : 2u;
  if (m.has_model_spec && spec_field == 1u) emit_model_spec(s, T, m, 1, T.spec_len[mi]);
  for (uint32_t i = 0; i < m.n_tensors; ++i) {
    const uint32_t ti = m.first_tensor + i;
    const EncTensor& t = T.tensors[ti];
    put_key(s, map_field, 2);
    put_varint(s, T.entry_len[ti]);
    if (t.name.len) emit_string(s, 1, T.strings, t.name);
    if (T.tensor_len[ti] != 0) {  // TensorProto::default() <=> zero encoded length
      put_key(s, 2, 2);
      put_varint(s, T.tensor_len[ti]);
      emit_tensor_body(s, T, t, on_body);
    }
  }
  if (m.has_model_spec && spec_field == 2u) emit_model_spec(s, T, m, 2, T.spec_len[mi]);
  if (m.kind == TFS_MSG_PREDICT_REQUEST)
    for (uint32_t i = 0; i < m.n_filters; ++i) emit_string(s, 3, T.strings, T.filters[m.first_filter + i]);
}

// K6a: nested lengths bottom-up (once), then the message length and the offset of its largest bulk body.
__device__ void msg_layout_body(const EncTables& T, uint32_t mi) {
  const EncMsg& m = T.msgs[mi];
  for (uint32_t i = 0; i < m.n_tensors; ++i) {
    const uint32_t ti = m.first_tensor + i;
    const EncTensor& t = T.tensors[ti];
    if (is_input_kind(m.kind)) break;  // features were laid out by k_feature_layout
    const uint64_t tl = tensor_body_len(T, t);
    T.tensor_len[ti] = tl;
    uint64_t el = 0;
    if (t.name.len) el += str_field_len(t.name.len);
    if (tl != 0) el += 1u + varint_len(tl) + tl;
    T.entry_len[ti] = el;
  }
  T.spec_len[mi] = m.has_model_spec ? spec_body_len(m) : 0;
  CountSink s;
  uint64_t best_len = 0, best_off = 0;
  emit_message(s, T, mi, [&](uint32_t fi, uint64_t off) {
    const EncField& f = T.fields[fi];
    if (!f.is_varint && f.body_len > best_len) {
      best_len = f.body_len;
      best_off = off;
    }
  });
  T.msg_len[mi] = s.pos;
  T.msg_pay_off[mi] = best_off;
}

// K6b: exclusive scan of the message slots -> msg_off; total and overflow flag.  One CTA of NW warps.
template <int NW>
__device__ void msg_scan_body(const EncTables& T, uint32_t flags, uint32_t dst_skew, uint64_t dst_capacity, uint64_t* smem) {
  const uint32_t prefix = (flags & TFS_F_GRPC_PREFIX) ? 5u : 0u;
  uint64_t carry = 0;
  for (uint32_t base = 0; base < T.n_msgs; base += NW * 32) {
    const uint32_t mi = base + threadIdx.x;
    uint64_t slot = 0, pad = 0;
    if (mi < T.n_msgs) {
      const uint64_t len = T.msg_len[mi] + prefix;
      if (flags & TFS_F_ALIGN_PAYLOAD) {
        pad = (16u - ((dst_skew + prefix + T.msg_pay_off[mi]) & 15u)) & 15u;
        slot = (pad + len + 15u) & ~static_cast<uint64_t>(15);
      } else {
        slot = len;
      }
    }
    uint64_t total;
    const uint64_t excl = block_exclusive_scan<uint64_t, NW>(slot, smem, &total);
    if (mi < T.n_msgs) T.msg_off[mi] = carry + excl + pad;
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    T.result[0] = carry;
    T.result[1] = carry > dst_capacity ? 1u : 0u;
  }
}

// K6c: write every framing byte, record where each payload body goes.
__device__ void write_headers_body(const EncTables& T, uint32_t mi, uint8_t* __restrict__ dst, uint32_t flags) {
  ByteSink s{dst, T.msg_off[mi]};
  if (flags & TFS_F_GRPC_PREFIX) {  // tonic encode_client: 0x00, u32 big-endian length
    const uint32_t l = static_cast<uint32_t>(T.msg_len[mi]);
    s.put(0);
    s.put(static_cast<uint8_t>(l >> 24));
    s.put(static_cast<uint8_t>(l >> 16));
    s.put(static_cast<uint8_t>(l >> 8));
    s.put(static_cast<uint8_t>(l));
  }
  emit_message(s, T, mi, [&](uint32_t fi, uint64_t off) {
    EncField& f = T.fields[fi];
    f.dst_off = off;
    if (!f.is_varint) T.segs[f.first_tile].dst = dst + off;
  });
}

__global__ void __launch_bounds__(128) k_msg_layout(EncTables T) {
  const uint32_t mi = blockIdx.x * blockDim.x + threadIdx.x;
  if (mi < T.n_msgs) msg_layout_body(T, mi);
}
__global__ void __launch_bounds__(1024) k_msg_scan(EncTables T, uint32_t flags, uint32_t dst_skew, uint64_t dst_capacity) {
  __shared__ uint64_t smem[33];
  msg_scan_body<32>(T, flags, dst_skew, dst_capacity, smem);
}
__global__ void __launch_bounds__(128) k_write_headers(EncTables T, uint8_t* __restrict__ dst, uint32_t flags) {
  const uint32_t mi = blockIdx.x * blockDim.x + threadIdx.x;
  if (mi >= T.n_msgs || T.result[1] != 0) return;
  write_headers_body(T, mi, dst, flags);
}

// Small batches (<= kSmallBatch messages): layout, scan and framing bytes in ONE launch, one WARP per message.
// The traversal of a message is a serial chain of dependent reads of its descriptor records; out of global memory every
// link costs an L2 round trip (the first version, one thread per message in a single CTA, took 50-70 us for 256
// two-tensor requests).  Here the lanes of the warp first copy the message's records (EncMsg, its tensors, their fields)
// into shared memory, lane 0 walks them there, and the per-tensor / per-message work arrays live there too.  Between
// layout and writing every warp needs the lengths of all earlier messages: the CTAs (at most 32, all resident) meet at a
// counter in global memory.
constexpr uint32_t kSmallBatch = 256;
constexpr uint32_t kFrameWarps = 8;      // messages per CTA
constexpr uint32_t kShTensors = 4;       // a message with more map entries / payload fields than this is walked in global memory
constexpr uint32_t kShFields = 8;
struct FrameShadow {
  EncMsg msg;
  EncTensor tensors[kShTensors];
  EncField fields[kShFields];
  uint64_t tensor_len[kShTensors], entry_len[kShTensors];
  uint64_t spec_len, msg_len, msg_pay_off, msg_off;
};
static_assert(sizeof(EncMsg) % 8 == 0 && sizeof(EncTensor) % 8 == 0 && sizeof(EncField) % 8 == 0, "records are copied as 8-byte words");

__device__ __forceinline__ void warp_copy_words(void* dst, const void* src, uint32_t bytes) {
  unsigned long long* d = static_cast<unsigned long long*>(dst);
  const unsigned long long* s = static_cast<const unsigned long long*>(src);
  for (uint32_t i = lane_id(); i < bytes / 8u; i += 32u) d[i] = s[i];
}

__global__ void __launch_bounds__(kFrameWarps * 32) k_frame_small(EncTables T, uint8_t* __restrict__ dst, uint32_t flags, uint32_t dst_skew,
                                                                  uint64_t dst_capacity, int write) {
  __shared__ FrameShadow shadow[kFrameWarps];
  const uint32_t mi = blockIdx.x * kFrameWarps + warp_id();
  const bool live = mi < T.n_msgs;
  FrameShadow& sh = shadow[warp_id()];
  EncTables L = T;  // this warp's view: the records of ITS message are in shared memory, indexed like the global tables
  bool shadowed = false;
  uint32_t f_lo = 0, f_n = 0;
  if (live) {
    const EncMsg m = T.msgs[mi];
    // eligible: few tensors whose field ranges follow each other (how every builder lays them out)
    bool ok = m.n_tensors <= kShTensors && !is_input_kind(m.kind);
    if (ok && m.n_tensors) {
      f_lo = T.tensors[m.first_tensor].first_field;
      uint32_t next = f_lo;
      for (uint32_t i = 0; i < m.n_tensors && ok; ++i) {
        const EncTensor& t = T.tensors[m.first_tensor + i];
        ok = t.first_field == next;
        next += t.n_fields;
      }
      f_n = next - f_lo;
      ok = ok && f_n <= kShFields;
    }
    if (ok) {
      shadowed = true;
      warp_copy_words(&sh.msg, &T.msgs[mi], sizeof(EncMsg));
      warp_copy_words(sh.tensors, &T.tensors[m.first_tensor], sizeof(EncTensor) * m.n_tensors);
      if (f_n) warp_copy_words(sh.fields, &T.fields[f_lo], sizeof(EncField) * f_n);
      __syncwarp();
      L.msgs = &sh.msg - mi;
      L.tensors = sh.tensors - m.first_tensor;
      L.fields = sh.fields - f_lo;
      L.tensor_len = sh.tensor_len - m.first_tensor;
      L.entry_len = sh.entry_len - m.first_tensor;
    }
    L.spec_len = &sh.spec_len - mi;
    L.msg_len = &sh.msg_len - mi;
    L.msg_pay_off = &sh.msg_pay_off - mi;
    L.msg_off = &sh.msg_off - mi;
    if (lane_id() == 0) {
      msg_layout_body(L, mi);
      T.msg_len[mi] = sh.msg_len;  // for the other warps' prefix and for the host
      T.msg_pay_off[mi] = sh.msg_pay_off;
      __threadfence();
    }
  }
  // ---- every message's length is needed from here on: all CTAs of this (small) grid meet
  __syncthreads();
  if (gridDim.x > 1) {
    if (threadIdx.x == 0) {
      // two counters that live in the context and are zero between launches: the last CTA through clears them again
      atomicAdd(T.ticket, 1u);
      while (atomicAdd(T.ticket, 0u) < gridDim.x) __nanosleep(50);
      __threadfence();
      if (atomicAdd(T.ticket + 1, 1u) == gridDim.x - 1u) {  // everybody has left the loop above
        T.ticket[0] = 0u;
        T.ticket[1] = 0u;
        __threadfence();
      }
    }
    __syncthreads();
  }
  if (!live) return;
  // ---- this message's offset = the slots of all earlier messages; the grand total for the overflow check
  const uint32_t prefix = (flags & TFS_F_GRPC_PREFIX) ? 5u : 0u;
  uint64_t before = 0, all = 0, my_pad = 0;
  for (uint32_t base = 0; base < T.n_msgs; base += 32u) {
    const uint32_t i = base + lane_id();
    uint64_t slot = 0, pad = 0;
    if (i < T.n_msgs) {
      const uint64_t len = __ldcg(&T.msg_len[i]) + prefix;
      if (flags & TFS_F_ALIGN_PAYLOAD) {
        pad = (16u - ((dst_skew + prefix + __ldcg(&T.msg_pay_off[i])) & 15u)) & 15u;
        slot = (pad + len + 15u) & ~static_cast<uint64_t>(15);
      } else {
        slot = len;
      }
    }
    if (i < mi) before += slot;
    if (i == mi) my_pad = pad;
    all += slot;
  }
  before = warp_sum_u64(before);
  all = warp_sum_u64(all);
  my_pad = warp_sum_u64(my_pad);
  if (lane_id() == 0) {
    sh.msg_off = before + my_pad;
    T.msg_off[mi] = sh.msg_off;
    if (mi == 0) {
      T.result[0] = all;
      T.result[1] = all > dst_capacity ? 1u : 0u;
    }
    if (write && all <= dst_capacity) write_headers_body(L, mi, dst, flags);
  }
  __syncwarp();
  // what later kernels read from the field records (body_len, dst_off) goes back to the global table
  if (shadowed && f_n) warp_copy_words(&T.fields[f_lo], sh.fields, sizeof(EncField) * f_n);
}

// =====================================================================================================
// host: validate + compile the descriptors into one upload blob
// =====================================================================================================
// Host table that keeps its storage between batches (no allocation, no page faults on the request path); `n` entries valid.
template <class T>
struct GrowArray {
  using value_type = T;
  std::vector<T> store;
  size_t n = 0;
  void set_size(size_t k) {
    if (store.size() < k) store.resize(k);
    n = k;
  }
  size_t size() const { return n; }
  T* data() { return store.data(); }
  const T* data() const { return store.data(); }
  T& operator[](size_t i) { return store[i]; }
  const T& operator[](size_t i) const { return store[i]; }
  const T& back() const { return store[n - 1]; }
};

struct Compiled {
  // the records sit in ONE block, msgs | tensors | fields (16-byte aligned each): the context's pinned record block when
  // the batch is going to be encoded (uploaded from where it was compiled), `own` when only the bound is wanted
  EncMsg* msgs = nullptr;
  EncTensor* tensors = nullptr;
  EncField* fields = nullptr;
  size_t n_msgs = 0, n_tensors = 0, n_fields = 0, rec_bytes = 0;
  std::vector<uint64_t> own;
  GrowArray<StrRef> filters;
  std::vector<uint8_t> strings;
  GrowArray<uint32_t> vfield_first_tile, vfield_index;                    // (n + 1) tile prefixes / field index per varint field
  GrowArray<uint32_t> rfield_first_tile, rfield_first_out, rfield_index;  // the same for ragged fields (+ output tiles)
  GrowArray<CopySeg> segs;
  GrowArray<uint64_t> field_bound;          // scratch of compile_batch: upper bound of every field's body
  GrowArray<uint8_t> tensor_used, field_used;
  uint64_t n_copy_tiles = 0;
  uint64_t bound = 0;  // upper bound of the encoded bytes
  // true when the host can already tell that every bulk segment will be congruent mod 16 with its destination
  // (TFS_F_ALIGN_PAYLOAD, one bulk field per message, 16-byte aligned plain-byte sources): TMA ring only
  bool all_tma = true;
  bool any_bulk = false;   // some copy segment is >= 4 KiB of plain bytes (a candidate for the TMA ring)
  bool any_small = false;  // some copy segment has <= kSmallSegBytes of output (no tiles: k_copy_small)
  int cast_kind = 0;       // staged cast kernel needed: 1 = widening segments only, 2 = narrowing ones too (launch_copy)
  bool has_input = false;  // some message is an Input / ClassificationRequest: features are framed by their own kernels
  void reset() {  // a new batch in the same storage
    strings.clear();
    n_copy_tiles = bound = 0;
    all_tma = true;
    any_bulk = any_small = has_input = false;
    cast_kind = 0;
  }
};

StrRef add_str(std::vector<uint8_t>& strings, const tfs_str_t& s) {
  StrRef r{static_cast<uint32_t>(strings.size()), s.len};
  if (s.len) strings.insert(strings.end(), reinterpret_cast<const uint8_t*>(s.ptr), reinterpret_cast<const uint8_t*>(s.ptr) + s.len);
  return r;
}

// Which conversions a payload field may ask for (mirrors the ABI comment on tfs_field_desc_t).
bool compile_field(const tfs_field_desc_t& f, EncField* o, uint64_t* max_body) {
  std::memset(o, 0, sizeof(*o));
  o->src = f.data;
  o->n_elems = f.n_elems;
  o->kind = f.kind;
  o->op_param = f.op_param;
  const int32_t st = f.src_type, dt = f.dst_type;
  if (f.n_elems != 0 && f.data == nullptr) return false;
  const bool is_f32_kind = f.kind == TFS_KIND_FLOAT_VAL || f.kind == TFS_KIND_SCOMPLEX_VAL;
  if (f.op != TFS_OP_NONE && !(st == TFS_T_U8 && is_f32_kind)) return false;
  if (f.op < TFS_OP_NONE || f.op > TFS_OP_MUL) return false;
  auto copy_class = [&](int32_t op, uint64_t out_bytes) {
    o->is_varint = 0;
    o->mode = op;
    o->body_len = out_bytes;
    *max_body = out_bytes;
    return true;
  };
  auto varint_class = [&](int32_t mode, uint32_t max_len) {
    // the varint kernels load whole elements: the array has to be aligned to its element size
    if (reinterpret_cast<uintptr_t>(f.data) & (tfs_type_size(st) - 1u)) return false;
    o->is_varint = 1;
    o->mode = mode;
    *max_body = f.n_elems * max_len;
    return true;
  };
  switch (f.kind) {
    case TFS_KIND_FLOAT_VAL:
    case TFS_KIND_SCOMPLEX_VAL:
      if (dt != TFS_T_NONE && dt != TFS_T_F32) return false;
      if (st == TFS_T_F32) return copy_class(CP_BYTES, f.n_elems * 4);
      if (st == TFS_T_U8) return copy_class(f.op == TFS_OP_DIV ? CP_U8_F32_DIV : f.op == TFS_OP_MUL ? CP_U8_F32_MUL : CP_U8_F32, f.n_elems * 4);
      return false;
    case TFS_KIND_DOUBLE_VAL:
    case TFS_KIND_DCOMPLEX_VAL:
      if (st != TFS_T_F64 || (dt != TFS_T_NONE && dt != TFS_T_F64)) return false;
      return copy_class(CP_BYTES, f.n_elems * 8);
    case TFS_KIND_BOOL_VAL:
      if ((st != TFS_T_BOOL && st != TFS_T_U8) || (dt != TFS_T_NONE && dt != TFS_T_BOOL)) return false;
      return copy_class(CP_BOOL, f.n_elems);
    case TFS_KIND_TENSOR_CONTENT: {
      uint32_t w = tfs_type_size(st);
      if (w == 0) return false;
      if (dt == TFS_T_NONE || dt == st) return copy_class(CP_BYTES, f.n_elems * w);
      if (st == TFS_T_F32 && dt == TFS_T_F16) return copy_class(CP_F32_F16, f.n_elems * 2);
      if (st == TFS_T_F32 && dt == TFS_T_BF16) return copy_class(CP_F32_BF16, f.n_elems * 2);
      return false;
    }
    case TFS_KIND_INT_VAL:
      if (dt != TFS_T_NONE && dt != TFS_T_I32) return false;
      switch (st) {
        case TFS_T_I32: return varint_class(VM_I32, 10);
        case TFS_T_I16: return varint_class(VM_I16, 10);
        case TFS_T_I8: return varint_class(VM_I8, 10);
        case TFS_T_U16: return varint_class(VM_U16, 3);
        case TFS_T_U8: return varint_class(VM_U8, 2);
        default: return false;
      }
    case TFS_KIND_INT64_VAL:
      if (st != TFS_T_I64 || (dt != TFS_T_NONE && dt != TFS_T_I64)) return false;
      return varint_class(VM_I64, 10);
    case TFS_KIND_UINT64_VAL:
      if (st != TFS_T_U64 || (dt != TFS_T_NONE && dt != TFS_T_U64)) return false;
      return varint_class(VM_I64, 10);
    case TFS_KIND_UINT32_VAL:
      if (st != TFS_T_U32 || (dt != TFS_T_NONE && dt != TFS_T_U32)) return false;
      return varint_class(VM_U32, 5);
    case TFS_KIND_HALF_VAL:
      if (st == TFS_T_I32 && (dt == TFS_T_NONE || dt == TFS_T_I32)) return varint_class(VM_I32, 10);
      if ((st == TFS_T_U16 || st == TFS_T_F16 || st == TFS_T_BF16) && (dt == TFS_T_NONE || dt == st || dt == TFS_T_I32))
        return varint_class(VM_U16, 3);
      if (st == TFS_T_F32 && dt == TFS_T_F16) return varint_class(VM_F32_F16, 3);
      if (st == TFS_T_F32 && dt == TFS_T_BF16) return varint_class(VM_F32_BF16, 3);
      return false;
    case TFS_KIND_FLOAT_LIST:  // FloatList.value: packed fixed32
      if (st != TFS_T_F32 || (dt != TFS_T_NONE && dt != TFS_T_F32)) return false;
      return copy_class(CP_BYTES, f.n_elems * 4);
    case TFS_KIND_INT64_LIST:  // Int64List.value: packed varint
      if (st != TFS_T_I64 || (dt != TFS_T_NONE && dt != TFS_T_I64)) return false;
      return varint_class(VM_I64, 10);
    case TFS_KIND_BYTES_LIST:  // BytesList.value = 1: the record key byte (1 << 3 | 2) is this kind's own number
    case TFS_KIND_RESOURCE_HANDLE_VAL:  // repeated messages, every element the encoded body of one (tensor.proto:72-76):
    case TFS_KIND_VARIANT_VAL:          // on the wire `key len body` per element, exactly the shape of repeated bytes
    case TFS_KIND_STRING_VAL: {  // repeated bytes, flattened: (n + 1) u64 offsets, then the bytes
      if (st != TFS_T_RAGGED || (dt != TFS_T_NONE && dt != TFS_T_RAGGED) || (reinterpret_cast<uintptr_t>(f.data) & 7u) != 0) return false;
      o->is_varint = FC_RAGGED;
      o->aux = f.aux;
      uint32_t vl = 1;
      for (uint64_t v = f.aux; v >= 0x80; v >>= 7) ++vl;  // no element is longer than all of them together
      *max_body = f.n_elems * (1 + vl) + f.aux;
      return true;
    }
    default: return false;
  }
}

// What compiling a range of messages yields besides the records themselves (which land at the descriptors' own indices).
struct alignas(128) RangeOut {  // (one per worker, side by side in a vector: own cache lines)
  const char* fail = nullptr;  // first descriptor error of the range (TFS_ERR_INVALID_ARG)
  uint64_t bound = 0;
  bool has_input = false;
  std::vector<uint8_t> strings;  // a worker's strings; StrRef offsets are relative to it until the ranges are joined
};

// Messages [m0, m1) -> c->msgs / tensors / fields.  `strings` receives their names (offsets relative to its start).
// Ranges of one batch run on several threads: every record index belongs to one message, the used-marks are atomic.
void compile_range(const tfs_batch_t* b, uint32_t flags, Compiled* c, uint32_t m0, uint32_t m1, std::vector<uint8_t>& strings,
                   uint64_t* field_bound, uint8_t* tensor_used, uint8_t* field_used, RangeOut* r) {
#define RANGE_FAIL(msg) \
  do {                  \
    r->fail = (msg);    \
    return;             \
  } while (0)
  const uint32_t prefix = (flags & TFS_F_GRPC_PREFIX) ? 5u : 0u;
  uint64_t bound_sum = 0;
  bool any_input = false;
  for (uint32_t mi = m0; mi < m1; ++mi) {
    const tfs_msg_desc_t& m = b->msgs[mi];
    EncMsg& o = c->msgs[mi];
    std::memset(&o, 0, sizeof(o));
    if (m.kind < TFS_MSG_TENSOR_PROTO || m.kind > TFS_MSG_METADATA_RESPONSE) RANGE_FAIL("bad message kind");
    if (m.kind == TFS_MSG_METADATA_RESPONSE) RANGE_FAIL("GetModelMetadataResponse is decode only");
    if (m.kind == TFS_MSG_METADATA_REQUEST && m.n_tensors) RANGE_FAIL("GetModelMetadataRequest carries no tensors");
    const bool input_kind = m.kind == TFS_MSG_INPUT || m.kind == TFS_MSG_CLASSIFY_REQUEST;  // n_filters counts Examples there
    if (static_cast<uint64_t>(m.first_tensor) + m.n_tensors > b->n_tensors) RANGE_FAIL("tensor range out of bounds");
    if (!input_kind && static_cast<uint64_t>(m.first_filter) + m.n_filters > b->n_filters) RANGE_FAIL("filter range out of bounds");
    if (m.kind == TFS_MSG_INPUT && m.has_model_spec) RANGE_FAIL("Input has no model_spec");
    if (m.version_choice < TFS_VERSION_NONE || m.version_choice > TFS_VERSION_LABEL) RANGE_FAIL("bad version_choice");
    if (m.kind == TFS_MSG_TENSOR_PROTO && (m.n_tensors != 1 || m.has_model_spec || m.n_filters))
      RANGE_FAIL("a bare TensorProto message has exactly one tensor and nothing else");
    if (m.kind == TFS_MSG_PREDICT_RESPONSE && m.n_filters) RANGE_FAIL("PredictResponse has no output_filter");
    o.kind = m.kind;
    o.has_model_spec = m.has_model_spec ? 1 : 0;
    o.version_choice = m.version_choice;
    o.version = m.version_choice == TFS_VERSION_NUMBER ? m.version : 0;
    o.first_tensor = m.first_tensor;
    o.n_tensors = m.n_tensors;
    o.first_filter = input_kind ? 0 : m.first_filter;
    o.n_filters = m.n_filters;
    uint64_t bound = 64 + prefix + (input_kind ? 24ull * m.n_filters : 0);
    int32_t last_example = 0;
    if (m.has_model_spec) {
      o.model_name = add_str(strings, m.model_name);
      o.signature_name = add_str(strings, m.signature_name);
      if (m.version_choice == TFS_VERSION_LABEL) o.version_label = add_str(strings, m.version_label);
      bound += m.model_name.len + m.signature_name.len + m.version_label.len + 48;
    }
    for (uint32_t i = 0; i < m.n_filters && !input_kind; ++i) bound += b->filters[m.first_filter + i].len + 8;
    for (uint32_t i = 0; i < m.n_tensors; ++i) {
      const uint32_t ti = m.first_tensor + i;
      if (__atomic_fetch_add(&tensor_used[ti], 1, __ATOMIC_RELAXED)) RANGE_FAIL("a tensor descriptor is referenced twice");
      const tfs_tensor_desc_t& t = b->tensors[ti];
      EncTensor& ot = c->tensors[ti];
      std::memset(&ot, 0, sizeof(ot));
      if (t.n_dims > TFS_MAX_DIMS) RANGE_FAIL("rank exceeds TFS_MAX_DIMS");
      if (!t.has_shape && (t.n_dims || t.unknown_rank)) RANGE_FAIL("dims without has_shape");
      if (static_cast<uint64_t>(t.first_field) + t.n_fields > b->n_fields) RANGE_FAIL("field range out of bounds");
      if (input_kind) {  // a Feature: no dtype / shape, at most one list, Example index ascending
        if (t.dtype || t.has_shape || t.n_fields > 1) RANGE_FAIL("a feature has no dtype / shape and at most one list");
        if (t.version_number < last_example || static_cast<uint32_t>(t.version_number) >= m.n_filters)
          RANGE_FAIL("feature: Example index not ascending or out of range");
        last_example = t.version_number;
      }
      ot.name = m.kind == TFS_MSG_TENSOR_PROTO ? StrRef{0, 0} : add_str(strings, t.name);
      ot._pad = mi;  // the message this descriptor belongs to
      any_input |= input_kind;
      ot.dtype = t.dtype;
      ot.version_number = t.version_number;
      ot.has_shape = t.has_shape ? 1 : 0;
      ot.unknown_rank = t.unknown_rank ? 1 : 0;
      ot.n_dims = t.n_dims;
      ot.first_field = t.first_field;
      ot.n_fields = t.n_fields;
      for (uint32_t d = 0; d < t.n_dims; ++d) ot.dims[d] = t.dims[d];
      bound += t.name.len + 64 + 13ull * t.n_dims;
      uint32_t seen = 0;
      for (uint32_t k = 0; k < t.n_fields; ++k) {
        const uint32_t fi = t.first_field + k;
        if (__atomic_fetch_add(&field_used[fi], 1, __ATOMIC_RELAXED)) RANGE_FAIL("a field descriptor is referenced twice");
        const int32_t fk = b->fields[fi].kind;
        if (input_kind != (fk >= TFS_KIND_BYTES_LIST && fk <= TFS_KIND_INT64_LIST))
          RANGE_FAIL("list kinds belong to features, *_val kinds to tensors");
        const int32_t slot = input_kind ? 0 : tfs_slot_of_kind(fk);
        // a kind may come more than once in a TENSOR: the arrays are concatenated in descriptor order (batch assembly)
        if (slot < 0 || (input_kind && ((seen >> slot) & 1u))) RANGE_FAIL("unknown payload kind, or two lists in one feature");
        seen |= 1u << slot;
        if (!compile_field(b->fields[fi], &c->fields[fi], &field_bound[fi])) RANGE_FAIL("payload field: conversion not allowed by the ABI");
        bound += field_bound[fi] + 16;
      }
      // prost emits in ascending field number: order this tensor's fields by kind
      if (t.n_fields > 1)  // (std::stable_sort allocates a scratch buffer even for one element)
        std::stable_sort(c->fields + t.first_field, c->fields + t.first_field + t.n_fields,
                         [](const EncField& x, const EncField& y) { return x.kind < y.kind; });
    }
    bound_sum += (flags & TFS_F_ALIGN_PAYLOAD) ? align_up(bound + 16, 16) : bound;
  }
  r->bound = bound_sum;
  r->has_input = any_input;
#undef RANGE_FAIL
}

// fn(k) for k in [0, n) on n - 1 short-lived threads and the calling one; a thread that cannot be started is run inline
template <class F>
void run_ranges(uint32_t n, F fn) {
  std::vector<std::thread> crew;
  crew.reserve(n);
  for (uint32_t k = 1; k < n; ++k) {
    try {
      crew.emplace_back([&fn, k] { fn(k); });
    } catch (const std::system_error&) {
      fn(k);
    }
  }
  fn(0);
  for (std::thread& t : crew) t.join();
}

// a batch of tiny messages is host work: above this many messages the descriptors are compiled on
// several threads
constexpr uint32_t kParallelCompileMsgs = 8192;
constexpr uint32_t kCompileThreadsMax = 8;

tfs_status_t compile_batch(tfs_ctx* ctx, const tfs_batch_t* b, uint32_t flags, Compiled* c, bool pinned_records = false) {
  if (!b) return set_err(ctx, TFS_ERR_INVALID_ARG, "batch is null");
  if ((b->n_msgs && !b->msgs) || (b->n_tensors && !b->tensors) || (b->n_fields && !b->fields) || (b->n_filters && !b->filters))
    return set_err(ctx, TFS_ERR_INVALID_ARG, "batch array pointer is null");
  HostTrace tr(ctx, "  compile");
  c->n_msgs = b->n_msgs;
  c->n_tensors = b->n_tensors;
  c->n_fields = b->n_fields;
  const size_t o_t = align_up(sizeof(EncMsg) * c->n_msgs, 16), o_f = align_up(o_t + sizeof(EncTensor) * c->n_tensors, 16);
  c->rec_bytes = align_up(o_f + sizeof(EncField) * c->n_fields, 16);
  uint8_t* rec = nullptr;
  if (pinned_records) {  // no allocation, no page faults, and no copy before the upload (a batch of 10^6 requests: 300 MB)
    TFS_TRY(pinned_reserve(ctx, &ctx->h_rec, &ctx->h_rec_cap, c->rec_bytes + 16));
    rec = static_cast<uint8_t*>(ctx->h_rec);
  } else {
    c->own.resize(c->rec_bytes / 8 + 2);
    rec = reinterpret_cast<uint8_t*>(c->own.data());
  }
  c->msgs = reinterpret_cast<EncMsg*>(rec);
  c->tensors = reinterpret_cast<EncTensor*>(rec + o_t);
  c->fields = reinterpret_cast<EncField*>(rec + o_f);
  c->reset();
  c->filters.set_size(b->n_filters);
  c->field_bound.set_size(b->n_fields);
  c->tensor_used.set_size(b->n_tensors);
  c->field_used.set_size(b->n_fields);
  uint64_t* const field_bound = c->field_bound.data();
  uint8_t* const tensor_used = c->tensor_used.data();
  uint8_t* const field_used = c->field_used.data();
  if (b->n_fields) std::memset(field_bound, 0, 8ull * b->n_fields);
  if (b->n_fields) std::memset(field_used, 0, b->n_fields);
  if (b->n_tensors) std::memset(tensor_used, 0, b->n_tensors);
  for (uint32_t i = 0; i < b->n_filters; ++i) c->filters[i] = add_str(c->strings, b->filters[i]);
  uint32_t n_workers = 1;
  tr.mark("alloc");
  // (the work is per descriptor: 8 192 one-tensor requests and 64 ClassificationRequests of 192 features weigh the same)
  const uint64_t n_desc = static_cast<uint64_t>(b->n_msgs) + b->n_tensors + b->n_fields;
  // starting a thread costs ~0.1 ms, a descriptor ~30 ns: 64 ClassificationRequests of 192 features (25 k descriptors,
  // 0.4 ms) were measured SLOWER on two threads (0.6 ms), so the work has to be a few milliseconds before it is split
  if (n_desc >= 12ull * kParallelCompileMsgs && b->n_msgs >= 2)
    n_workers = std::min<uint32_t>({kCompileThreadsMax, std::max(1u, std::thread::hardware_concurrency() / 2),
                                    static_cast<uint32_t>(n_desc / (6ull * kParallelCompileMsgs)), b->n_msgs});
  if (const char* e = std::getenv("TFS_COMPILE_THREADS")) n_workers = std::min<uint32_t>(std::max(1, std::atoi(e)), std::max(1u, b->n_msgs));
  if (n_workers <= 1) {
    c->strings.reserve(c->strings.size() + 64ull * b->n_msgs + 16ull * b->n_tensors);
    RangeOut r;
    compile_range(b, flags, c, 0, b->n_msgs, c->strings, field_bound, tensor_used, field_used, &r);
    if (r.fail) return set_err(ctx, TFS_ERR_INVALID_ARG, r.fail);
    c->bound = r.bound;
    c->has_input = r.has_input;
  } else {
    std::vector<RangeOut> outs(n_workers);
    auto first_of = [&](uint32_t k) { return static_cast<uint32_t>(static_cast<uint64_t>(b->n_msgs) * k / n_workers); };
    run_ranges(n_workers, [&](uint32_t k) {
      outs[k].strings.reserve(64ull * (first_of(k + 1) - first_of(k)));
      compile_range(b, flags, c, first_of(k), first_of(k + 1), outs[k].strings, field_bound, tensor_used, field_used, &outs[k]);
    });
    for (uint32_t k = 0; k < n_workers; ++k)  // the error of the earliest message, as the one-thread walk reports it
      if (outs[k].fail) return set_err(ctx, TFS_ERR_INVALID_ARG, outs[k].fail);
    // join the string blocks and move every name reference of a range by where its block landed
    std::vector<uint32_t> base(n_workers);
    uint64_t total = c->strings.size();
    for (uint32_t k = 0; k < n_workers; ++k) {
      base[k] = static_cast<uint32_t>(total);
      total += outs[k].strings.size();
      c->bound += outs[k].bound;
      c->has_input |= outs[k].has_input;
    }
    if (total > 0xFFFFFFF0ull) return set_err(ctx, TFS_ERR_TOO_LARGE, "names of the batch exceed 4 GiB");
    c->strings.resize(total);
    run_ranges(n_workers, [&](uint32_t k) {
      if (!outs[k].strings.empty()) std::memcpy(c->strings.data() + base[k], outs[k].strings.data(), outs[k].strings.size());
      const uint32_t d = base[k];
      for (uint32_t mi = first_of(k); mi < first_of(k + 1); ++mi) {
        EncMsg& o = c->msgs[mi];
        if (o.model_name.len) o.model_name.off += d;
        if (o.signature_name.len) o.signature_name.off += d;
        if (o.version_label.len) o.version_label.off += d;
        for (uint32_t ti = o.first_tensor; ti < o.first_tensor + o.n_tensors; ++ti)
          if (c->tensors[ti].name.len) c->tensors[ti].name.off += d;
      }
    });
  }
  // descriptors no message refers to: the block is reused, give them a defined (empty) record
  for (uint32_t ti = 0; ti < b->n_tensors; ++ti)
    if (!tensor_used[ti]) std::memset(&c->tensors[ti], 0, sizeof(EncTensor));
  for (uint32_t fi = 0; fi < b->n_fields; ++fi)
    if (!field_used[fi]) std::memset(&c->fields[fi], 0, sizeof(EncField));
  tr.mark("records");
  // tile tables: every varint / ragged field gets a run of tiles, every copy-class field a segment; positions are running
  // sums in field order.  Large batches: the sums of equal ranges of fields first (on several threads), then the fill.
  struct TileSums {
    uint64_t vt = 0, rt = 0, ro = 0, ct = 0;  // varint tiles, ragged tiles, ragged output tiles, copy tiles
    uint64_t nseg = 0, nvf = 0, nrf = 0;      // copy segments, varint fields, ragged fields
    bool any_bulk = false, any_small = false;
    int cast_kind = 0;
  };
  // fields [f0, f1) starting at the sums `at`; FILL = false only counts.  Returns the sums at f1.
  auto tile_range_pass = [&](uint32_t f0, uint32_t f1, TileSums at, bool fill) -> TileSums {
    for (uint32_t fi = f0; fi < f1; ++fi) {
      EncField& f = c->fields[fi];
      if (!field_used[fi] || f.n_elems == 0) {
        if (fill) {
          f.n_elems = field_used[fi] ? f.n_elems : 0;
          f.n_tiles = 0;
        }
        continue;
      }
      if (f.is_varint == FC_RAGGED) {
        const uint64_t nt = (f.n_elems + kRagTile - 1) / kRagTile, no = (field_bound[fi] + kRagOutTile - 1) / kRagOutTile;
        if (fill) {
          f.first_tile = static_cast<uint32_t>(at.rt);
          f.n_tiles = static_cast<uint32_t>(nt);
          c->rfield_first_tile[at.nrf] = static_cast<uint32_t>(at.rt);
          c->rfield_first_out[at.nrf] = static_cast<uint32_t>(at.ro);
          c->rfield_index[at.nrf] = fi;
        }
        ++at.nrf;
        at.rt += nt;
        at.ro += no;
      } else if (f.is_varint) {
        const uint64_t nt = (f.n_elems + kVarintTile - 1) / kVarintTile;
        if (fill) {
          f.first_tile = static_cast<uint32_t>(at.vt);
          f.n_tiles = static_cast<uint32_t>(nt);
          c->vfield_first_tile[at.nvf] = static_cast<uint32_t>(at.vt);
          c->vfield_index[at.nvf] = fi;
        }
        ++at.nvf;
        at.vt += nt;
      } else {
        const uint32_t osz = (f.mode == CP_BYTES || f.mode == CP_BOOL) ? 1u : (f.mode == CP_F32_F16 || f.mode == CP_F32_BF16) ? 2u : 4u;
        const uint64_t n = f.body_len / osz;
        if (fill) {
          CopySeg s;
          s.src = static_cast<const uint8_t*>(f.src);
          s.dst = nullptr;
          s.op = f.mode;
          s.param = f.op_param;
          s.n = n;
          s.first_tile = at.ct;
          f.first_tile = static_cast<uint32_t>(at.nseg);
          f.n_tiles = 0;
          c->segs[at.nseg] = s;
        }
        ++at.nseg;
        if (f.body_len <= kSmallSegBytes) {  // no tiles: k_copy_small takes it
          at.any_small = true;
          continue;
        }
        // +1: byte segments are tiled on the 16-byte grid of the destination address, which can add a tile
        at.ct += (f.body_len + kCopyTileBytes - 1) / kCopyTileBytes + 1;
        at.any_bulk |= f.mode == CP_BYTES && f.body_len >= 4096;
        if (f.mode >= CP_U8_F32 && n >= 1024)  // (the kernel re-checks size and source alignment per segment)
          at.cast_kind = std::max(at.cast_kind, (f.mode == CP_F32_F16 || f.mode == CP_F32_BF16) ? 2 : 1);
      }
    }
    return at;
  };
  const uint32_t tile_workers = b->n_fields >= 8 * kParallelCompileMsgs ? std::min<uint32_t>(n_workers, b->n_fields) : 1;
  auto field_first = [&](uint32_t k) { return static_cast<uint32_t>(static_cast<uint64_t>(b->n_fields) * k / tile_workers); };
  std::vector<TileSums> at(tile_workers + 1);
  if (tile_workers > 1) {
    std::vector<TileSums> part(tile_workers);
    run_ranges(tile_workers, [&](uint32_t k) { part[k] = tile_range_pass(field_first(k), field_first(k + 1), TileSums{}, false); });
    for (uint32_t k = 0; k < tile_workers; ++k) {
      TileSums& n = at[k + 1];
      n = at[k];
      n.vt += part[k].vt, n.rt += part[k].rt, n.ro += part[k].ro, n.ct += part[k].ct;
      n.nseg += part[k].nseg, n.nvf += part[k].nvf, n.nrf += part[k].nrf;
      n.any_bulk |= part[k].any_bulk;
      n.any_small |= part[k].any_small;
      n.cast_kind = std::max(n.cast_kind, part[k].cast_kind);
    }
  }
  // one thread: the tables are sized for the most there can be, the sums come out of the fill itself
  const TileSums& cap = at[tile_workers];
  c->segs.set_size(tile_workers > 1 ? cap.nseg : b->n_fields);
  c->vfield_first_tile.set_size((tile_workers > 1 ? cap.nvf : b->n_fields) + 1);
  c->vfield_index.set_size(tile_workers > 1 ? cap.nvf : b->n_fields);
  c->rfield_first_tile.set_size((tile_workers > 1 ? cap.nrf : b->n_fields) + 1);
  c->rfield_first_out.set_size((tile_workers > 1 ? cap.nrf : b->n_fields) + 1);
  c->rfield_index.set_size(tile_workers > 1 ? cap.nrf : b->n_fields);
  TileSums end;
  if (tile_workers > 1) {
    run_ranges(tile_workers, [&](uint32_t k) { tile_range_pass(field_first(k), field_first(k + 1), at[k], true); });
    end = cap;
  } else {
    end = tile_range_pass(0, b->n_fields, TileSums{}, true);
  }
  if (end.rt > 0x7FFFFFFFull || end.ro > 0x7FFFFFFFull) return set_err(ctx, TFS_ERR_INVALID_ARG, "batch too large (string_val tiles)");
  if (end.vt > 0x7FFFFFFFull) return set_err(ctx, TFS_ERR_INVALID_ARG, "batch too large (varint tiles)");
  c->segs.set_size(end.nseg);
  c->vfield_index.set_size(end.nvf);
  c->vfield_first_tile.set_size(end.nvf + 1);
  c->vfield_first_tile[end.nvf] = static_cast<uint32_t>(end.vt);
  c->rfield_index.set_size(end.nrf);
  c->rfield_first_tile.set_size(end.nrf + 1);
  c->rfield_first_out.set_size(end.nrf + 1);
  c->rfield_first_tile[end.nrf] = static_cast<uint32_t>(end.rt);
  c->rfield_first_out[end.nrf] = static_cast<uint32_t>(end.ro);
  c->n_copy_tiles = end.ct;
  c->any_bulk = end.any_bulk;
  c->any_small = end.any_small;
  c->cast_kind = end.cast_kind;
  // can the generic copy kernel be skipped?  Only the largest bulk body of a message is placed on a 16-byte
  // boundary by TFS_F_ALIGN_PAYLOAD, so every message may carry at most one bulk field, a plain byte copy
  // from a 16-byte aligned source, large enough for the TMA ring.
  // (TFS_F_ALIGN_PAYLOAD places tensor payloads; the lists of an Input are not aligned by it)
  tr.mark("tiles");
  c->all_tma = (flags & TFS_F_ALIGN_PAYLOAD) != 0 && !c->has_input;
  for (uint32_t mi = 0; mi < b->n_msgs && c->all_tma; ++mi) {
    uint32_t bulk = 0;
    const EncMsg& m = c->msgs[mi];
    for (uint32_t ti = m.first_tensor; ti < m.first_tensor + m.n_tensors; ++ti) {
      const EncTensor& t = c->tensors[ti];
      for (uint32_t fi = t.first_field; fi < t.first_field + t.n_fields; ++fi) {
        const EncField& f = c->fields[fi];
        if (f.is_varint || f.n_elems == 0 || f.body_len <= kSmallSegBytes) continue;  // (small bodies need no alignment)
        ++bulk;
        // (a widening / narrowing field that the staged cast kernel takes needs no alignment at all)
        const uint32_t in_sz = (f.mode == CP_F32_F16 || f.mode == CP_F32_BF16) ? 4u : (f.mode == CP_F16_F32 || f.mode == CP_BF16_F32) ? 2u : 1u;
        const uint32_t out_sz = (f.mode == CP_F32_F16 || f.mode == CP_F32_BF16) ? 2u : 4u;
        const bool cast_ok = f.mode >= CP_U8_F32 && f.body_len / out_sz >= 1024 && (reinterpret_cast<uintptr_t>(f.src) & (in_sz - 1u)) == 0;
        if (!cast_ok && (f.mode != CP_BYTES || f.body_len < 4096 || (reinterpret_cast<uintptr_t>(f.src) & 15u) != 0)) c->all_tma = false;
      }
    }
    if (bulk > 1) c->all_tma = false;
  }
  return TFS_OK;
}

struct Uploaded {
  EncTables T;
  uint64_t n_copy_tiles;
  bool has_input;
};

template <class V>
size_t place(size_t* cursor, const V& v) {
  size_t off = align_up(*cursor, 16);
  *cursor = off + v.size() * sizeof(typename V::value_type);
  return off;
}

tfs_status_t upload(tfs_ctx* ctx, const Compiled& c, Uploaded* u) {
  // device table block: the records (already in the pinned record block, msgs | tensors | fields) and the small tables
  const size_t o_msgs = 0, o_tensors = reinterpret_cast<const uint8_t*>(c.tensors) - reinterpret_cast<const uint8_t*>(c.msgs),
               o_fields = reinterpret_cast<const uint8_t*>(c.fields) - reinterpret_cast<const uint8_t*>(c.msgs);
  size_t cur = c.rec_bytes;
  const size_t o_filters = place(&cur, c.filters), o_strings = place(&cur, c.strings), o_vft = place(&cur, c.vfield_first_tile),
               o_vfi = place(&cur, c.vfield_index), o_segs = place(&cur, c.segs), o_rft = place(&cur, c.rfield_first_tile),
               o_rfo = place(&cur, c.rfield_first_out), o_rfi = place(&cur, c.rfield_index);
  const size_t total = align_up(cur, 16) + 16, small = total - c.rec_bytes;
  // Small tables go up in ONE copy.  Measured (tfs_encode_host, 1.2 GB of payload arriving in chunks on the copy stream):
  // one small host -> device copy overtakes the queued chunks, a second one behind it waits for all of them, and the
  // framing kernel with it - 47.8 instead of 29.3 ms per call.  Only a large record block (10^5 messages and more) is
  // uploaded from where it was compiled.
  const bool oneup = total <= (256u << 10);
  TFS_TRY(pinned_reserve(ctx, &ctx->h_up, &ctx->h_up_cap, oneup ? total : small));
  TFS_TRY(dev_reserve(ctx, &ctx->d_tables, total));
  uint8_t* h = static_cast<uint8_t*>(ctx->h_up) - (oneup ? 0 : c.rec_bytes);  // (offsets below are those of the device block)
  auto put = [&](size_t off, const void* p, size_t n) {
    if (n) std::memcpy(h + off, p, n);
  };
  put(o_filters, c.filters.data(), c.filters.size() * sizeof(StrRef));
  put(o_strings, c.strings.data(), c.strings.size());
  put(o_vft, c.vfield_first_tile.data(), c.vfield_first_tile.size() * 4);
  put(o_vfi, c.vfield_index.data(), c.vfield_index.size() * 4);
  put(o_segs, c.segs.data(), c.segs.size() * sizeof(CopySeg));
  put(o_rft, c.rfield_first_tile.data(), c.rfield_first_tile.size() * 4);
  put(o_rfo, c.rfield_first_out.data(), c.rfield_first_out.size() * 4);
  put(o_rfi, c.rfield_index.data(), c.rfield_index.size() * 4);
  if (oneup) {
    if (c.rec_bytes) std::memcpy(h, c.msgs, c.rec_bytes);
    TFS_CUDA(ctx, cudaMemcpyAsync(ctx->d_tables.p, h, total, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    TFS_CUDA(ctx, cudaMemcpyAsync(ctx->d_tables.p, c.msgs, c.rec_bytes, cudaMemcpyHostToDevice, ctx->stream));
    TFS_CUDA(ctx, cudaMemcpyAsync(static_cast<uint8_t*>(ctx->d_tables.p) + c.rec_bytes, ctx->h_up, small, cudaMemcpyHostToDevice, ctx->stream));
  }

  const uint32_t n_msgs = static_cast<uint32_t>(c.n_msgs), n_tensors = static_cast<uint32_t>(c.n_tensors);
  const uint32_t n_vtiles = c.vfield_first_tile.back();
  const uint32_t n_rtiles = c.rfield_first_tile.back(), n_routs = c.rfield_first_out.back();
  size_t w = 0;
  auto carve = [&](size_t bytes) {
    size_t off = align_up(w, 16);
    w = off + bytes;
    return off;
  };
  const size_t w_vt = carve(sizeof(VTile) * static_cast<size_t>(n_vtiles)), w_tf = carve(4ull * n_vtiles);
  const size_t w_tb = carve(4ull * n_vtiles), w_to = carve(8ull * n_vtiles), w_tl = carve(8ull * n_tensors),
               w_el = carve(8ull * n_tensors), w_eo = carve(8ull * n_tensors), w_fr = carve(8ull * n_tensors), w_fx = carve(4ull * n_tensors), w_sl = carve(8ull * n_msgs), w_ml = carve(8ull * n_msgs),
               w_mp = carve(8ull * n_msgs), w_mo = carve(8ull * n_msgs), w_res = carve(32), w_rb = carve(8ull * n_rtiles),
               w_ro = carve(8ull * n_rtiles);
  TFS_TRY(dev_reserve(ctx, &ctx->d_work, w + 16));
  uint8_t* d = static_cast<uint8_t*>(ctx->d_tables.p);
  uint8_t* k = static_cast<uint8_t*>(ctx->d_work.p);
  EncTables& T = u->T;
  T.msgs = reinterpret_cast<EncMsg*>(d + o_msgs);
  T.tensors = reinterpret_cast<EncTensor*>(d + o_tensors);
  T.fields = reinterpret_cast<EncField*>(d + o_fields);
  T.filters = reinterpret_cast<StrRef*>(d + o_filters);
  T.strings = d + o_strings;
  T.vfield_first_tile = reinterpret_cast<uint32_t*>(d + o_vft);
  T.vfield_index = reinterpret_cast<uint32_t*>(d + o_vfi);
  T.segs = reinterpret_cast<CopySeg*>(d + o_segs);
  T.vtiles = reinterpret_cast<VTile*>(k + w_vt);
  T.tile_field = reinterpret_cast<uint32_t*>(k + w_tf);
  T.tile_bytes = reinterpret_cast<uint32_t*>(k + w_tb);
  T.tile_off = reinterpret_cast<uint64_t*>(k + w_to);
  T.tensor_len = reinterpret_cast<uint64_t*>(k + w_tl);
  T.entry_len = reinterpret_cast<uint64_t*>(k + w_el);
  T.entry_off = reinterpret_cast<uint64_t*>(k + w_eo);
  T.feat_rec = reinterpret_cast<uint64_t*>(k + w_fr);
  T.feat_ex = reinterpret_cast<int32_t*>(k + w_fx);
  T.spec_len = reinterpret_cast<uint64_t*>(k + w_sl);
  T.msg_len = reinterpret_cast<uint64_t*>(k + w_ml);
  T.msg_pay_off = reinterpret_cast<uint64_t*>(k + w_mp);
  T.msg_off = reinterpret_cast<uint64_t*>(k + w_mo);
  T.result = reinterpret_cast<uint64_t*>(k + w_res);
  T.ticket = ctx->d_ctrl;  // k_frame_small's meeting counters
  T.rtile_bytes = reinterpret_cast<uint64_t*>(k + w_rb);
  T.rtile_off = reinterpret_cast<uint64_t*>(k + w_ro);
  T.rfield_first_tile = reinterpret_cast<uint32_t*>(d + o_rft);
  T.rfield_first_out = reinterpret_cast<uint32_t*>(d + o_rfo);
  T.rfield_index = reinterpret_cast<uint32_t*>(d + o_rfi);
  T.n_rfields = static_cast<uint32_t>(c.rfield_index.size());
  T.n_rtiles = n_rtiles;
  T.n_routs = n_routs;
  // result[0..1] are always written by the layout kernel; result[2] ("ragged offsets invalid") only accumulates
  if (n_rtiles) TFS_CUDA(ctx, cudaMemsetAsync(T.result, 0, 32, ctx->stream));
  T.n_msgs = n_msgs;
  T.n_tensors = n_tensors;
  T.n_fields = static_cast<uint32_t>(c.n_fields);
  T.n_filters = static_cast<uint32_t>(c.filters.size());
  T.n_vfields = static_cast<uint32_t>(c.vfield_index.size());
  T.n_vtiles = n_vtiles;
  T.n_segs = static_cast<uint32_t>(c.segs.size());
  u->n_copy_tiles = c.n_copy_tiles;
  u->has_input = c.has_input;
  return TFS_OK;
}

// size + layout (+ framing bytes when `dst` is given): everything that replaces prost's encoded_len / the
// non-payload part of encode_raw
tfs_status_t run_sizing(tfs_ctx* ctx, const Uploaded& u, uint32_t flags, uint32_t dst_skew, uint64_t dst_capacity, uint8_t* dst) {
  const EncTables& T = u.T;
  if (T.n_vtiles) {
    {
      k_tile_map<<<(T.n_vfields + 7) / 8, 256, 0, ctx->stream>>>(T);
      ctx->launches++;
      ProfScope prof(ctx, PK_VARINT_SIZE);
      k_varint_size<<<T.n_vtiles, kVarintThreads, 0, ctx->stream>>>(T);
    }
    ctx->launches++;
    const uint32_t warps_per_block = 256 / 32;
    k_field_scan<<<(T.n_vfields + warps_per_block - 1) / warps_per_block, 256, 0, ctx->stream>>>(T);
    ctx->launches++;
  }
  if (T.n_rtiles) {  // repeated bytes (string_val): record sizes per tile, then the record area length per field
    k_ragged_size<<<T.n_rtiles, kRagThreads, 0, ctx->stream>>>(T);
    ctx->launches++;
    k_ragged_scan<<<(T.n_rfields + 7) / 8, 256, 0, ctx->stream>>>(T);
    ctx->launches++;
  }
  if (u.has_input && T.n_tensors) {  // Input / ClassificationRequest: lengths of every feature, one thread each
    k_feature_layout<<<(T.n_tensors + 255) / 256, 256, 0, ctx->stream>>>(T);
    ctx->launches++;
  }
  if (T.n_msgs <= kSmallBatch) {
    {
      ProfScope prof(ctx, PK_FRAME);
      k_frame_small<<<(T.n_msgs + kFrameWarps - 1) / kFrameWarps, kFrameWarps * 32, 0, ctx->stream>>>(T, dst, flags, dst_skew, dst_capacity, dst != nullptr);
    }
    ctx->launches++;
  } else {
    k_msg_layout<<<(T.n_msgs + 127) / 128, 128, 0, ctx->stream>>>(T);
    ctx->launches++;
    k_msg_scan<<<1, 1024, 0, ctx->stream>>>(T, flags, dst_skew, dst_capacity);
    ctx->launches++;
    if (dst) {
      k_write_headers<<<(T.n_msgs + 127) / 128, 128, 0, ctx->stream>>>(T, dst, flags);
      ctx->launches++;
    }
  }
  if (u.has_input && T.n_tensors && dst) {  // the map entries themselves, at the places the per-message pass reserved
    k_feature_write<<<(T.n_tensors + 255) / 256, 256, 0, ctx->stream>>>(T, dst);
    ctx->launches++;
  }
  TFS_CUDA(ctx, cudaGetLastError());
  return TFS_OK;
}

// `behind_layout`: the copy runs on the copy stream behind ctx->enc_layout_ev (recorded after the framing kernels) and only
// that stream is waited for - the payload kernels queued after the event keep running (TFS_F_ASYNC)
tfs_status_t fetch_results(tfs_ctx* ctx, const EncTables& T, uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out,
                           bool* overflow, bool behind_layout = false) {
  // msg_len | msg_pay_off | msg_off | result were carved next to each other: one device -> host copy
  const size_t n = T.n_msgs;
  const uint8_t* base = reinterpret_cast<const uint8_t*>(T.msg_len);
  const size_t o_off = reinterpret_cast<const uint8_t*>(T.msg_off) - base, o_res = reinterpret_cast<const uint8_t*>(T.result) - base;
  const size_t bytes = o_res + 32;
  TFS_TRY(pinned_reserve(ctx, &ctx->h_down, &ctx->h_down_cap, bytes));
  uint8_t* h = static_cast<uint8_t*>(ctx->h_down);
  if (behind_layout) {
    TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->enc_layout_ev, 0));
    TFS_CUDA(ctx, cudaMemcpyAsync(h, base, bytes, cudaMemcpyDeviceToHost, ctx->copy_stream));
    TFS_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
  } else {
    TFS_CUDA(ctx, cudaMemcpyAsync(h, base, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  if (msg_off_out && n) std::memcpy(msg_off_out, h + o_off, 8 * n);
  if (msg_len_out && n) std::memcpy(msg_len_out, h, 8 * n);
  const uint64_t* res = reinterpret_cast<const uint64_t*>(h + o_res);
  if (total_out) *total_out = res[0];
  *overflow = res[1] != 0;
  if (T.n_rtiles && res[2] != 0) return set_err(ctx, TFS_ERR_INVALID_ARG, "string_val: the offsets in memory are not ascending from 0 to the stated total (aux)");
  return TFS_OK;
}

}  // namespace

// Encode a compiled batch whose field data pointers are device pointers.  When `sync_out` the results
// are fetched and the stream synchronised; otherwise the caller fetches later (host-staged path).
tfs_status_t encode_device(tfs_ctx* ctx, const tfs_batch_t* batch, void* dst_dev, uint64_t dst_capacity, uint32_t flags,
                           uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out, bool size_only) {
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  HostTrace tr(ctx, "encode");
  if (!ctx->enc_scratch) ctx->enc_scratch = std::make_shared<Compiled>();
  Compiled& c = *static_cast<Compiled*>(ctx->enc_scratch.get());
  TFS_TRY(compile_batch(ctx, batch, flags, &c, true));
  tr.mark("compile");
  Uploaded u;
  TFS_TRY(upload(ctx, c, &u));
  tr.mark("upload");
  const uint32_t dst_skew = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(dst_dev) & 15u);
  TFS_TRY(run_sizing(ctx, u, size_only ? 0u : flags, dst_skew, size_only ? ~0ull : dst_capacity,
                     size_only ? nullptr : static_cast<uint8_t*>(dst_dev)));
  const bool async = !size_only && (flags & TFS_F_ASYNC) != 0 && ctx->pipe.empty();
  if (async) {  // offsets, lengths and the overflow / validity words are final here
    if (!ctx->copy_stream) TFS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    if (!ctx->enc_layout_ev) TFS_CUDA(ctx, cudaEventCreateWithFlags(&ctx->enc_layout_ev, cudaEventDisableTiming));
    TFS_CUDA(ctx, cudaEventRecord(ctx->enc_layout_ev, ctx->stream));
  }
  if (!size_only) {
    const EncTables& T = u.T;
    uint8_t* dst = static_cast<uint8_t*>(dst_dev);
    if (T.n_vtiles) {
      {
        ProfScope prof(ctx, PK_VARINT_SCATTER);
        k_varint_scatter<<<T.n_vtiles, kVarintThreads, 0, ctx->stream>>>(T, dst);
      }
      ctx->launches++;
    }
    if (T.n_routs) {
      k_ragged_copy<<<T.n_routs, kRagThreads, 0, ctx->stream>>>(T, dst);
      ctx->launches++;
    }
    TFS_CUDA(ctx, cudaGetLastError());
    // gated on the overflow flag: when dst is too small segs[].dst were never filled in
    if (T.n_segs) {
      std::vector<HostSeg> hs;
      if (!ctx->pipe.empty()) {  // sources still arriving in chunks (tfs_encode_host): tell launch_copy where the segments are
        hs.reserve(c.segs.size());
        for (size_t i = 0; i < c.segs.size(); ++i) {
          const CopySeg& g = c.segs[i];
          const uint64_t next = i + 1 < c.segs.size() ? c.segs[i + 1].first_tile : c.n_copy_tiles;
          const uint32_t in = (g.op == CP_F32_F16 || g.op == CP_F32_BF16) ? 4u : (g.op == CP_F16_F32 || g.op == CP_BF16_F32) ? 2u : 1u;
          const uint32_t out = (g.op == CP_BYTES || g.op == CP_BOOL) ? 1u : (g.op == CP_F32_F16 || g.op == CP_F32_BF16) ? 2u : 4u;
          const uint64_t stb = static_cast<uint64_t>(kCopyTileBytes) / out * in;
          hs.push_back({reinterpret_cast<uintptr_t>(g.src), g.n * in, g.first_tile, next - g.first_tile, stb, stb});
        }
      }
      TFS_TRY(launch_copy(ctx, T.segs, T.n_segs, u.n_copy_tiles, T.result + 1, !c.all_tma, hs.empty() ? nullptr : &hs, c.any_bulk, c.cast_kind, c.any_small));
    }
  }
  bool overflow = false;
  tr.mark("launch");
  TFS_TRY(fetch_results(ctx, u.T, msg_off_out, msg_len_out, total_out, &overflow, async));
  tr.mark("fetch");
  if (overflow) return set_err(ctx, TFS_ERR_NOMEM, "dst_capacity is smaller than the encoded batch");
  return TFS_OK;
}

uint64_t encode_bound(const tfs_batch_t* batch, uint32_t flags) {
  tfs_ctx dummy;
  dummy.trace = std::getenv("TFS_TRACE") != nullptr;
  Compiled c;
  if (compile_batch(&dummy, batch, flags, &c) != TFS_OK) return 0;
  return c.bound + 64;
}

}  // namespace tfs

using namespace tfs;

extern "C" {

tfs_status_t tfs_encode_size(tfs_ctx_t* ctx, const tfs_batch_t* batch, uint64_t* msg_len_out) {
  if (!ctx || !batch || (batch->n_msgs && !msg_len_out)) return TFS_ERR_INVALID_ARG;
  uint64_t total = 0;
  return encode_device(ctx, batch, nullptr, 0, 0, nullptr, msg_len_out, &total, true);
}

uint64_t tfs_encode_bound(const tfs_batch_t* batch, uint32_t flags) { return batch ? encode_bound(batch, flags) : 0; }

tfs_status_t tfs_encode(tfs_ctx_t* ctx, const tfs_batch_t* batch, void* dst_dev, uint64_t dst_capacity, uint32_t flags,
                        uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out) {
  if (!ctx || !batch || (!dst_dev && dst_capacity)) return TFS_ERR_INVALID_ARG;
  return encode_device(ctx, batch, dst_dev, dst_capacity, flags, msg_off_out, msg_len_out, total_out, false);
}

}  // extern "C"
