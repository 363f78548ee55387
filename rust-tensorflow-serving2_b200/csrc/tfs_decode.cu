// tfs_decode.cu — batched TensorProto / PredictRequest / PredictResponse DECODE on sm_100a.
//
// Replaces, on the reference path (response of src/lib.rs:265, returned at src/lib.rs:266), what tonic's
// ProstDecoder does: prost `Message::decode` = merge loop + derived merge_field.
//   structure walk (decode_key / merge_field / hash_map::merge / skip_field)  -> K6d k_walk<COUNT>, k_walk<FILL>
//                                                                                 (one thread per message)
//   *::merge_repeated for varint types (decode_varint + push per element)      -> K4a k_varint_count (terminator ballots),
//                                                                                 K4b k_span_scan, K4c k_varint_gather
//   float/double::merge_repeated, bytes::merge                                 -> K5 k_copy (tfs_copy.cu)
// Element counts are known before any value is written (prost grows its Vecs by doubling instead).
// Integer / byte work bound by HBM bandwidth: no tensor cores.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "tfs_device.cuh"
#include "tfs_varint_dec.cuh"

struct tfs_decoded {
  uint64_t generation = 0;
  tfs_ctx* owner = nullptr;
  tfs::DecTables T{};
  uint64_t totals[24] = {0};
  // the result tables, in place in the pinned block they were copied into (owned by the handle, recycled on free)
  tfs::PinnedBlock down;
  const tfs_msg_out_t* msgs = nullptr;
  const tfs_tensor_out_t* tensors = nullptr;
  const tfs_span_t* filters = nullptr;
  const uint8_t* strings = nullptr;
  uint64_t n_msgs = 0, n_tensors = 0, n_filters = 0, n_strings = 0;
  uint32_t n_failed = 0;
  bool host_staged = false;  // wire was staged by tfs_decode_host
  uint64_t stage_generation = 0;
};

namespace tfs {

namespace {

enum { WT_VARINT = 0, WT_64 = 1, WT_LEN = 2, WT_SGROUP = 3, WT_EGROUP = 4, WT_32 = 5 };
// TOT_GENERIC counts the copy-class spans the TMA ring cannot take (not congruent mod 16, tiny, or widened)
enum Tot { TOT_TENSORS = 0, TOT_FILTERS, TOT_SPANS, TOT_STR, TOT_VTILES, TOT_CTILES, TOT_POOL, TOT_FAILED, TOT_GENERIC, TOT_OVERFLOW, TOT_RETRY, TOT_BULK, TOT_BAR0, TOT_BAR1, TOT_QT, TOT_QS, TOT_QF, TOT_QSTR, TOT_SMALL, TOT_GO, TOT_COUNT_ = 24 };  // TOT_BAR*: meeting counters of k_plan_warp; TOT_Q*: largest per-message counts seen (quota hints); TOT_SMALL: copy-class spans of at most kSmallSegBytes (no tiles, k_copy_small); TOT_GO: the plan kernel finished everything (tables final, pool large enough, copy segments bound) - a TMA ring queued behind it may run
constexpr size_t kTotalsBytes = sizeof(uint64_t) * TOT_COUNT_;

__device__ __forceinline__ uint32_t slot_of(uint32_t tag) {
  // 4..13 -> 0..9, 16 -> 10, 17 -> 11, 14 -> 12, 15 -> 13
  return tag <= 13 ? tag - 4u : tag >= 16 ? tag - 6u : tag - 2u;
}
// slots whose values are a TFS_T_RAGGED block: string_val, resource_handle_val, variant_val
__device__ __forceinline__ bool ragged_slot(uint32_t slot) { return slot == 4u || slot == 12u || slot == 13u; }
// message types the nested validator knows (what can sit inside a resource_handle_val / variant_val element)
enum NestedType : uint8_t { NT_VARIANT = 0, NT_TENSOR, NT_SHAPE, NT_DIM, NT_RESOURCE };

// Rust core::str::from_utf8 validity (no overlongs, no surrogates, <= U+10FFFF)
template <class Get>
__device__ bool utf8_ok(Get p, uint32_t n) {  // p(i) = i-th byte
  uint32_t i = 0;
  while (i < n) {
    const uint8_t b = p(i);
    if (b < 0x80) {
      i += 1;
    } else if (b >= 0xC2 && b <= 0xDF) {
      if (i + 1 >= n || (p(i + 1) & 0xC0) != 0x80) return false;
      i += 2;
    } else if (b >= 0xE0 && b <= 0xEF) {
      if (i + 2 >= n) return false;
      const uint8_t c1 = p(i + 1), c2 = p(i + 2);
      const uint8_t lo = b == 0xE0 ? 0xA0 : 0x80, hi = b == 0xED ? 0x9F : 0xBF;
      if (c1 < lo || c1 > hi || (c2 & 0xC0) != 0x80) return false;
      i += 3;
    } else if (b >= 0xF0 && b <= 0xF4) {
      if (i + 3 >= n) return false;
      const uint8_t c1 = p(i + 1), c2 = p(i + 2), c3 = p(i + 3);
      const uint8_t lo = b == 0xF0 ? 0x90 : 0x80, hi = b == 0xF4 ? 0x8F : 0xBF;
      if (c1 < lo || c1 > hi || (c2 & 0xC0) != 0x80 || (c3 & 0xC0) != 0x80) return false;
      i += 4;
    } else {
      return false;
    }
  }
  return true;
}

__device__ __forceinline__ uint32_t varint_tiles_of(const uint8_t* wire, uint64_t off, uint64_t len) {
  if (len == 0) return 0;
  const uint64_t a = reinterpret_cast<uintptr_t>(wire) + off;
  const uint64_t a0 = a & ~static_cast<uint64_t>(15);
  return static_cast<uint32_t>((a + len - a0 + kDecTileBytes - 1) / kDecTileBytes);
}

// =====================================================================================================
// K6d: structure walk, one thread per message.  FILL == false counts what the message needs
// (tensors, filters, spans, string bytes, tiles) and finds structural errors; FILL == true repeats
// the identical traversal and writes the records at the bases the scan of the counts produced.
// =====================================================================================================
template <bool FILL>
struct Walker {
  const DecTables& T;
  const uint8_t* w;   // wire base for the walk's own reads
  const uint8_t* wv;  // wire base the value kernels will read from (their tile grids follow ITS alignment)
  uint64_t beg, end, pos;
  uint32_t msg;
  int32_t err = 0;
  uint64_t err_pos = 0;
  bool soft_unsupported = false;
  bool verify = false;  // error path only: also look inside packed varint bodies (see walk_body)
  bool opaque = false;  // inside a resource_handle_val / variant_val element: validate only, record nothing
  // counters (COUNT) / cursors (FILL)
  uint32_t n_tensors = 0, n_filters = 0, n_spans = 0, n_str = 0, n_vt = 0, n_ct = 0, n_generic = 0, n_bulk = 0, n_small = 0;
  uint32_t seen_slots = 0;  // payload slots of the current tensor that already have a chunk
  // single-walk mode (k_plan_warp): the records go to per-message quota slots; a cursor that reaches its cap keeps counting
  // but writes to the slot's spare entry (tensors, spans) or nowhere (filters, string bytes), and `ovf` asks for a retry
  uint32_t cap_t = 0xFFFFFFFFu, cap_s = 0xFFFFFFFFu, cap_f = 0xFFFFFFFFu, cap_str = 0xFFFFFFFFu;
  bool ovf = false;
  // current tensor
  uint32_t cur = 0;    // its index
  uint32_t cur_w = 0;  // where its record is written (== cur unless over the quota)
  uint32_t cur_dims = 0;
  uint32_t content_span = 0xFFFFFFFFu;
  uint64_t tend = 0;  // end of the TensorProto being merged (the limit of its merge loop)

  // 16-byte window over the wire: header bytes are read through it, one aligned vector load per 16 bytes
  // instead of one dependent byte load per byte (the walk is a chain of dependent reads)
  uint4 win = make_uint4(0, 0, 0, 0);
  uint64_t win_base = ~0ull;

  __device__ Walker(const DecTables& t, uint32_t m) : T(t), w(t.walk_wire), wv(t.wire), msg(m) {}

  __device__ __forceinline__ uint8_t byte_at(uint64_t off) {
    const uint64_t a = reinterpret_cast<uintptr_t>(w) + off;
    const uint64_t base = a & ~static_cast<uint64_t>(15);
    if (base != win_base) {  // the aligned 16 bytes around a valid byte never leave its page
      win = *reinterpret_cast<const uint4*>(base);
      win_base = base;
    }
    const uint32_t k = static_cast<uint32_t>(a >> 2) & 3u;
    const uint32_t word = k == 0 ? win.x : k == 1 ? win.y : k == 2 ? win.z : win.w;
    return static_cast<uint8_t>(word >> ((static_cast<uint32_t>(a) & 3u) * 8u));
  }

  // the (up to) four bytes at `off`, read through the window without a branch per word; bytes that lie beyond the
  // window's 16 are junk
  __device__ __forceinline__ uint32_t bytes4_at(uint64_t off) {
    (void)byte_at(off);  // brings the window in
    const uint32_t a = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(w) + off);
    auto word_of = [&](uint32_t k) -> uint32_t {
      const uint32_t s01 = 0x3210u + (k & 1u) * 0x4444u;
      return __byte_perm(__byte_perm(win.x, win.y, s01), __byte_perm(win.z, win.w, s01), 0x3210u + (k >> 1) * 0x4444u);
    };
    const uint32_t k = (a >> 2) & 3u;
    return __funnelshift_r(word_of(k), word_of(min(k + 1u, 3u)), (a & 3u) * 8u);
  }
  __device__ __forceinline__ bool fail(int32_t code) {
    if (!err) {
      err = code;
      err_pos = pos - beg;
    }
    return false;
  }
  // prost decode_varint: <= 10 bytes, the 10th byte's high bits fall off, running out of buffer or a
  // 10th continuation byte is "invalid varint"
  __device__ bool varint(uint64_t* v) {
    uint64_t r = 0;
    const uint64_t rem = end - pos;
    const uint32_t n = rem < 10 ? static_cast<uint32_t>(rem) : 10u;
    for (uint32_t i = 0; i < n; ++i) {
      const uint8_t b = byte_at(pos++);
      r |= static_cast<uint64_t>(b & 0x7F) << (7u * i);
      if (b < 0x80) {
        *v = r;
        return true;
      }
    }
    return fail(TFS_ERR_INVALID_VARINT);
  }
  __device__ bool key(uint32_t* tag, uint32_t* wt) {
    uint64_t k;
    if (!varint(&k)) return false;
    if (k > 0xFFFFFFFFull) return fail(TFS_ERR_INVALID_KEY);
    const uint32_t x = static_cast<uint32_t>(k & 7u);
    if (x > 5u) return fail(TFS_ERR_WIRE_TYPE_VALUE);
    const uint32_t t = static_cast<uint32_t>(k) >> 3;
    if (t < 1u) return fail(TFS_ERR_INVALID_TAG);
    *tag = t;
    *wt = x;
    return true;
  }
  __device__ __forceinline__ bool check_wt(uint32_t expected, uint32_t actual) { return expected == actual ? true : fail(TFS_ERR_WIRE_TYPE); }
  // merge_loop prologue / epilogue
  __device__ bool loop_begin(uint64_t* region_end) {
    uint64_t len;
    if (!varint(&len)) return false;
    if (len > end - pos) return fail(TFS_ERR_BUFFER_UNDERFLOW);
    *region_end = pos + len;
    return true;
  }
  __device__ __forceinline__ bool loop_end(uint64_t region_end) { return pos == region_end ? true : fail(TFS_ERR_LENGTH_EXCEEDED); }

  __device__ bool skip_simple(uint32_t wt) {
    uint64_t len = 0;
    switch (wt) {
      case WT_VARINT: {
        uint64_t v;
        if (!varint(&v)) return false;
        break;
      }
      case WT_32: len = 4; break;
      case WT_64: len = 8; break;
      case WT_LEN:
        if (!varint(&len)) return false;
        break;
      default: return fail(TFS_ERR_END_GROUP);  // a stray end-group
    }
    if (len > end - pos) return fail(TFS_ERR_BUFFER_UNDERFLOW);
    pos += len;
    return true;
  }
  // prost skip_field with its recursion budget `ctx`; groups are unwound iteratively
  __device__ bool skip(uint32_t wt, uint32_t tag, uint32_t ctx) {
    if (ctx == 0) return fail(TFS_ERR_RECURSION);
    if (wt != WT_SGROUP) return skip_simple(wt);
    uint32_t stack[100];
    uint32_t depth = 1;
    stack[0] = tag;
    while (depth > 0) {
      uint32_t it, iw;
      if (!key(&it, &iw)) return false;
      if (iw == WT_EGROUP) {
        if (it != stack[depth - 1]) return fail(TFS_ERR_END_GROUP);
        --depth;
        continue;
      }
      if (ctx - depth == 0) return fail(TFS_ERR_RECURSION);
      if (iw == WT_SGROUP) {
        stack[depth++] = it;
        continue;
      }
      if (!skip_simple(iw)) return false;
    }
    return true;
  }

  // ---- outputs ----
  __device__ tfs_span_t add_string(uint64_t off, uint32_t len) {
    tfs_span_t sp;
    sp.off = n_str;
    sp.len = len;
    sp._pad = 0;
    if (FILL && static_cast<uint64_t>(n_str) + len > cap_str) ovf = true;
    if (FILL && static_cast<uint64_t>(n_str) + len <= cap_str) {
      uint8_t* dst = T.strings + n_str;
      for (uint32_t i = 0; i < len;) {
        const uint32_t in_window = 16u - (static_cast<uint32_t>(reinterpret_cast<uintptr_t>(w) + off + i) & 15u);
        const uint32_t n = min(min(4u, in_window), len - i);
        const uint32_t v = bytes4_at(off + i);
        dst[i] = static_cast<uint8_t>(v);
        if (n > 1) dst[i + 1] = static_cast<uint8_t>(v >> 8);
        if (n > 2) dst[i + 2] = static_cast<uint8_t>(v >> 16);
        if (n > 3) dst[i + 3] = static_cast<uint8_t>(v >> 24);
        i += n;
      }
    }
    n_str += len;
    return sp;
  }
  // string::merge: LEN wire type, length check, UTF-8 check
  __device__ bool string_field(uint32_t wt, tfs_span_t* out) {
    if (!check_wt(WT_LEN, wt)) return false;
    uint64_t len;
    if (!varint(&len)) return false;
    if (len > end - pos) return fail(TFS_ERR_BUFFER_UNDERFLOW);
    const uint64_t off = pos;
    pos += len;
    if (!utf8_ok([&](uint32_t i) { return byte_at(off + i); }, static_cast<uint32_t>(len))) return fail(TFS_ERR_UTF8);
    *out = add_string(off, static_cast<uint32_t>(len));
    return true;
  }

  __device__ void begin_tensor() {
    cur = n_tensors++;
    cur_w = cur;
    if (cur >= cap_t) {
      cur_w = cap_t;
      ovf = true;
    }
    cur_dims = 0;
    content_span = 0xFFFFFFFFu;
    seen_slots = 0;
    if (FILL) T.tensors[cur_w].msg_index = msg;
  }
  __device__ void add_span(uint32_t slot, uint32_t cls, uint64_t off, uint64_t len, uint64_t n_elems, uint32_t out_mult) {
    if ((verify || opaque) && cls == SC_VARINT && !err) {  // an over-long varint inside this body comes BEFORE whatever failed later
      uint32_t run = 0;
      for (uint64_t q = off; q < off + len; ++q) {
        if (byte_at(q) & 0x80) {
          if (++run == 10) {
            const uint64_t keep = pos;
            pos = q + 1;
            fail(TFS_ERR_INVALID_VARINT);
            pos = keep;
            break;
          }
        } else {
          run = 0;
        }
      }
    }
    if (opaque) {  // (nobody will look at this body again: it was checked just above, or needs no check)
      return;
    }
    const uint32_t vt = cls == SC_VARINT ? varint_tiles_of(wv, off, len) : 0u;
    // at most kSmallSegBytes of output: the span owns no copy tiles, k_copy_small takes it (one warp per span)
    const bool small = cls != SC_VARINT && len * out_mult <= kSmallSegBytes;
    if (small && len) n_small++;
    if (cls == SC_STR && len && !small) n_generic++;  // string bytes land wherever the earlier elements end: generic kernel
    if (cls != SC_VARINT && len >= 4096) n_bulk++;  // large enough for the TMA ring (if it turns out congruent mod 16)
    if (cls == SC_COPY && len && !small) {
      // Can the TMA ring take this span?  Its pool region starts on a 64-byte boundary, so the first chunk of
      // a slot is congruent mod 16 with its source iff the source address is 16-byte aligned.  A conservative
      // count (later chunks, tiny or widened spans count as "no") decides whether the generic kernel is launched.
      const bool first = !((seen_slots >> slot) & 1u);
      if (!first || len < 4096 || out_mult != 1 || ((reinterpret_cast<uintptr_t>(wv) + off) & 15u) != 0) n_generic++;
    }
    seen_slots |= 1u << slot;
    // +1: byte segments are tiled on the 16-byte grid of the destination address, which can add a tile
    const uint32_t ct = (cls != SC_VARINT && !small) ? static_cast<uint32_t>((len * out_mult + kCopyTileBytes - 1) / kCopyTileBytes) + 1u : 0u;
    if (FILL) {
      DecSpan s;
      s.wire_off = off;
      s.len = len;
      s.elem_base = 0;
      s.n_elems = n_elems;
      s.tensor = cur;
      s.slot = slot;
      s.msg = msg;
      s.cls = cls;
      s.first_tile = cls == SC_VARINT ? n_vt : n_ct;
      s.elem_tile0 = n_vt;
      uint32_t at = n_spans;
      if (at >= cap_s) {
        at = cap_s;
        ovf = true;
      }
      T.spans[at] = s;
      // both numberings stay monotone over the span table (a span owns no tiles of the other class)
      T.span_vtile0[at] = n_vt;
      T.span_ctile0[at] = n_ct;
    }
    n_spans++;
    n_vt += vt;
    n_ct += ct;
  }

  // {int32,int64,uint32,uint64,bool}::merge_repeated
  __device__ bool repeated_varint(uint32_t slot, uint32_t wt) {
    if (wt == WT_LEN) {
      uint64_t rend;
      if (!loop_begin(&rend)) return false;
      const uint64_t off = pos, len = rend - pos;
      if (len) {
        // does the last varint of the body end exactly at the limit?  prost reads on past it.
        if (byte_at(rend - 1) & 0x80) {
          // error path only: prost fails at the FIRST bad varint, so look for an over-long one (ten
          // continuation bytes in a row) earlier in the body before classifying the trailing one
          uint32_t run = 0;
          for (uint64_t q = off; q < rend; ++q) {
            if (byte_at(q) & 0x80) {
              if (++run == 10) {
                pos = q + 1;
                return fail(TFS_ERR_INVALID_VARINT);
              }
            } else {
              run = 0;
            }
          }
          uint32_t k = 0;  // continuation bytes at the end of the body
          while (k < 10 && k < len && (byte_at(rend - 1 - k) & 0x80)) ++k;
          pos = rend;
          if (k >= 10) return fail(TFS_ERR_INVALID_VARINT);
          for (uint32_t j = k; j < 10; ++j) {
            if (pos >= end) return fail(TFS_ERR_INVALID_VARINT);
            if (!(byte_at(pos++) & 0x80)) return fail(TFS_ERR_LENGTH_EXCEEDED);
          }
          return fail(TFS_ERR_INVALID_VARINT);
        }
        add_span(slot, SC_VARINT, off, len, 0, 1);
        if (opaque && err) return false;  // (the body was checked on the spot: nobody looks at it again)
      }
      pos = rend;
      return true;
    }
    if (!check_wt(WT_VARINT, wt)) return false;
    const uint64_t off = pos;
    uint64_t v;
    if (!varint(&v)) return false;
    add_span(slot, SC_VARINT, off, pos - off, 0, 1);
    return true;
  }
  // {float,double}::merge_repeated
  __device__ bool repeated_fixed(uint32_t slot, uint32_t wt, uint32_t width) {
    if (wt == WT_LEN) {
      uint64_t rend;
      if (!loop_begin(&rend)) return false;
      const uint64_t off = pos, len = rend - pos;
      const uint64_t rem = len % width;
      if (rem) {  // the last element straddles the limit, or the buffer ends inside it
        pos = rend - rem;
        if (end - pos >= width) {
          pos += width;
          return fail(TFS_ERR_LENGTH_EXCEEDED);
        }
        return fail(TFS_ERR_BUFFER_UNDERFLOW);
      }
      if (len) add_span(slot, SC_COPY, off, len, len / width, 1);
      pos = rend;
      return true;
    }
    if (!check_wt(width == 4 ? WT_32 : WT_64, wt)) return false;
    if (end - pos < width) return fail(TFS_ERR_BUFFER_UNDERFLOW);
    add_span(slot, SC_COPY, pos, width, 1, 1);
    pos += width;
    return true;
  }

  __device__ bool dim_field(uint32_t tag, uint32_t wt, uint32_t ctx, int64_t* size) {
    if (tag == 1) {
      if (!check_wt(WT_VARINT, wt)) return false;
      uint64_t v;
      if (!varint(&v)) return false;
      *size = static_cast<int64_t>(v);
      return true;
    }
    if (tag == 2) {
      tfs_span_t sp;
      const uint32_t before = n_str;
      if (!string_field(wt, &sp)) return false;
      n_str = before;  // Dim.name is not carried; a non-empty one makes the message UNSUPPORTED
      if (sp.len) soft_unsupported = true;
      return true;
    }
    return skip(wt, tag, ctx);
  }
  __device__ bool shape_field(uint32_t tag, uint32_t wt, uint32_t ctx) {
    if (tag == 2) {  // repeated Dim
      if (!check_wt(WT_LEN, wt)) return false;
      uint64_t rend;
      if (!loop_begin(&rend)) return false;
      int64_t size = 0;
      while (pos < rend) {
        uint32_t t2, w2;
        if (!key(&t2, &w2)) return false;
        if (!dim_field(t2, w2, ctx - 1, &size)) return false;
      }
      if (!loop_end(rend)) return false;
      if (cur_dims < TFS_MAX_DIMS) {
        if (FILL) T.tensors[cur_w].dims[cur_dims] = size;
      } else {
        soft_unsupported = true;
      }
      cur_dims++;
      if (FILL) T.tensors[cur_w].n_dims = cur_dims;
      return true;
    }
    if (tag == 3) {
      if (!check_wt(WT_VARINT, wt)) return false;
      uint64_t v;
      if (!varint(&v)) return false;
      if (FILL) T.tensors[cur_w].unknown_rank = v != 0;
      return true;
    }
    return skip(wt, tag, ctx);
  }
  // string::merge without keeping the string (inside an opaque element)
  __device__ bool string_opaque(uint32_t wt) {
    if (!check_wt(WT_LEN, wt)) return false;
    uint64_t len;
    if (!varint(&len)) return false;
    if (len > end - pos) return fail(TFS_ERR_BUFFER_UNDERFLOW);
    const uint64_t off = pos;
    pos += len;
    if (len > 0xFFFFFFFFull) return fail(TFS_ERR_UNSUPPORTED);
    if (!utf8_ok([&](uint32_t i) { return byte_at(off + i); }, static_cast<uint32_t>(len))) return fail(TFS_ERR_UTF8);
    return true;
  }
  __device__ bool bytes_opaque(uint32_t wt) {
    if (!check_wt(WT_LEN, wt)) return false;
    uint64_t len;
    if (!varint(&len)) return false;
    if (len > end - pos) return fail(TFS_ERR_BUFFER_UNDERFLOW);
    pos += len;
    return true;
  }
  __device__ bool varint_opaque(uint32_t wt) {
    if (!check_wt(WT_VARINT, wt)) return false;
    uint64_t v;
    return varint(&v);
  }
  // What prost's merge does to the body of a resource_handle_val / variant_val element, without keeping anything: the merge
  // loops of the nested messages (resource_handle.proto:13-30; tensor.proto:86-94 and, through `tensors`, tensor.proto:14-84,
  // tensor_shape.proto:13-46 again) unrolled into one loop over an explicit stack - a device thread has no stack to recurse
  // a hundred frames deep.  `budget` = the recursion budget the ROOT message's fields are merged with.
  __device__ bool validate_nested(uint8_t root, uint64_t root_end, uint32_t budget) {
    uint64_t ends[100];
    uint8_t types[100];
    uint32_t depth = 1, b = budget;  // b: the budget of the innermost open message's fields
    ends[0] = root_end;
    types[0] = root;
    while (depth) {
      const uint64_t e = ends[depth - 1];
      if (pos >= e) {  // merge_loop's end: the region must end exactly here
        if (pos != e) return fail(TFS_ERR_LENGTH_EXCEEDED);
        --depth;
        ++b;
        continue;
      }
      uint32_t tag, wt;
      if (!key(&tag, &wt)) return false;
      const uint8_t ty = types[depth - 1];
      uint8_t child = 0xFF;  // a message-typed field of this message?
      if (ty == NT_VARIANT && tag == 3) child = NT_TENSOR;
      else if (ty == NT_TENSOR && tag == 2) child = NT_SHAPE;
      else if (ty == NT_TENSOR && tag == 14) child = NT_RESOURCE;
      else if (ty == NT_TENSOR && tag == 15) child = NT_VARIANT;
      else if (ty == NT_SHAPE && tag == 2) child = NT_DIM;
      if (child != 0xFF) {  // message::merge: wire type, recursion limit, length; its fields get one less
        if (!check_wt(WT_LEN, wt)) return false;
        if (b == 0) return fail(TFS_ERR_RECURSION);
        uint64_t ce;
        if (!loop_begin(&ce)) return false;
        ends[depth] = ce;
        types[depth] = child;
        ++depth;
        --b;
        continue;
      }
      bool ok;
      switch (ty) {
        case NT_VARIANT: ok = tag == 1 ? string_opaque(wt) : tag == 2 ? bytes_opaque(wt) : skip(wt, tag, b); break;
        case NT_RESOURCE: ok = (tag == 1 || tag == 2 || tag == 3 || tag == 5) ? string_opaque(wt) : tag == 4 ? varint_opaque(wt) : skip(wt, tag, b); break;
        case NT_SHAPE: ok = tag == 3 ? varint_opaque(wt) : skip(wt, tag, b); break;
        case NT_DIM: ok = tag == 1 ? varint_opaque(wt) : tag == 2 ? string_opaque(wt) : skip(wt, tag, b); break;
        default:  // NT_TENSOR
          switch (tag) {
            case 1: case 3: ok = varint_opaque(wt); break;
            case 4: case 8: ok = bytes_opaque(wt); break;
            case 5: case 9: ok = repeated_fixed(0, wt, 4); break;
            case 6: case 12: ok = repeated_fixed(0, wt, 8); break;
            case 7: case 10: case 11: case 13: case 16: case 17: ok = repeated_varint(0, wt); break;
            default: ok = skip(wt, tag, b); break;
          }
      }
      if (!ok) return false;
    }
    return true;
  }

  // TensorProto::merge_field; `ctx` is the recursion budget its fields are merged with
  __device__ bool tensor_field(uint32_t tag, uint32_t wt, uint32_t ctx) {
    switch (tag) {
      case 1:
      case 3: {
        if (!check_wt(WT_VARINT, wt)) return false;
        uint64_t v;
        if (!varint(&v)) return false;
        if (FILL) {
          if (tag == 1) T.tensors[cur_w].dtype = static_cast<int32_t>(v);
          else T.tensors[cur_w].version_number = static_cast<int32_t>(v);
        }
        return true;
      }
      case 2: {
        if (!check_wt(WT_LEN, wt)) return false;
        uint64_t rend;
        if (!loop_begin(&rend)) return false;
        if (FILL) T.tensors[cur_w].has_shape = 1;
        while (pos < rend) {
          uint32_t t2, w2;
          if (!key(&t2, &w2)) return false;
          if (!shape_field(t2, w2, ctx - 1)) return false;
        }
        return loop_end(rend);
      }
      case 4: {  // bytes::merge: the last occurrence replaces earlier ones
        if (!check_wt(WT_LEN, wt)) return false;
        uint64_t len;
        if (!varint(&len)) return false;
        if (len > end - pos) return fail(TFS_ERR_BUFFER_UNDERFLOW);
        if (FILL && content_span != 0xFFFFFFFFu) {
          const uint32_t at = content_span < cap_s ? content_span : cap_s;
          T.spans[at].len = 0;
          T.spans[at].n_elems = 0;
        }
        content_span = n_spans;
        add_span(0, SC_COPY, pos, len, len, (T.flags & TFS_F_CONTENT_HALF_TO_F32) ? 2u : 1u);
        pos += len;
        return true;
      }
      case 5: case 9: return repeated_fixed(slot_of(tag), wt, 4);
      case 6: case 12: return repeated_fixed(slot_of(tag), wt, 8);
      case 7: case 10: case 11: case 13: case 16: case 17: return repeated_varint(slot_of(tag), wt);
      case 8: {  // repeated bytes: every occurrence appends one element (bytes::merge_repeated)
        if (!check_wt(WT_LEN, wt)) return false;
        uint64_t len;
        if (!varint(&len)) return false;
        if (len > end - pos) return fail(TFS_ERR_BUFFER_UNDERFLOW);
        add_span(slot_of(8), SC_STR, pos, len, 1, 1);
        pos += len;
        // A string tensor is a run of these.  While the next field is another string_val with a one-byte length
        // that stays inside the tensor, take it here: the same bookkeeping as above without a trip through the
        // key / wire-type / field dispatch per element.  Anything else (and every error) is left to the general path.
        // The loop keeps what it touches in registers: the walker itself lives in local memory (its methods are
        // real calls), and with one or two warps per batch every instruction of this chain is paid in full.
        {
          static_assert(sizeof(DecSpan) == 64 && offsetof(DecSpan, elem_base) == 16 && offsetof(DecSpan, tensor) == 32 &&
                        offsetof(DecSpan, first_tile) == 48, "string spans are written as four 16-byte stores");
          const uint64_t wa = reinterpret_cast<uintptr_t>(w), te = tend;
          uint64_t p = pos, lwb = win_base;
          uint4 lw = win;
          uint32_t ns = n_spans, nsm = n_small;
          const uint32_t nc = n_ct;  // (an element with a one-byte length is a small span: no copy tiles)
          const uint32_t nv = n_vt;
          const uint4 ids = make_uint4(cur, slot_of(8), msg, SC_STR);
          uint4* const spans = reinterpret_cast<uint4*>(T.spans);
          uint32_t* const sv = T.span_vtile0;
          uint32_t* const sc = T.span_ctile0;
          // word `k` of the window without a branch (the lanes of a warp sit at different offsets: a branch per
          // word would run all four arms one after the other)
          auto word_of = [&](uint32_t k) -> uint32_t {
            const uint32_t s01 = 0x3210u + (k & 1u) * 0x4444u;
            return __byte_perm(__byte_perm(lw.x, lw.y, s01), __byte_perm(lw.z, lw.w, s01), 0x3210u + (k >> 1) * 0x4444u);
          };
          // the (up to) four bytes at `off`, through the local copy of the window; bytes beyond the window are junk
          auto rd4 = [&](uint64_t off) -> uint32_t {
            const uint64_t a = wa + off, base = a & ~static_cast<uint64_t>(15);
            if (base != lwb) {
              lw = *reinterpret_cast<const uint4*>(base);
              lwb = base;
            }
            const uint32_t k = static_cast<uint32_t>(a >> 2) & 3u;
            return __funnelshift_r(word_of(k), word_of(min(k + 1u, 3u)), (static_cast<uint32_t>(a) & 3u) * 8u);
          };
          while (p + 2 <= te) {
            const uint32_t kl = rd4(p);  // key byte, length byte
            if ((kl & 0xFFu) != 0x42u) break;
            const uint32_t l = ((wa + p) & 15u) != 15u ? (kl >> 8) & 0xFFu : rd4(p + 1) & 0xFFu;
            if (l >= 0x80u || l > te - (p + 2)) break;
            if (FILL && ns >= cap_s) ovf = true;
            if (FILL && ns < cap_s) {
              const uint64_t off = p + 2;
              uint4* d = spans + 4ull * ns;
              d[0] = make_uint4(static_cast<uint32_t>(off), static_cast<uint32_t>(off >> 32), l, 0u);  // wire_off, len
              d[1] = make_uint4(0u, 0u, 1u, 0u);                                                     // elem_base, n_elems
              d[2] = ids;                                                                            // tensor, slot, msg, cls
              d[3] = make_uint4(nc, 0u, nv, 0u);                                                     // first_tile, elem_tile0
              sv[ns] = nv;
              sc[ns] = nc;
            }
            ns++;
            nsm += l != 0;  // as add_span counts a small span
            p += 2 + l;
          }
          pos = p;
          win = lw;
          win_base = lwb;
          n_spans = ns;
          n_small = nsm;
        }
        return true;
      }
      case 14: case 15: {
        // repeated ResourceHandleProto / VariantTensorDataProto (tensor.proto:72-76): message::merge_repeated.  Every element
        // is validated as prost's merge would walk it - the TensorProtos inside a variant, the variants inside those, ...
        // down to the recursion limit - and recorded like a string_val element: one span holding the element's body.
        if (!check_wt(WT_LEN, wt)) return false;
        if (ctx == 0) return fail(TFS_ERR_RECURSION);
        uint64_t rend;
        if (!loop_begin(&rend)) return false;
        const uint64_t off = pos, len = rend - pos;
        opaque = true;
        const bool ok = validate_nested(tag == 14 ? NT_RESOURCE : NT_VARIANT, rend, ctx - 1);
        opaque = false;
        if (!ok) return false;
        add_span(slot_of(tag), SC_STR, off, len, 1, 1);
        pos = rend;
        return true;
      }
      default: return skip(wt, tag, ctx);
    }
  }
  __device__ bool tensor_message(uint32_t wt, uint32_t ctx) {  // message::merge of a TensorProto
    if (!check_wt(WT_LEN, wt)) return false;
    uint64_t rend;
    if (!loop_begin(&rend)) return false;
    tend = rend;
    while (pos < rend) {
      uint32_t t2, w2;
      if (!key(&t2, &w2)) return false;
      if (!tensor_field(t2, w2, ctx - 1)) return false;
    }
    return loop_end(rend);
  }
  // hash_map::merge — the map field's own wire type is not looked at
  __device__ bool map_entry(uint32_t ctx) {
    uint64_t rend;
    if (!loop_begin(&rend)) return false;
    begin_tensor();
    while (pos < rend) {
      uint32_t t2, w2;
      if (!key(&t2, &w2)) return false;
      if (t2 == 1) {
        tfs_span_t sp;
        if (!string_field(w2, &sp)) return false;
        if (FILL) T.tensors[cur_w].key = sp;
      } else if (t2 == 2) {
        if (!tensor_message(w2, ctx - 1)) return false;
      } else if (!skip(w2, t2, ctx - 1)) {
        return false;
      }
    }
    return loop_end(rend);
  }
  // hash_map::merge of one `metadata` entry of a GetModelMetadataResponse (get_model_metadata.proto:29): key = 1 (string),
  // value = 2 (google.protobuf.Any { string type_url = 1; bytes value = 2 }).  The entry becomes a tensor record whose
  // string_val has two elements, [type_url, value]; repeated occurrences merge as prost merges them (the last one wins),
  // so the two spans are emitted when the entry ends.
  __device__ bool any_entry(uint32_t ctx) {
    uint64_t rend;
    if (!loop_begin(&rend)) return false;
    begin_tensor();
    uint64_t url_off = 0, url_len = 0, val_off = 0, val_len = 0;
    while (pos < rend) {
      uint32_t t2, w2;
      if (!key(&t2, &w2)) return false;
      if (t2 == 1) {
        tfs_span_t sp;
        if (!string_field(w2, &sp)) return false;
        if (FILL) T.tensors[cur_w].key = sp;
      } else if (t2 == 2) {  // message::merge of the Any
        if (!check_wt(WT_LEN, w2)) return false;
        if (ctx == 0) return fail(TFS_ERR_RECURSION);
        uint64_t aend;
        if (!loop_begin(&aend)) return false;
        while (pos < aend) {
          uint32_t t3, w3;
          if (!key(&t3, &w3)) return false;
          if (t3 == 1 || t3 == 2) {
            if (!check_wt(WT_LEN, w3)) return false;
            uint64_t len;
            if (!varint(&len)) return false;
            if (len > end - pos) return fail(TFS_ERR_BUFFER_UNDERFLOW);
            const uint64_t off = pos;
            pos += len;
            if (t3 == 1) {
              if (!utf8_ok([&](uint32_t i) { return byte_at(off + i); }, static_cast<uint32_t>(len))) return fail(TFS_ERR_UTF8);
              url_off = off, url_len = len;
            } else {
              val_off = off, val_len = len;
            }
          } else if (!skip(w3, t3, ctx - 2)) {  // (the Any's own fields are merged two levels below the map field)
            return false;
          }
        }
        if (!loop_end(aend)) return false;
      } else if (!skip(w2, t2, ctx - 1)) {
        return false;
      }
    }
    if (!loop_end(rend)) return false;
    add_span(slot_of(8), SC_STR, url_off ? url_off : beg, url_len, 1, 1);  // (an absent field is an empty element; any valid offset)
    add_span(slot_of(8), SC_STR, val_off ? val_off : beg, val_len, 1, 1);
    return true;
  }
  __device__ bool model_spec(uint32_t wt, uint32_t ctx, tfs_msg_out_t* mo) {
    if (!check_wt(WT_LEN, wt)) return false;
    uint64_t rend;
    if (!loop_begin(&rend)) return false;
    mo->has_model_spec = 1;
    while (pos < rend) {
      uint32_t t2, w2;
      if (!key(&t2, &w2)) return false;
      switch (t2) {
        case 1:
          if (!string_field(w2, &mo->model_name)) return false;
          break;
        case 3:
          if (!string_field(w2, &mo->signature_name)) return false;
          break;
        case 4:
          if (!string_field(w2, &mo->version_label)) return false;
          mo->version_choice = TFS_VERSION_LABEL;
          mo->version = 0;
          break;
        case 2: {  // Int64Value: oneof merge keeps the value when the variant is already `version`
          if (!check_wt(WT_LEN, w2)) return false;
          uint64_t r2;
          if (!loop_begin(&r2)) return false;
          if (mo->version_choice != TFS_VERSION_NUMBER) mo->version = 0;
          while (pos < r2) {
            uint32_t t3, w3;
            if (!key(&t3, &w3)) return false;
            if (t3 == 1) {
              if (!check_wt(WT_VARINT, w3)) return false;
              uint64_t v;
              if (!varint(&v)) return false;
              mo->version = static_cast<int64_t>(v);
            } else if (!skip(w3, t3, ctx - 2)) {
              return false;
            }
          }
          if (!loop_end(r2)) return false;
          mo->version_choice = TFS_VERSION_NUMBER;
          mo->version_label.off = 0;
          mo->version_label.len = 0;
          break;
        }
        default:
          if (!skip(w2, t2, ctx - 1)) return false;
      }
    }
    return loop_end(rend);
  }

  __device__ bool run(tfs_msg_out_t* mo) {
    if (T.flags & TFS_F_GRPC_PREFIX) {  // tonic Streaming: flag byte 0, u32 big-endian length
      const uint64_t n = end - beg;
      if (n < 5) return fail(TFS_ERR_BUFFER_UNDERFLOW);
      if (byte_at(beg) != 0) return fail(TFS_ERR_GRPC_FLAG);
      const uint64_t l = (static_cast<uint64_t>(byte_at(beg + 1)) << 24) | (static_cast<uint64_t>(byte_at(beg + 2)) << 16) |
                         (static_cast<uint64_t>(byte_at(beg + 3)) << 8) | byte_at(beg + 4);
      pos = beg + 1;
      if (l != n - 5) return fail(l > n - 5 ? TFS_ERR_BUFFER_UNDERFLOW : TFS_ERR_LENGTH_EXCEEDED);
      pos = beg + 5;
    }
    const uint64_t body = pos;
    if (T.msg_kind == TFS_MSG_TENSOR_PROTO) {
      begin_tensor();
      tend = end;
    }
    while (pos < end) {
      uint32_t tag, wt;
      if (!key(&tag, &wt)) return false;
      bool ok;
      if (T.msg_kind == TFS_MSG_TENSOR_PROTO) {
        ok = tensor_field(tag, wt, 100);
      } else if (T.msg_kind == TFS_MSG_PREDICT_REQUEST) {
        if (tag == 1) ok = model_spec(wt, 100, mo);
        else if (tag == 2) ok = map_entry(100);
        else if (tag == 3) {
          tfs_span_t sp;
          ok = string_field(wt, &sp);
          if (ok) {
            if (FILL && n_filters < cap_f) T.filters[n_filters] = sp;
            if (FILL && n_filters >= cap_f) ovf = true;
            n_filters++;
          }
        } else ok = skip(wt, tag, 100);
      } else if (T.msg_kind == TFS_MSG_METADATA_REQUEST || T.msg_kind == TFS_MSG_METADATA_RESPONSE) {
        if (tag == 1) ok = model_spec(wt, 100, mo);
        else if (tag == 2 && T.msg_kind == TFS_MSG_METADATA_RESPONSE) ok = any_entry(100);
        else if (tag == 2) {  // repeated string metadata_field
          tfs_span_t sp;
          ok = string_field(wt, &sp);
          if (ok) {
            if (FILL && n_filters < cap_f) T.filters[n_filters] = sp;
            if (FILL && n_filters >= cap_f) ovf = true;
            n_filters++;
          }
        } else ok = skip(wt, tag, 100);
      } else {
        if (tag == 1) ok = map_entry(100);
        else if (tag == 2) ok = model_spec(wt, 100, mo);
        else ok = skip(wt, tag, 100);
      }
      if (!ok) return false;
    }
    (void)body;
    if (soft_unsupported) {
      pos = beg;
      return fail(TFS_ERR_UNSUPPORTED);
    }
    return true;
  }
};

template <bool FILL>
__device__ void walk_body(const DecTables& T, uint32_t mi) {
  Walker<FILL> wk(T, mi);
  wk.beg = T.msg_off[mi];
  wk.end = wk.beg + T.msg_len[mi];
  wk.pos = wk.beg;
  tfs_msg_out_t mo;
  memset(&mo, 0, sizeof(mo));
  if (!FILL) {
    const bool in_bounds = wk.end <= T.wire_bytes && wk.end >= wk.beg;
    // the per-message counters (string bytes, spans, tiles) and string lengths are 32-bit: a message of 4 GiB or more
    // (more than a gRPC frame can carry) is left to prost instead of wrapping silently
    const bool too_long = T.msg_len[mi] > 0xFFFFFFF0ull;
    bool ok = !in_bounds ? wk.fail(TFS_ERR_BUFFER_UNDERFLOW) : too_long ? wk.fail(TFS_ERR_UNSUPPORTED) : wk.run(&mo);
    if (!ok && in_bounds && !too_long) {
      // prost stops at the FIRST error in stream order.  The walk does not look inside packed varint bodies (K4a does,
      // but only for messages that pass the walk), so a failed message is walked once more with those bodies checked.
      Walker<false> v(T, mi);
      v.beg = wk.beg;
      v.end = wk.end;
      v.pos = wk.beg;
      v.verify = true;
      tfs_msg_out_t scratch;
      memset(&scratch, 0, sizeof(scratch));
      v.run(&scratch);
      if (v.err) {
        wk.err = v.err;
        wk.err_pos = v.err_pos;
      }
    }
    T.msgs[mi].status = ok ? TFS_OK : wk.err;
    T.msgs[mi].err_offset = ok ? 0 : wk.err_pos;
    T.cnt_tensors[mi] = ok ? wk.n_tensors : 0u;
    T.cnt_filters[mi] = ok ? wk.n_filters : 0u;
    T.cnt_spans[mi] = ok ? wk.n_spans : 0u;
    T.cnt_strbytes[mi] = ok ? wk.n_str : 0u;
    T.cnt_vtiles[mi] = ok ? wk.n_vt : 0u;
    T.cnt_ctiles[mi] = ok ? wk.n_ct : 0u;
    if (ok && wk.n_generic) atomicAdd(reinterpret_cast<unsigned long long*>(&T.totals[TOT_GENERIC]), static_cast<unsigned long long>(wk.n_generic));
    if (ok && wk.n_bulk) atomicAdd(reinterpret_cast<unsigned long long*>(&T.totals[TOT_BULK]), static_cast<unsigned long long>(wk.n_bulk));
    if (ok && wk.n_small) atomicAdd(reinterpret_cast<unsigned long long*>(&T.totals[TOT_SMALL]), static_cast<unsigned long long>(wk.n_small));
  } else {
    const int32_t st = T.msgs[mi].status;
    const uint64_t eo = T.msgs[mi].err_offset;
    // cnt_* now hold the exclusive prefixes
    wk.n_tensors = T.cnt_tensors[mi];
    wk.n_filters = T.cnt_filters[mi];
    wk.n_spans = T.cnt_spans[mi];
    wk.n_str = T.cnt_strbytes[mi];
    wk.n_vt = T.cnt_vtiles[mi];
    wk.n_ct = T.cnt_ctiles[mi];
    mo.first_tensor = wk.n_tensors;
    mo.first_filter = wk.n_filters;
    const uint32_t t0 = wk.n_tensors, f0 = wk.n_filters;
    if (st == TFS_OK) {
      wk.run(&mo);
      mo.n_tensors = wk.n_tensors - t0;
      mo.n_filters = wk.n_filters - f0;
    } else {
      mo.status = st;
      mo.err_offset = eo;
    }
    T.msgs[mi] = mo;
  }
}

template <bool FILL>
// `lanes` (a power of two) threads of each warp take a message: the lanes of a warp wait for each other at every load of
// this serial walk, and a batch too small to fill the GPU loses nothing by spreading over more warps.
__global__ void __launch_bounds__(128) k_walk(DecTables T, uint32_t lanes) {
  const uint32_t lane = threadIdx.x & 31u, warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (lane >= lanes) return;
  const uint32_t mi = warp * lanes + lane;
  if (mi < T.n_msgs) walk_body<FILL>(T, mi);
}

// exclusive scan of `a[0..n)` in place by one CTA of NW warps, eight consecutive entries per thread and round (the arrays
// are 64-byte aligned); returns the total in every thread
template <int NW>
__device__ uint64_t cta_scan_u32(uint32_t* a, uint32_t n, uint64_t* smem) {
  uint64_t carry = 0;
  for (uint32_t base = 0; base < n; base += NW * 32 * 8) {
    const uint32_t i = base + threadIdx.x * 8u;
    uint32_t x[8];
    if (i + 8u <= n) {
      const uint4 lo = *reinterpret_cast<const uint4*>(a + i), hi = *reinterpret_cast<const uint4*>(a + i + 4);
      x[0] = lo.x, x[1] = lo.y, x[2] = lo.z, x[3] = lo.w, x[4] = hi.x, x[5] = hi.y, x[6] = hi.z, x[7] = hi.w;
    } else {
#pragma unroll
      for (uint32_t j = 0; j < 8; ++j) x[j] = i + j < n ? a[i + j] : 0u;
    }
    uint64_t v = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) v += x[j];
    uint64_t total;
    uint64_t run = carry + block_exclusive_scan<uint64_t, NW>(v, smem, &total);
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j) {
      if (i + j < n) a[i + j] = static_cast<uint32_t>(run);
      run += x[j];
    }
    carry += total;
    __syncthreads();
  }
  return carry;
}

// the six per-message counters -> exclusive offsets and totals; one CTA per counter
__global__ void __launch_bounds__(1024) k_scan_counts(DecTables T) {
  __shared__ uint64_t smem[33];
  uint32_t* const arr[6] = {T.cnt_tensors, T.cnt_filters, T.cnt_spans, T.cnt_strbytes, T.cnt_vtiles, T.cnt_ctiles};
  const int tot[6] = {TOT_TENSORS, TOT_FILTERS, TOT_SPANS, TOT_STR, TOT_VTILES, TOT_CTILES};
  const uint64_t t = cta_scan_u32<32>(arr[blockIdx.x], T.n_msgs, smem);
  if (threadIdx.x == 0) {
    T.totals[tot[blockIdx.x]] = t;
    if (blockIdx.x == 0) {
      T.totals[TOT_POOL] = 0;
      T.totals[TOT_FAILED] = 0;
      // TOT_GENERIC / TOT_BULK / TOT_SMALL were accumulated by the walk (the host zeroes the totals before pass 1)
    }
  }
}

constexpr uint32_t kSmallBatch = 256;

// K4b: per varint span, exclusive scan of its tile counts (one warp per span) -> tile_base, n_elems
__global__ void __launch_bounds__(256) k_span_scan(DecTables T, uint32_t n_spans, uint64_t n_vtiles) {
  const uint32_t si = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (si >= n_spans) return;
  if (T.spans[si].cls != SC_VARINT) return;
  const uint64_t t0 = T.span_vtile0[si];
  const uint64_t t1 = si + 1 < n_spans ? T.span_vtile0[si + 1] : n_vtiles;
  const uint32_t nt = static_cast<uint32_t>(t1 - t0);
  uint64_t carry = 0;
  for (uint32_t base = 0; base < nt; base += 128) {
    const uint32_t i0 = base + lane_id() * 4u;
    uint32_t x[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) x[j] = (i0 + j < nt) ? T.tile_count[t0 + i0 + j] : 0u;
    const uint32_t s = x[0] + x[1] + x[2] + x[3];
    const uint32_t incl = warp_inclusive_scan(s);
    uint64_t o = carry + (incl - s);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (i0 + j < nt) T.tile_base[t0 + i0 + j] = o;
      o += x[j];
    }
    carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
  }
  if (lane_id() == 0) T.spans[si].n_elems = carry;
}

// Element type a slot is decoded to, and its size
__device__ __forceinline__ int32_t slot_dst_type(uint32_t slot, int32_t dtype, uint32_t flags, uint64_t content_len) {
  const bool half16 = dtype == 19 || dtype == 14;
  switch (slot) {
    case 0: return (half16 && (flags & TFS_F_CONTENT_HALF_TO_F32) && content_len % 2 == 0) ? TFS_T_F32 : TFS_T_U8;
    case 1: case 5: return TFS_T_F32;
    case 2: case 8: return TFS_T_F64;
    case 3: return TFS_T_I32;
    case 4: case 12: case 13: return TFS_T_RAGGED;
    case 6: return TFS_T_I64;
    case 7: return TFS_T_BOOL;
    case 9: return (half16 && (flags & TFS_F_HALF_TO_F32)) ? TFS_T_F32 : TFS_T_I32;
    case 10: return TFS_T_U32;
    case 11: return TFS_T_U64;
    default: return TFS_T_U8;
  }
}
__device__ __forceinline__ uint32_t type_size_dev(int32_t t) {
  switch (t) {
    case TFS_T_U8: case TFS_T_I8: case TFS_T_BOOL: return 1;
    case TFS_T_U16: case TFS_T_I16: case TFS_T_F16: case TFS_T_BF16: return 2;
    case TFS_T_I32: case TFS_T_U32: case TFS_T_F32: return 4;
    default: return 8;
  }
}
// bytes a slot occupies in the pool
__device__ __forceinline__ uint64_t slot_pool_bytes(uint32_t slot, int32_t type, uint64_t n) {
  if (slot == 0) return type == TFS_T_F32 ? n * 2 : n;  // n = wire bytes of tensor_content
  return n * type_size_dev(type);
}

// K6e: per message, in wire order: element base of every span inside its (tensor, slot) vector, element
// totals per slot, final status (walk errors + varint errors found by K4a), pool bytes per tensor.
// `s0`, `s1`: the message's range in the span table.
__device__ __forceinline__ uint32_t span_end_of(const DecTables& T, uint32_t mi) {
  return mi + 1 < T.n_msgs ? T.cnt_spans[mi + 1] : static_cast<uint32_t>(T.totals[TOT_SPANS]);
}
__device__ void finalize_body(const DecTables& T, uint32_t mi, uint32_t s0, uint32_t s1) {
  tfs_msg_out_t& mo = T.msgs[mi];
  const unsigned long long e = T.err[mi];
  if (mo.status == TFS_OK && e != ~0ull) {
    mo.status = static_cast<int32_t>(e & 0xFF);
    mo.err_offset = e >> 8;
  }
  // (dst_off[] of a ragged slot doubles as the running byte count of its elements until the layout below assigns it;
  // the tensor records start out zeroed)
  // A run of spans of one (tensor, slot) - every element of a string_val, every chunk of a packed field - keeps its
  // running counts in registers: a round trip through global memory per span would make each iteration wait for the last.
  uint32_t run_tensor = ~0u, run_slot = 0;
  uint64_t run_elems = 0, run_str = 0;
  constexpr uint32_t kAhead = 4;  // spans whose fields are loaded before the first of them is used
  uint32_t nx_tensor[kAhead], nx_slot[kAhead];
  uint64_t nx_n[kAhead], nx_len[kAhead];
  for (uint32_t sb = s0; sb < s1; sb += kAhead) {
#pragma unroll
    for (uint32_t j = 0; j < kAhead; ++j) {
      const DecSpan& q = T.spans[min(sb + j, s1 - 1)];
      nx_tensor[j] = q.tensor;
      nx_slot[j] = q.slot;
      nx_n[j] = q.n_elems;
      nx_len[j] = q.len;
    }
#pragma unroll
    for (uint32_t j = 0; j < kAhead; ++j) {
      const uint32_t si = sb + j;
      if (si >= s1) break;
      DecSpan& s = T.spans[si];
      const uint32_t st = nx_tensor[j], ss = nx_slot[j];
      if (st != run_tensor || ss != run_slot) {
        if (run_tensor != ~0u) {
          T.tensors[run_tensor].n_elems[run_slot] = run_elems;
          if (ragged_slot(run_slot)) T.tensors[run_tensor].dst_off[run_slot] = run_str;
        }
        run_tensor = st;
        run_slot = ss;
        run_elems = T.tensors[st].n_elems[ss];
        run_str = ragged_slot(ss) ? T.tensors[st].dst_off[ss] : 0;
      }
      s.elem_base = run_elems;
      run_elems += nx_n[j];
      if (ragged_slot(ss)) {  // string_val / resource_handle_val / variant_val: the slots whose spans are SC_STR
        s.elem_tile0 = run_str;
        run_str += nx_len[j];
      }
    }
  }
  if (run_tensor != ~0u) {
    T.tensors[run_tensor].n_elems[run_slot] = run_elems;
    if (ragged_slot(run_slot)) T.tensors[run_tensor].dst_off[run_slot] = run_str;
  }
  for (uint32_t ti = mo.first_tensor; ti < mo.first_tensor + mo.n_tensors; ++ti) {
    tfs_tensor_out_t& t = T.tensors[ti];
    uint64_t bytes = 0;
    for (uint32_t sl = 0; sl < TFS_N_SLOTS; ++sl) {
      const int32_t ty = slot_dst_type(sl, t.dtype, T.flags, t.n_elems[sl]);
      t.dst_type[sl] = ty;
      const uint64_t str_bytes = t.dst_off[sl];  // ragged slots: the running byte count of the slot's elements was kept here
      t.dst_off[sl] = bytes;  // relative to the tensor's base for now
      // ragged slots: (n + 1) u64 offsets, then the bytes of all elements (TFS_T_RAGGED)
      const uint64_t need = ragged_slot(sl) ? (t.n_elems[sl] ? 8ull * (t.n_elems[sl] + 1) + str_bytes : 0ull) : slot_pool_bytes(sl, ty, t.n_elems[sl]);
      bytes += (need + 63u) & ~static_cast<uint64_t>(63);
    }
    T.tensor_pool[ti] = bytes;
  }
  if (mo.status != TFS_OK) {  // a failed message exposes nothing
    mo.n_tensors = 0;
    mo.n_filters = 0;
    mo.has_model_spec = 0;
    atomicAdd(reinterpret_cast<unsigned long long*>(&T.totals[TOT_FAILED]), 1ull);
  }
}

__global__ void __launch_bounds__(128) k_dec_finalize(DecTables T) {
  const uint32_t mi = blockIdx.x * blockDim.x + threadIdx.x;
  if (mi < T.n_msgs) finalize_body(T, mi, T.cnt_spans[mi], span_end_of(T, mi));
}

// exclusive scan of the per-tensor pool bytes, then absolute slot offsets. One CTA of NW warps.
template <int NW>
__device__ void pool_layout_body(const DecTables& T, uint32_t n_tensors, uint64_t pool_capacity, uint64_t* smem) {
  uint64_t carry = 0;
  for (uint32_t base = 0; base < n_tensors; base += NW * 32) {
    const uint32_t i = base + threadIdx.x;
    const uint64_t v = i < n_tensors ? T.tensor_pool[i] : 0ull;
    uint64_t total;
    const uint64_t ex = block_exclusive_scan<uint64_t, NW>(v, smem, &total);
    if (i < n_tensors) {
      const uint64_t b = carry + ex;
      for (uint32_t sl = 0; sl < TFS_N_SLOTS; ++sl) T.tensors[i].dst_off[sl] += b;
    }
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    T.totals[TOT_POOL] = carry;
    T.totals[TOT_OVERFLOW] = carry > pool_capacity ? 1ull : 0ull;  // gates the value kernels of a one-call decode
  }
}
// Large batches: the scan alone in one CTA (tensor_pool[] becomes each tensor's base), the 14 slot offsets of every tensor
// are then shifted by a grid of their own (one CTA doing both took 46 us for 2 000 tensors: 28 000 strided accesses).
__global__ void __launch_bounds__(1024) k_pool_layout(DecTables T, uint32_t n_tensors, uint64_t pool_capacity) {
  __shared__ uint64_t smem[33];
  uint64_t carry = 0;
  for (uint32_t base = 0; base < n_tensors; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint64_t v = i < n_tensors ? T.tensor_pool[i] : 0ull;
    uint64_t total;
    const uint64_t ex = block_exclusive_scan<uint64_t, 32>(v, smem, &total);
    if (i < n_tensors) T.tensor_pool[i] = carry + ex;
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    T.totals[TOT_POOL] = carry;
    T.totals[TOT_OVERFLOW] = carry > pool_capacity ? 1ull : 0ull;  // gates the value kernels of a one-call decode
  }
}
__global__ void __launch_bounds__(256) k_pool_apply(DecTables T, uint32_t n_tensors) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t i = g >> 4, sl = g & 15u;
  if (i < n_tensors && sl < TFS_N_SLOTS) T.tensors[i].dst_off[sl] += T.tensor_pool[i];
}

// Small batches: finalize and pool layout in ONE single-CTA launch
__global__ void __launch_bounds__(kSmallBatch) k_finalize_small(DecTables T, uint32_t n_tensors, uint64_t pool_capacity) {
  __shared__ uint64_t smem[kSmallBatch / 32 + 1];
  const uint32_t mi = (threadIdx.x & 31u) * (kSmallBatch / 32) + (threadIdx.x >> 5);  // dealt round the warps (k_plan_small)
  if (mi < T.n_msgs) finalize_body(T, mi, T.cnt_spans[mi], span_end_of(T, mi));  // reads what this message's walk wrote and the counted varint spans
  __syncthreads();
  pool_layout_body<kSmallBatch / 32>(T, n_tensors, pool_capacity, smem);
}

// at decode time: bind one copy-class span / string element to its place in the caller's pool.  In two steps, all the
// loads and then all the stores, so that a caller with several spans in hand can have their loads in flight together.
struct SpanBinding {
  CopySeg g;
  uint64_t* off_slot;  // string element: where its start offset goes (nullptr otherwise) ...
  uint64_t* end_slot;  // ... and, for the last element of a string_val, where the total goes
  uint64_t off_value, end_value;
};
__device__ __forceinline__ SpanBinding bind_prepare(const DecTables& T, uint32_t si, uint8_t* pool) {
  const DecSpan& s = T.spans[si];
  SpanBinding b;
  b.off_slot = nullptr;
  b.end_slot = nullptr;
  b.off_value = b.end_value = 0;
  CopySeg& g = b.g;
  g.first_tile = T.span_ctile0[si];
  g.src = T.wire + s.wire_off;
  g.dst = nullptr;
  g.n = 0;
  g.op = CP_BYTES;
  g.param = 0.f;
  if (s.cls == SC_COPY && s.len != 0 && T.msgs[s.msg].status == TFS_OK) {
    const tfs_tensor_out_t& t = T.tensors[s.tensor];
    const int32_t ty = t.dst_type[s.slot];
    if (s.slot == 0) {
      if (ty == TFS_T_F32) {
        g.op = t.dtype == 19 ? CP_F16_F32 : CP_BF16_F32;
        g.n = s.len / 2;
      } else {
        g.n = s.len;
      }
      g.dst = pool + t.dst_off[0];
    } else {
      g.n = s.len;  // fixed-width values are copied as bytes
      g.dst = pool + t.dst_off[s.slot] + s.elem_base * type_size_dev(ty);
    }
  } else if (s.cls == SC_STR && T.msgs[s.msg].status == TFS_OK) {
    // one string_val element: its start offset into the offsets array (the pool is 16-byte aligned and slots start
    // on 64-byte boundaries, so the offsets are aligned words), its bytes behind the offsets
    const tfs_tensor_out_t& t = T.tensors[s.tensor];
    const uint64_t n = t.n_elems[s.slot];
    uint8_t* block = pool + t.dst_off[s.slot];
    uint64_t* offs = reinterpret_cast<uint64_t*>(block);
    b.off_slot = offs + s.elem_base;
    b.off_value = s.elem_tile0;
    if (s.elem_base + 1 == n) {
      b.end_slot = offs + n;
      b.end_value = s.elem_tile0 + s.len;
    }
    g.n = s.len;
    g.dst = block + 8 * (n + 1) + s.elem_tile0;
  }
  return b;
}
__device__ __forceinline__ void bind_commit(const DecTables& T, uint32_t si, const SpanBinding& b) {
  if (b.off_slot) *b.off_slot = b.off_value;
  if (b.end_slot) *b.end_slot = b.end_value;
  T.segs[si] = b.g;
}
__device__ __forceinline__ void bind_span(const DecTables& T, uint32_t si, uint8_t* pool) { bind_commit(T, si, bind_prepare(T, si, pool)); }

// host and device agree on where the result blob (tensors | filters | strings) keeps its three arrays
__host__ __device__ inline void blob_layout(uint64_t n_tensors, uint64_t n_filters, uint64_t n_str, uint64_t* p_flt, uint64_t* p_str,
                                            uint64_t* bytes) {
  *p_flt = (sizeof(tfs_tensor_out_t) * n_tensors + 63u) & ~static_cast<uint64_t>(63);
  *p_str = (*p_flt + sizeof(tfs_span_t) * n_filters + 63u) & ~static_cast<uint64_t>(63);
  *bytes = *p_str + n_str;
}

// Batches of up to 1024 messages, planned optimistically in ONE launch: one WARP per message, one structure walk.
//   walk     lane 0 walks its message ONCE, writing tensor / span / filter records and string bytes into per-message QUOTA
//            slots (sized from what earlier batches on this context needed; a message that outgrows its slots keeps
//            counting, and the host retries with larger quotas).  The first version walked every message twice (count, then
//            fill at scanned offsets) with one thread per message in a single CTA: 72 us for 32 requests, 140 us for 256.
//   meet     the CTAs of this small grid (<= 32, all resident) meet at a counter; every warp then sums the counts of the
//            messages before its own (<= 256 of them) and the totals;
//   compact  the warp copies its records from the quota slots to their final, dense places, turning message-local indices
//            and offsets into batch-wide ones;
//   finish   a batch without varint bodies is finalized here too (element bases, pool layout after a second meeting, copy
//            segments bound to the pool): one host synchronisation per decode instead of two.
// totals[TOT_RETRY] = 1 asks the host for larger tables.
constexpr uint32_t kPlanWarps = 8;
struct PlanQuota {
  tfs_tensor_out_t* tensors;  // [n_msgs * (q_t + 1)]: the last entry of each message's run is the spare one
  DecSpan* spans;             // [n_msgs * (q_s + 1)]
  uint32_t* span_vtile0;      // [n_msgs * (q_s + 1)]
  uint32_t* span_ctile0;
  tfs_span_t* filters;        // [n_msgs * q_f]
  uint8_t* strings;           // [n_msgs * q_str]
  uint32_t* cnt[6];           // per-message COUNTS: tensors, filters, spans, string bytes, varint tiles, copy tiles
  uint32_t q_t, q_s, q_f, q_str;
};
static_assert(sizeof(tfs_tensor_out_t) % 8 == 0 && sizeof(DecSpan) % 8 == 0 && sizeof(tfs_span_t) % 8 == 0, "records are moved as 8-byte words");

__device__ __forceinline__ void grid_meet(uint64_t* counter) {  // all CTAs of a grid that is resident as a whole
  __threadfence();
  __syncthreads();
  if (gridDim.x > 1) {
    if (threadIdx.x == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(counter), 1ull);
      while (atomicAdd(reinterpret_cast<unsigned long long*>(counter), 0ull) < gridDim.x) __nanosleep(40);
      __threadfence();
    }
    __syncthreads();
  }
}
__device__ __forceinline__ unsigned long long warp_sum64(unsigned long long v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
  return v;
}
__device__ __forceinline__ void warp_move_words(void* dst, const void* src, uint64_t bytes) {
  unsigned long long* d = static_cast<unsigned long long*>(dst);
  const unsigned long long* s2 = static_cast<const unsigned long long*>(src);
  for (uint64_t i = lane_id(); i < bytes / 8u; i += 32u) d[i] = s2[i];
}

__global__ void __launch_bounds__(kPlanWarps * 32) k_plan_warp(DecTables T, PlanQuota Q, uint64_t blob_cap, uint32_t tensor_cap, uint32_t span_cap,
                                                                 uint64_t pool_capacity, uint8_t* pool) {
  const uint32_t mi = blockIdx.x * kPlanWarps + warp_id();
  const bool live = mi < T.n_msgs;
  const uint32_t lane = lane_id();
  // ---- walk (lane 0), into this message's quota slots
  tfs_msg_out_t mo;
  memset(&mo, 0, sizeof(mo));
  uint32_t c_t = 0, c_f = 0, c_s = 0, c_str = 0, c_vt = 0, c_ct = 0;
  const uint32_t bt = mi * (Q.q_t + 1), bs = mi * (Q.q_s + 1), bf = mi * Q.q_f, bstr = mi * Q.q_str;
  if (live) {
    // the walk only stores what differs from the default: clear the tensor records first
    unsigned long long* z = reinterpret_cast<unsigned long long*>(Q.tensors + bt);
    for (uint32_t i = lane; i < (Q.q_t + 1) * (sizeof(tfs_tensor_out_t) / 8); i += 32) z[i] = 0ull;
    __syncwarp();
    if (lane == 0) {
      DecTables W = T;
      W.tensors = Q.tensors;
      W.spans = Q.spans;
      W.span_vtile0 = Q.span_vtile0;
      W.span_ctile0 = Q.span_ctile0;
      W.filters = Q.filters;
      W.strings = Q.strings;
      Walker<true> wk(W, mi);
      wk.beg = T.msg_off[mi];
      wk.end = wk.beg + T.msg_len[mi];
      wk.pos = wk.beg;
      wk.n_tensors = bt;
      wk.cap_t = bt + Q.q_t;
      wk.n_spans = bs;
      wk.cap_s = bs + Q.q_s;
      wk.n_filters = bf;
      wk.cap_f = bf + Q.q_f;
      wk.n_str = bstr;
      wk.cap_str = bstr + Q.q_str;
      const bool in_bounds = wk.end <= T.wire_bytes && wk.end >= wk.beg;
      const bool too_long = T.msg_len[mi] > 0xFFFFFFF0ull;  // (see walk_body: 32-bit per-message counters)
      bool ok = !in_bounds ? wk.fail(TFS_ERR_BUFFER_UNDERFLOW) : too_long ? wk.fail(TFS_ERR_UNSUPPORTED) : wk.run(&mo);
      if (!ok && in_bounds && !too_long) {  // prost stops at the FIRST error in stream order: look inside the packed bodies too
        Walker<false> v(T, mi);
        v.beg = wk.beg;
        v.end = wk.end;
        v.pos = wk.beg;
        v.verify = true;
        tfs_msg_out_t scratch;
        memset(&scratch, 0, sizeof(scratch));
        v.run(&scratch);
        if (v.err) {
          wk.err = v.err;
          wk.err_pos = v.err_pos;
        }
      }
      if (ok) {
        c_t = wk.n_tensors - bt;
        c_f = wk.n_filters - bf;
        c_s = wk.n_spans - bs;
        c_str = wk.n_str - bstr;
        c_vt = wk.n_vt;
        c_ct = wk.n_ct;
        if (wk.n_generic) atomicAdd(reinterpret_cast<unsigned long long*>(&T.totals[TOT_GENERIC]), static_cast<unsigned long long>(wk.n_generic));
        if (wk.n_bulk) atomicAdd(reinterpret_cast<unsigned long long*>(&T.totals[TOT_BULK]), static_cast<unsigned long long>(wk.n_bulk));
        if (wk.n_small) atomicAdd(reinterpret_cast<unsigned long long*>(&T.totals[TOT_SMALL]), static_cast<unsigned long long>(wk.n_small));
        atomicMax(reinterpret_cast<unsigned long long*>(&T.totals[TOT_QT]), static_cast<unsigned long long>(c_t));
        atomicMax(reinterpret_cast<unsigned long long*>(&T.totals[TOT_QS]), static_cast<unsigned long long>(c_s));
        atomicMax(reinterpret_cast<unsigned long long*>(&T.totals[TOT_QF]), static_cast<unsigned long long>(c_f));
        atomicMax(reinterpret_cast<unsigned long long*>(&T.totals[TOT_QSTR]), static_cast<unsigned long long>(c_str));
        if (wk.ovf) T.totals[TOT_RETRY] = 1;  // a quota was too small: nothing of this launch is used
      } else {
        memset(&mo, 0, sizeof(mo));
        mo.status = wk.err;
        mo.err_offset = wk.err_pos;
      }
      Q.cnt[0][mi] = c_t;
      Q.cnt[1][mi] = c_f;
      Q.cnt[2][mi] = c_s;
      Q.cnt[3][mi] = c_str;
      Q.cnt[4][mi] = c_vt;
      Q.cnt[5][mi] = c_ct;
    }
  }
  grid_meet(&T.totals[TOT_BAR0]);
  // ---- this message's bases = counts of the messages before it; the totals (every warp, redundantly: n_msgs <= 256)
  unsigned long long pre[6] = {0, 0, 0, 0, 0, 0}, tot[6] = {0, 0, 0, 0, 0, 0};
  for (uint32_t base = 0; base < T.n_msgs; base += 32u) {
    const uint32_t i = base + lane;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const unsigned long long v = i < T.n_msgs ? __ldcg(&Q.cnt[k][i]) : 0u;
      tot[k] += v;
      if (i < mi) pre[k] += v;
    }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    pre[k] = warp_sum64(pre[k]);
    tot[k] = warp_sum64(tot[k]);
  }
  const uint64_t n_tensors = tot[0], n_filters = tot[1], n_spans = tot[2], n_str = tot[3], n_vt = tot[4];
  uint64_t p_flt, p_str, bytes;
  blob_layout(n_tensors, n_filters, n_str, &p_flt, &p_str, &bytes);
  const bool retry = __ldcg(&T.totals[TOT_RETRY]) != 0 || bytes > blob_cap || n_tensors > tensor_cap || n_spans > span_cap ||
                     n_str > 0xFFFFFFF0ull || tot[5] > 0xFFFFFFF0ull || n_vt > 0xFFFFFFF0ull;
  if (mi == 0 && lane == 0) {
    T.totals[TOT_TENSORS] = n_tensors;
    T.totals[TOT_FILTERS] = n_filters;
    T.totals[TOT_SPANS] = n_spans;
    T.totals[TOT_STR] = n_str;
    T.totals[TOT_VTILES] = n_vt;
    T.totals[TOT_CTILES] = tot[5];
    T.totals[TOT_POOL] = 0;
    if (retry) T.totals[TOT_RETRY] = 1;
  }
  if (retry) return;  // (uniform over the grid)
  DecTables L = T;
  uint8_t* blob = reinterpret_cast<uint8_t*>(T.tensors);
  L.filters = reinterpret_cast<tfs_span_t*>(blob + p_flt);
  L.strings = blob + p_str;
  const uint32_t t_base = static_cast<uint32_t>(pre[0]), f_base = static_cast<uint32_t>(pre[1]), s_base = static_cast<uint32_t>(pre[2]);
  const uint32_t str_base = static_cast<uint32_t>(pre[3]), vt_base = static_cast<uint32_t>(pre[4]), ct_base = static_cast<uint32_t>(pre[5]);
  if (live) {
    // exclusive prefixes, where the later kernels (and finalize below) expect them
    if (lane == 0) {
      T.cnt_tensors[mi] = t_base;
      T.cnt_filters[mi] = f_base;
      T.cnt_spans[mi] = s_base;
      T.cnt_strbytes[mi] = str_base;
      T.cnt_vtiles[mi] = vt_base;
      T.cnt_ctiles[mi] = ct_base;
    }
    // ---- compact: quota slots -> dense tables, message-local indices -> batch-wide ones
    c_t = __shfl_sync(0xFFFFFFFFu, c_t, 0);
    c_f = __shfl_sync(0xFFFFFFFFu, c_f, 0);
    c_s = __shfl_sync(0xFFFFFFFFu, c_s, 0);
    c_str = __shfl_sync(0xFFFFFFFFu, c_str, 0);
    warp_move_words(L.tensors + t_base, Q.tensors + bt, sizeof(tfs_tensor_out_t) * static_cast<uint64_t>(c_t));
    warp_move_words(L.spans + s_base, Q.spans + bs, sizeof(DecSpan) * static_cast<uint64_t>(c_s));
    warp_move_words(L.filters + f_base, Q.filters + bf, sizeof(tfs_span_t) * static_cast<uint64_t>(c_f));
    for (uint32_t i = lane; i < c_str; i += 32) L.strings[str_base + i] = Q.strings[bstr + i];
    for (uint32_t i = lane; i < c_s; i += 32) {
      L.span_vtile0[s_base + i] = Q.span_vtile0[bs + i] + vt_base;
      L.span_ctile0[s_base + i] = Q.span_ctile0[bs + i] + ct_base;
    }
    __syncwarp();
    for (uint32_t i = lane; i < c_t; i += 32) L.tensors[t_base + i].key.off = L.tensors[t_base + i].key.off - bstr + str_base;
    for (uint32_t i = lane; i < c_f; i += 32) L.filters[f_base + i].off = L.filters[f_base + i].off - bstr + str_base;
    for (uint32_t i = lane; i < c_s; i += 32) {
      DecSpan& sp = L.spans[s_base + i];
      sp.tensor = sp.tensor - bt + t_base;
      sp.first_tile += sp.cls == SC_VARINT ? vt_base : ct_base;
      if (sp.cls == SC_VARINT) sp.elem_tile0 += vt_base;  // (a string span keeps its byte offset there: set by finalize)
    }
    if (lane == 0) {
      if (mo.status == TFS_OK) {
        mo.model_name.off = mo.model_name.off - bstr + str_base;
        mo.signature_name.off = mo.signature_name.off - bstr + str_base;
        mo.version_label.off = mo.version_label.off - bstr + str_base;
        mo.first_tensor = t_base;
        mo.n_tensors = c_t;
        mo.first_filter = f_base;
        mo.n_filters = c_f;
      }
      T.msgs[mi] = mo;
    }
    __syncwarp();
  }
  if (n_vt != 0) return;  // with varint bodies the element totals are only known after K4a / K4b: k_finalize_small follows
  // ---- finish: element bases + pool bytes per tensor (own message), pool layout (needs every tensor's bytes), binding
  if (live && lane == 0) finalize_body(L, mi, s_base, s_base + c_s);
  grid_meet(&T.totals[TOT_BAR1]);
  unsigned long long before = 0, all = 0;
  for (uint64_t base = 0; base < n_tensors; base += 32u) {
    const uint64_t i = base + lane;
    const unsigned long long v = i < n_tensors ? __ldcg(&T.tensor_pool[i]) : 0ull;
    all += v;
    if (i < t_base) before += v;
  }
  before = warp_sum64(before);
  all = warp_sum64(all);
  if (mi == 0 && lane == 0) {
    T.totals[TOT_POOL] = all;
    T.totals[TOT_OVERFLOW] = all > pool_capacity ? 1ull : 0ull;  // gates the value kernels of a one-call decode
  }
  // the copy segments are bound below, by every warp: the go-ahead for a TMA ring queued behind this kernel is given by
  // the kernel boundary itself (TOT_GO only says that this point was reached with a pool that is large enough)
  if (mi == 0 && lane == 0) T.totals[TOT_GO] = (pool && all <= pool_capacity) ? 1ull : 0ull;
  if (!live) return;
  // (a failed message exposes nothing: finalize zeroed its counts in msgs[], but its tensor records still take part in the layout)
  unsigned long long run = before;
  for (uint32_t i = 0; i < c_t; ++i) {  // the tensors of one message are few: lanes split the slots
    const unsigned long long here = __ldcg(&T.tensor_pool[t_base + i]);
    if (lane < TFS_N_SLOTS) L.tensors[t_base + i].dst_off[lane] += run;
    run += here;
  }
  __syncwarp();
  if (pool && all <= pool_capacity)
    for (uint32_t i = lane; i < c_s; i += 32) bind_span(L, s_base + i, pool);
}

// at decode time: bind every copy-class span to its place in the caller's pool
__global__ void __launch_bounds__(256) k_dec_bind(DecTables T, uint32_t n_spans, uint8_t* pool, const uint64_t* __restrict__ gate) {
  const uint32_t si = blockIdx.x * blockDim.x + threadIdx.x;
  if (si >= n_spans || (gate && *gate != 0)) return;
  bind_span(T, si, pool);
}

// =====================================================================================================
// host orchestration
// =====================================================================================================
size_t carve(size_t* cur, size_t bytes, size_t align = 64) {
  size_t off = align_up(*cur, align);
  *cur = off + bytes;
  return off;
}

// value kernels: bind the copy spans to the pool, gather the varint spans, copy the bulk spans
tfs_status_t launch_values(tfs_ctx* ctx, tfs_decoded* dec, uint8_t* pool, const uint64_t* gate, const std::vector<HostSeg>* host_segs = nullptr,
                           bool spans_bound = false, bool tma_done = false) {
  const DecTables& T = dec->T;
  const uint64_t n_spans = dec->totals[TOT_SPANS], n_vt = dec->totals[TOT_VTILES], n_ct = dec->totals[TOT_CTILES];
  if (n_spans && !spans_bound) {
    k_dec_bind<<<static_cast<uint32_t>((n_spans + 255) / 256), 256, 0, ctx->stream>>>(T, static_cast<uint32_t>(n_spans), pool, gate);
    ctx->launches++;
  }
  if (n_vt) {
    // the varint kernels read whole bodies: when the wire is still arriving in chunks, wait for the last one
    if (!ctx->pipe.empty()) TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipe.back().ready, 0));
    k_dtile_bind<<<static_cast<uint32_t>((n_vt + 255) / 256), 256, 0, ctx->stream>>>(T, static_cast<uint32_t>(n_vt), pool, gate);
    ctx->launches++;
    {
      ProfScope prof(ctx, PK_VARINT_GATHER);
      k_varint_gather<<<static_cast<uint32_t>(n_vt), kGatherThreads, 0, ctx->stream>>>(T, pool, gate);
    }
    ctx->launches++;
  }
  TFS_CUDA(ctx, cudaGetLastError());
  // the generic copy kernel is only needed for spans the TMA ring cannot take (counted by the walk)
  const bool need_generic = dec->totals[TOT_GENERIC] != 0 || (reinterpret_cast<uintptr_t>(pool) & 15u) != 0;
  // tensor_content of a 16-bit float tensor widened to f32 (TFS_F_CONTENT_HALF_TO_F32): the staged cast kernel
  const int cast_kind = (T.flags & TFS_F_CONTENT_HALF_TO_F32) ? 1 : 0;
  const bool any_small = dec->totals[TOT_SMALL] != 0;
  if (n_ct || any_small)
    TFS_TRY(launch_copy(ctx, T.segs, static_cast<uint32_t>(n_spans), n_ct, gate, need_generic, host_segs, dec->totals[TOT_BULK] != 0 && !tma_done, cast_kind, any_small));
  return TFS_OK;
}

}  // namespace

// Scan (and, when `pool_dev` is given, decode the values too) with two host synchronisations: one for the table
// sizes, one for the results.  pool_dev == nullptr: scan only (the caller allocates the pool from the view).
tfs_status_t decode_scan_device(tfs_ctx* ctx, const void* wire_dev, uint64_t wire_bytes, const uint64_t* msg_off,
                                const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, void* pool_dev,
                                uint64_t pool_capacity, tfs_decoded** out, const void* walk_wire_dev) {
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  if ((msg_kind < TFS_MSG_TENSOR_PROTO || msg_kind > TFS_MSG_PREDICT_RESPONSE) && msg_kind != TFS_MSG_METADATA_REQUEST && msg_kind != TFS_MSG_METADATA_RESPONSE)
    return set_err(ctx, TFS_ERR_INVALID_ARG, "bad msg_kind");
  if (reinterpret_cast<uintptr_t>(pool_dev) & 15u) return set_err(ctx, TFS_ERR_INVALID_ARG, "the value pool must be 16-byte aligned (it holds typed arrays)");
  tfs_decoded* dec = new (std::nothrow) tfs_decoded();
  if (!dec) return TFS_ERR_NOMEM;
  dec->owner = ctx;
  dec->generation = ++ctx->dec_generation;
  HostTrace tr(ctx, "decode");
  DecTables& T = dec->T;
  std::memset(&T, 0, sizeof(T));
  auto bail = [&](tfs_status_t st) {
    if (dec->down.p) cudaStreamSynchronize(ctx->stream);  // copies into the block may still be queued
    down_release(dec->down);
    delete dec;
    return st;
  };
  // the handle's pinned result block: at least `bytes` (a block that has to grow loses its contents)
  auto need_down = [&](size_t bytes) -> tfs_status_t {
    if (dec->down.cap >= bytes) return TFS_OK;
    if (dec->down.p) {
      TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      down_release(dec->down);
      dec->down = PinnedBlock{};
    }
    return down_acquire(ctx, bytes, &dec->down);
  };
#define DEC_TRY(expr)                         \
  do {                                        \
    tfs_status_t st__ = (expr);               \
    if (st__ != TFS_OK) return bail(st__);    \
  } while (0)
#define DEC_CUDA(call)                                                  \
  do {                                                                  \
    cudaError_t e__ = (call);                                           \
    if (e__ != cudaSuccess) return bail(cuda_fail(ctx, e__, #call));    \
  } while (0)

  // ---- pass 1 tables: message ranges, counters, totals + per-message records (contiguous: one D2H), error words
  // (message ranges, error words = all ones, totals = zero) are adjacent and come up in ONE host -> device copy
  size_t cur = 0;
  const size_t o_off = carve(&cur, 8ull * n_msgs), o_len = carve(&cur, 8ull * n_msgs), o_err = carve(&cur, 8ull * n_msgs),
               o_tot = carve(&cur, kTotalsBytes);
  const size_t up_bytes = cur;
  const size_t o_msgs = carve(&cur, sizeof(tfs_msg_out_t) * static_cast<size_t>(n_msgs)), o_ct = carve(&cur, 4ull * n_msgs),
               o_cf = carve(&cur, 4ull * n_msgs), o_cs = carve(&cur, 4ull * n_msgs), o_cb = carve(&cur, 4ull * n_msgs),
               o_cv = carve(&cur, 4ull * n_msgs), o_cc = carve(&cur, 4ull * n_msgs);
  const size_t res1_bytes = o_msgs + sizeof(tfs_msg_out_t) * static_cast<size_t>(n_msgs) - o_tot;  // totals .. msgs: one D2H
  DEC_TRY(dev_reserve(ctx, &ctx->d_dec1, cur + 64));
  DEC_TRY(pinned_reserve(ctx, &ctx->h_up, &ctx->h_up_cap, up_bytes + 64));
  uint8_t* d1 = static_cast<uint8_t*>(ctx->d_dec1.p);
  {
    uint8_t* h = static_cast<uint8_t*>(ctx->h_up);
    if (n_msgs) {
      std::memcpy(h + o_off, msg_off, 8ull * n_msgs);
      std::memcpy(h + o_len, msg_len, 8ull * n_msgs);
      std::memset(h + o_err, 0xFF, 8ull * n_msgs);
    }
    std::memset(h + o_tot, 0, kTotalsBytes);
    DEC_CUDA(cudaMemcpyAsync(d1, h, up_bytes, cudaMemcpyHostToDevice, ctx->stream));
  }
  T.wire = static_cast<const uint8_t*>(wire_dev);
  T.walk_wire = walk_wire_dev ? static_cast<const uint8_t*>(walk_wire_dev) : T.wire;
  T.wire_bytes = wire_bytes;
  T.msg_off = reinterpret_cast<const uint64_t*>(d1 + o_off);
  T.msg_len = reinterpret_cast<const uint64_t*>(d1 + o_len);
  T.n_msgs = n_msgs;
  T.msg_kind = msg_kind;
  T.flags = flags;
  T.cnt_tensors = reinterpret_cast<uint32_t*>(d1 + o_ct);
  T.cnt_filters = reinterpret_cast<uint32_t*>(d1 + o_cf);
  T.cnt_spans = reinterpret_cast<uint32_t*>(d1 + o_cs);
  T.cnt_strbytes = reinterpret_cast<uint32_t*>(d1 + o_cb);
  T.cnt_vtiles = reinterpret_cast<uint32_t*>(d1 + o_cv);
  T.cnt_ctiles = reinterpret_cast<uint32_t*>(d1 + o_cc);
  T.totals = reinterpret_cast<uint64_t*>(d1 + o_tot);
  T.msgs = reinterpret_cast<tfs_msg_out_t*>(d1 + o_msgs);
  T.err = reinterpret_cast<unsigned long long*>(d1 + o_err);

  const uint32_t mblocks = (n_msgs + 127) / 128;
  // walk kernels: as few messages per warp as keeps the grid within what the GPU holds at once (3 CTAs of 4 warps per SM)
  uint32_t walk_lanes = 1;
  while (walk_lanes < 32 && (n_msgs + walk_lanes - 1) / walk_lanes > static_cast<uint64_t>(ctx->sm_count) * 12) walk_lanes *= 2;
  const uint32_t wblocks = static_cast<uint32_t>(((n_msgs + walk_lanes - 1) / walk_lanes + 3) / 4);
  const bool small = n_msgs <= kSmallBatch;
  const uint64_t pool_cap_arg = pool_dev ? pool_capacity : ~0ull;
  const size_t h1 = 0, h2 = align_up(res1_bytes, 64);
  uint64_t n_tensors = 0, n_filters = 0, n_spans = 0, n_str = 0, n_vt = 0;
  uint64_t p_flt = 0, p_str = 0, res2_bytes = 0;  // the result blob: tensors | filters | strings (one D2H)
  size_t blob_copied = 0;                           // bytes of the blob that already are in h_down
  bool tables_behind_event = false;                 // the result tables come back on the copy stream, behind ctx->dec_tables_ev
  bool tma_ahead = false;                           // the TMA ring was queued behind the plan kernel, before the counts were known
  bool planned = false;                             // small path, no varint bodies: the tables are final after sync #1
  uint8_t* d2 = nullptr;
  auto load_totals = [&](const uint8_t* hd) {
    std::memcpy(dec->totals, hd, kTotalsBytes);
    n_tensors = dec->totals[TOT_TENSORS];
    n_filters = dec->totals[TOT_FILTERS];
    n_spans = dec->totals[TOT_SPANS];
    n_str = dec->totals[TOT_STR];
    n_vt = dec->totals[TOT_VTILES];
  };

  // per-message quota slots of the single-walk plan kernel, from what earlier batches on this context needed per message
  auto quota_bytes = [&]() -> uint64_t {
    return static_cast<uint64_t>(n_msgs) * ((ctx->dec_q_t + 1ull) * sizeof(tfs_tensor_out_t) + (ctx->dec_q_s + 1ull) * (sizeof(DecSpan) + 8) +
                                            ctx->dec_q_f * sizeof(tfs_span_t) + ctx->dec_q_str + 64);
  };
  // one warp per message, the CTAs meet at a counter: the grid has to be resident as a whole (<= one CTA per SM), and every
  // warp sums the counts of all messages before its own (fine up to a thousand); a batch whose largest message has millions
  // of string elements would need huge slots: two walks instead
  bool plan_warp = n_msgs <= 1024 && (n_msgs + kPlanWarps - 1) / kPlanWarps <= static_cast<uint32_t>(ctx->sm_count) && quota_bytes() <= (256ull << 20);
  if (plan_warp) {
    // ---- optimistic plan: tables laid out from this context's hints, ONE walk per message, one launch
    for (int attempt = 0;; ++attempt) {
      const uint64_t t_cap = std::max<uint64_t>(ctx->dec_hint_tensors, 4ull * n_msgs + 16), f_cap = std::max<uint64_t>(ctx->dec_hint_filters, 4ull * n_msgs + 16),
                     s_cap = std::max<uint64_t>(ctx->dec_hint_spans, 8ull * n_msgs + 32), str_cap = std::max<uint64_t>(ctx->dec_hint_str, 256ull * n_msgs + 1024);
      uint64_t q_flt, q_str, blob_cap;
      blob_layout(t_cap, f_cap, str_cap, &q_flt, &q_str, &blob_cap);
      cur = 0;
      const size_t p_blob = carve(&cur, blob_cap), p_spn = carve(&cur, sizeof(DecSpan) * s_cap), p_sv = carve(&cur, 4 * (s_cap + 1)),
                   p_sc = carve(&cur, 4 * (s_cap + 1)), p_seg = carve(&cur, sizeof(CopySeg) * s_cap), p_tp = carve(&cur, 8 * t_cap);
      PlanQuota Q{};
      Q.q_t = ctx->dec_q_t;
      Q.q_s = ctx->dec_q_s;
      Q.q_f = ctx->dec_q_f;
      Q.q_str = ctx->dec_q_str;
      const size_t nm = n_msgs;
      const size_t p_qt = carve(&cur, nm * (Q.q_t + 1ull) * sizeof(tfs_tensor_out_t)), p_qs = carve(&cur, nm * (Q.q_s + 1ull) * sizeof(DecSpan)),
                   p_qv = carve(&cur, nm * (Q.q_s + 1ull) * 4), p_qc = carve(&cur, nm * (Q.q_s + 1ull) * 4),
                   p_qf = carve(&cur, nm * Q.q_f * sizeof(tfs_span_t) + 16), p_qstr = carve(&cur, nm * Q.q_str + 16), p_qn = carve(&cur, 6 * 4 * nm + 16);
      DEC_TRY(dev_reserve(ctx, &ctx->d_dec2, cur + 64));
      d2 = static_cast<uint8_t*>(ctx->d_dec2.p);
      Q.tensors = reinterpret_cast<tfs_tensor_out_t*>(d2 + p_qt);
      Q.spans = reinterpret_cast<DecSpan*>(d2 + p_qs);
      Q.span_vtile0 = reinterpret_cast<uint32_t*>(d2 + p_qv);
      Q.span_ctile0 = reinterpret_cast<uint32_t*>(d2 + p_qc);
      Q.filters = reinterpret_cast<tfs_span_t*>(d2 + p_qf);
      Q.strings = d2 + p_qstr;
      for (int k = 0; k < 6; ++k) Q.cnt[k] = reinterpret_cast<uint32_t*>(d2 + p_qn) + static_cast<size_t>(k) * nm;
      T.tensors = reinterpret_cast<tfs_tensor_out_t*>(d2 + p_blob);
      T.filters = nullptr;  // placed by blob_layout() once the counts are known (kernel and host agree)
      T.strings = nullptr;
      T.spans = reinterpret_cast<DecSpan*>(d2 + p_spn);
      T.span_vtile0 = reinterpret_cast<uint32_t*>(d2 + p_sv);
      T.span_ctile0 = reinterpret_cast<uint32_t*>(d2 + p_sc);
      T.segs = reinterpret_cast<CopySeg*>(d2 + p_seg);
      T.tensor_pool = reinterpret_cast<uint64_t*>(d2 + p_tp);
      {
        ProfScope prof(ctx, PK_WALK);
        k_plan_warp<<<(n_msgs + kPlanWarps - 1) / kPlanWarps, kPlanWarps * 32, 0, ctx->stream>>>(
            T, Q, blob_cap, static_cast<uint32_t>(std::min<uint64_t>(t_cap, 0xFFFFFFF0ull)), static_cast<uint32_t>(std::min<uint64_t>(s_cap, 0xFFFFFFF0ull)),
            pool_cap_arg, static_cast<uint8_t*>(pool_dev));
      }
      ctx->launches++;
      DEC_CUDA(cudaGetLastError());
      // The last plan on this context needed nothing but the TMA ring (bulk payloads, the C2 shape): queue it NOW, behind the
      // plan kernel, instead of after the host has seen the counts - it reads them from device memory and stays idle unless
      // the plan kernel says go.  Saves the GPU's idle time across the synchronisation (~20 us per decode).
      tma_ahead = false;
      if (pool_dev && ctx->pipe.empty() && ctx->dec_spec_tma) {
        // (the results then come back on the copy stream, behind an event recorded in front of the ring: the call must not
        // wait for the payload)
        if (!ctx->copy_stream) DEC_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        if (!ctx->dec_tables_ev) DEC_CUDA(cudaEventCreateWithFlags(&ctx->dec_tables_ev, cudaEventDisableTiming));
        DEC_CUDA(cudaEventRecord(ctx->dec_tables_ev, ctx->stream));
        DEC_TRY(launch_copy_tma_deferred(ctx, T.segs, T.totals + TOT_GO, T.totals + TOT_SPANS, T.totals + TOT_CTILES, ctx->dec_hint_spans > 64));
        tma_ahead = true;
      }
      const size_t guess = static_cast<size_t>(std::min<uint64_t>(blob_cap, std::max<uint64_t>(ctx->dec_hint_blob, 4096 + 512ull * n_msgs)));
      DEC_TRY(need_down(h2 + guess + 64));
      uint8_t* hd = static_cast<uint8_t*>(dec->down.p);
      cudaStream_t rs1 = ctx->stream;
      if (tma_ahead) {
        rs1 = ctx->copy_stream;
        DEC_CUDA(cudaStreamWaitEvent(rs1, ctx->dec_tables_ev, 0));
      }
      DEC_CUDA(cudaMemcpyAsync(hd + h1, d1 + o_tot, res1_bytes, cudaMemcpyDeviceToHost, rs1));
      DEC_CUDA(cudaMemcpyAsync(hd + h2, d2 + p_blob, guess, cudaMemcpyDeviceToHost, rs1));
      DEC_CUDA(cudaStreamSynchronize(rs1));  // sync #1 (the only one when the batch has no varint bodies)
      load_totals(hd + h1);
      // next time every message gets what the largest one needed now, and a quarter more
      auto grow = [](uint32_t* q, uint64_t seen) {
        const uint64_t want = seen + seen / 4 + 1;
        if (want > *q) *q = static_cast<uint32_t>(std::min<uint64_t>(want, 0x7FFFFFFFull));
      };
      grow(&ctx->dec_q_t, dec->totals[TOT_QT]);
      grow(&ctx->dec_q_s, dec->totals[TOT_QS]);
      grow(&ctx->dec_q_f, dec->totals[TOT_QF]);
      grow(&ctx->dec_q_str, dec->totals[TOT_QSTR]);
      blob_layout(n_tensors, n_filters, n_str, &p_flt, &p_str, &res2_bytes);
      ctx->dec_hint_tensors = std::max(ctx->dec_hint_tensors, n_tensors + n_tensors / 4);
      ctx->dec_hint_filters = std::max(ctx->dec_hint_filters, n_filters + n_filters / 4);
      ctx->dec_hint_spans = std::max(ctx->dec_hint_spans, n_spans + n_spans / 4);
      ctx->dec_hint_str = std::max(ctx->dec_hint_str, n_str + n_str / 4);
      ctx->dec_hint_blob = std::max<uint64_t>(ctx->dec_hint_blob, res2_bytes + res2_bytes / 4);
      if (dec->totals[TOT_RETRY]) {
        if (attempt >= 1) return bail(set_err(ctx, TFS_ERR_NOMEM, "decode plan did not fit its own tables"));
        DEC_CUDA(cudaMemsetAsync(d1 + o_tot, 0, kTotalsBytes, ctx->stream));
        continue;
      }
      blob_copied = guess;
      break;
    }
    if (n_vt > 0xFFFFFFF0ull || n_str > 0xFFFFFFF0ull || dec->totals[TOT_CTILES] > 0xFFFFFFF0ull)
      return bail(set_err(ctx, TFS_ERR_NOMEM, "batch too large (32-bit table offsets)"));
    T.filters = reinterpret_cast<tfs_span_t*>(reinterpret_cast<uint8_t*>(T.tensors) + p_flt);
    T.strings = reinterpret_cast<uint8_t*>(T.tensors) + p_str;
    planned = n_vt == 0;
    if (n_vt) {
      cur = 0;
      const size_t p_tc = carve(&cur, 4 * n_vt), p_ts = carve(&cur, sizeof(TileRef) * n_vt), p_td = carve(&cur, sizeof(TileDst) * n_vt),
                   p_tb = carve(&cur, 8 * n_vt);
      DEC_TRY(dev_reserve(ctx, &ctx->d_dec3, cur + 64));
      uint8_t* d3 = static_cast<uint8_t*>(ctx->d_dec3.p);
      T.tile_count = reinterpret_cast<uint32_t*>(d3 + p_tc);
      T.dtiles = reinterpret_cast<TileRef*>(d3 + p_ts);
      T.dtile_dst = reinterpret_cast<TileDst*>(d3 + p_td);
      T.tile_base = reinterpret_cast<uint64_t*>(d3 + p_tb);
    }
  } else {
    {
      ProfScope prof(ctx, PK_WALK);
      k_walk<false><<<wblocks, 128, 0, ctx->stream>>>(T, walk_lanes);
    }
    ctx->launches++;
    k_scan_counts<<<6, 1024, 0, ctx->stream>>>(T);
    ctx->launches++;
    DEC_CUDA(cudaGetLastError());
    DEC_TRY(need_down(kTotalsBytes));
    DEC_CUDA(cudaMemcpyAsync(dec->down.p, T.totals, kTotalsBytes, cudaMemcpyDeviceToHost, ctx->stream));
    tr.mark("setup");
    DEC_CUDA(cudaStreamSynchronize(ctx->stream));  // sync #1: the sizes of the pass-2 tables
    tr.mark("sync1");
    load_totals(static_cast<const uint8_t*>(dec->down.p));
    if (n_tensors > 0xFFFFFFF0ull || n_spans > 0xFFFFFFF0ull || n_vt > 0xFFFFFFF0ull || n_str > 0xFFFFFFF0ull || dec->totals[TOT_CTILES] > 0xFFFFFFF0ull)
      return bail(set_err(ctx, TFS_ERR_NOMEM, "batch too large (32-bit table offsets)"));
    // ---- pass 2 tables; tensors | filters | strings are contiguous: one D2H
    blob_layout(n_tensors, n_filters, n_str, &p_flt, &p_str, &res2_bytes);
    DEC_TRY(need_down(h2 + res2_bytes + 64));  // now, while the stream is idle: growing the block later would wait for the value kernels
    cur = 0;
    const size_t p_blob = carve(&cur, res2_bytes), p_spn = carve(&cur, sizeof(DecSpan) * n_spans), p_sv = carve(&cur, 4 * (n_spans + 1)),
                 p_sc = carve(&cur, 4 * (n_spans + 1)), p_tc = carve(&cur, 4 * n_vt), p_ts = carve(&cur, sizeof(TileRef) * n_vt),
                 p_td = carve(&cur, sizeof(TileDst) * n_vt), p_tb = carve(&cur, 8 * n_vt), p_seg = carve(&cur, sizeof(CopySeg) * n_spans),
                 p_tp = carve(&cur, 8 * n_tensors);
    DEC_TRY(dev_reserve(ctx, &ctx->d_dec2, cur + 64));
    d2 = static_cast<uint8_t*>(ctx->d_dec2.p);
    T.tensors = reinterpret_cast<tfs_tensor_out_t*>(d2 + p_blob);
    T.filters = reinterpret_cast<tfs_span_t*>(d2 + p_blob + p_flt);
    T.strings = d2 + p_blob + p_str;
    T.spans = reinterpret_cast<DecSpan*>(d2 + p_spn);
    T.span_vtile0 = reinterpret_cast<uint32_t*>(d2 + p_sv);
    T.span_ctile0 = reinterpret_cast<uint32_t*>(d2 + p_sc);
    T.tile_count = reinterpret_cast<uint32_t*>(d2 + p_tc);
    T.dtiles = reinterpret_cast<TileRef*>(d2 + p_ts);
    T.dtile_dst = reinterpret_cast<TileDst*>(d2 + p_td);
    T.tile_base = reinterpret_cast<uint64_t*>(d2 + p_tb);
    T.segs = reinterpret_cast<CopySeg*>(d2 + p_seg);
    T.tensor_pool = reinterpret_cast<uint64_t*>(d2 + p_tp);
    if (n_tensors) DEC_CUDA(cudaMemsetAsync(T.tensors, 0, sizeof(tfs_tensor_out_t) * n_tensors, ctx->stream));
    if (n_msgs) {
      ProfScope prof(ctx, PK_WALK);
      k_walk<true><<<wblocks, 128, 0, ctx->stream>>>(T, walk_lanes);
      ctx->launches++;
    }
  }

  // wire still arriving in chunks (tfs_decode_host_into): fetch the span table so that every chunk's copy launch can be
  // limited to the tiles that can end inside the chunk
  std::vector<HostSeg> hsegs;
  auto fetch_host_segs = [&]() -> tfs_status_t {
    if (ctx->pipe.empty() || !pool_dev || n_spans == 0 || n_spans > 65536) return TFS_OK;
    std::vector<DecSpan> sp(n_spans);
    std::vector<uint32_t> ct(n_spans);
    TFS_CUDA(ctx, cudaMemcpyAsync(sp.data(), T.spans, sizeof(DecSpan) * n_spans, cudaMemcpyDeviceToHost, ctx->stream));
    TFS_CUDA(ctx, cudaMemcpyAsync(ct.data(), T.span_ctile0, 4 * n_spans, cudaMemcpyDeviceToHost, ctx->stream));
    TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint64_t n_ct = dec->totals[TOT_CTILES];
    hsegs.reserve(n_spans);
    for (uint64_t i = 0; i < n_spans; ++i) {
      if (sp[i].cls == SC_VARINT) continue;
      const uint64_t next = i + 1 < n_spans ? ct[i + 1] : n_ct;
      // tensor_content may be widened 2 -> 4 bytes (half the source bytes per tile): bound both ways
      hsegs.push_back({reinterpret_cast<uintptr_t>(T.wire) + sp[i].wire_off, sp[i].len, ct[i], next - ct[i],
                       static_cast<uint64_t>(kCopyTileBytes) / 2, static_cast<uint64_t>(kCopyTileBytes)});
    }
    return TFS_OK;
  };
  if (!planned) {
    if (n_vt) {
      if (!ctx->pipe.empty()) DEC_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->pipe.back().ready, 0));  // whole bodies are read
      DEC_CUDA(cudaMemsetAsync(T.tile_count, 0, 4 * n_vt, ctx->stream));
      k_dtile_map<<<static_cast<uint32_t>((n_spans + 7) / 8), 256, 0, ctx->stream>>>(T, static_cast<uint32_t>(n_spans), static_cast<uint32_t>(n_vt));
      ctx->launches++;
      {
        ProfScope prof(ctx, PK_VARINT_COUNT);
        k_varint_count<<<static_cast<uint32_t>(n_vt), kCountThreads, 0, ctx->stream>>>(T);
      }
      ctx->launches++;
      k_span_scan<<<static_cast<uint32_t>((n_spans + 7) / 8), 256, 0, ctx->stream>>>(T, static_cast<uint32_t>(n_spans), n_vt);
      ctx->launches++;
    }
    if (small) {
      k_finalize_small<<<1, kSmallBatch, 0, ctx->stream>>>(T, static_cast<uint32_t>(n_tensors), pool_cap_arg);
      ctx->launches++;
    } else {
      k_dec_finalize<<<mblocks, 128, 0, ctx->stream>>>(T);
      ctx->launches++;
      k_pool_layout<<<1, 1024, 0, ctx->stream>>>(T, static_cast<uint32_t>(n_tensors), pool_cap_arg);
      ctx->launches++;
      if (n_tensors) {
        k_pool_apply<<<static_cast<uint32_t>((n_tensors * 16 + 255) / 256), 256, 0, ctx->stream>>>(T, static_cast<uint32_t>(n_tensors));
        ctx->launches++;
      }
    }
    DEC_CUDA(cudaGetLastError());
    // one-call decode: the value kernels follow at once, gated on the device by "the pool is large enough".  The result
    // tables are final HERE: they come back on the copy stream behind this event, so that the call returns while the value
    // kernels are still running (the device-pool forms are asynchronous; the host forms wait for the stream themselves)
    if (pool_dev && ctx->pipe.empty()) {
      if (!ctx->copy_stream) DEC_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
      if (!ctx->dec_tables_ev) DEC_CUDA(cudaEventCreateWithFlags(&ctx->dec_tables_ev, cudaEventDisableTiming));
      DEC_CUDA(cudaEventRecord(ctx->dec_tables_ev, ctx->stream));
      tables_behind_event = true;
    }
    if (pool_dev) {
      DEC_TRY(fetch_host_segs());
      DEC_TRY(launch_values(ctx, dec, static_cast<uint8_t*>(pool_dev), T.totals + TOT_OVERFLOW, hsegs.empty() ? nullptr : &hsegs));
    }
    blob_copied = 0;
  } else if (pool_dev && !dec->totals[TOT_OVERFLOW]) {
    // planned: the host already knows the layout, the value kernels need no gate
    DEC_TRY(fetch_host_segs());
    DEC_TRY(launch_values(ctx, dec, static_cast<uint8_t*>(pool_dev), nullptr, hsegs.empty() ? nullptr : &hsegs, true, tma_ahead));
  }
  ctx->dec_spec_tma = planned && pool_dev && dec->totals[TOT_BULK] != 0 && !dec->totals[TOT_OVERFLOW];  // (the hint for the next batch)

  // ---- results to the host: (totals | msgs) and the blob (tensors | filters | strings)
  tr.mark("launch");
  // (a block that had to grow starts empty: what the plan path already fetched is fetched again)
  const bool regrown = dec->down.cap < h2 + res2_bytes + 64;
  DEC_TRY(need_down(h2 + res2_bytes + 64));
  uint8_t* hd = static_cast<uint8_t*>(dec->down.p);
  if (!planned || regrown || blob_copied < res2_bytes) {
    cudaStream_t rs = ctx->stream;
    if (tables_behind_event && !regrown) {
      rs = ctx->copy_stream;
      DEC_CUDA(cudaStreamWaitEvent(rs, ctx->dec_tables_ev, 0));
    }
    if (!planned || regrown) DEC_CUDA(cudaMemcpyAsync(hd + h1, d1 + o_tot, res1_bytes, cudaMemcpyDeviceToHost, rs));
    if (res2_bytes) DEC_CUDA(cudaMemcpyAsync(hd + h2, T.tensors, res2_bytes, cudaMemcpyDeviceToHost, rs));
    DEC_CUDA(cudaStreamSynchronize(rs));  // sync #2: the caller needs the element counts
    std::memcpy(dec->totals, hd + h1, kTotalsBytes);
  }
  tr.mark("sync2");
  dec->msgs = reinterpret_cast<const tfs_msg_out_t*>(hd + h1 + (o_msgs - o_tot));
  dec->tensors = reinterpret_cast<const tfs_tensor_out_t*>(hd + h2);
  dec->filters = reinterpret_cast<const tfs_span_t*>(hd + h2 + p_flt);
  dec->strings = hd + h2 + p_str;
  dec->n_msgs = n_msgs;
  dec->n_tensors = n_tensors;
  dec->n_filters = n_filters;
  dec->n_strings = n_str;
  dec->n_failed = static_cast<uint32_t>(dec->totals[TOT_FAILED]);
  ctx->err_has_location = false;
  if (dec->n_failed) {  // the first failed message, for tfs_last_error_detail / tfs_last_error_location
    for (uint32_t i = 0; i < n_msgs; ++i) {
      if (dec->msgs[i].status == TFS_OK) continue;
      ctx->err_has_location = true;
      ctx->err_msg_index = i;
      ctx->err_byte_offset = dec->msgs[i].err_offset;
      ctx->err = std::string("message ") + std::to_string(i) + ": " + tfs_strerror(dec->msgs[i].status) + " near byte " +
                 std::to_string(dec->msgs[i].err_offset) + " (" + std::to_string(dec->n_failed) + " of " + std::to_string(n_msgs) + " messages failed)";
      break;
    }
  }
  tr.mark("copyout");
  *out = dec;
#undef DEC_TRY
#undef DEC_CUDA
  if (pool_dev && dec->totals[TOT_OVERFLOW]) return set_err(ctx, TFS_ERR_NOMEM, "value pool is smaller than tfs_decoded_view().pool_bytes; nothing was written");
  return dec->n_failed ? TFS_ERR_BATCH : TFS_OK;
}

tfs_status_t decode_device(tfs_ctx* ctx, tfs_decoded* dec, void* pool_dev, uint64_t pool_bytes) {
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  if (dec->owner != ctx || dec->generation != ctx->dec_generation)
    return set_err(ctx, TFS_ERR_INVALID_ARG, "stale decode handle: another tfs_decode_scan ran on this ctx since");
  if (reinterpret_cast<uintptr_t>(pool_dev) & 15u) return set_err(ctx, TFS_ERR_INVALID_ARG, "the value pool must be 16-byte aligned (it holds typed arrays)");
  if (pool_bytes < dec->totals[TOT_POOL] || (!pool_dev && dec->totals[TOT_POOL]))
    return set_err(ctx, TFS_ERR_NOMEM, "value pool is smaller than tfs_decoded_view().pool_bytes");
  return launch_values(ctx, dec, static_cast<uint8_t*>(pool_dev), nullptr);
}

uint64_t decoded_pool_bytes(const tfs_decoded* dec) { return dec->totals[TOT_POOL]; }
void decoded_mark_host_staged(tfs_decoded* dec, uint64_t stage_generation) {
  dec->host_staged = true;
  dec->stage_generation = stage_generation;
}
bool decoded_stage_is_current(const tfs_decoded* dec, uint64_t stage_generation) {
  return dec->host_staged && dec->stage_generation == stage_generation;
}

}  // namespace tfs

using namespace tfs;

extern "C" {

tfs_status_t tfs_decode_scan(tfs_ctx_t* ctx, const void* wire_dev, uint64_t wire_bytes, const uint64_t* msg_off,
                             const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, tfs_decoded_t** out) {
  if (!ctx || !out || (n_msgs && (!msg_off || !msg_len)) || (wire_bytes && !wire_dev)) return TFS_ERR_INVALID_ARG;
  *out = nullptr;
  return decode_scan_device(ctx, wire_dev, wire_bytes, msg_off, msg_len, n_msgs, msg_kind, flags, nullptr, 0, out, nullptr);
}

tfs_status_t tfs_decode_into(tfs_ctx_t* ctx, const void* wire_dev, uint64_t wire_bytes, const uint64_t* msg_off, const uint64_t* msg_len,
                             uint32_t n_msgs, int32_t msg_kind, uint32_t flags, void* pool_dev, uint64_t pool_capacity,
                             tfs_decoded_t** out) {
  if (!ctx || !out || !pool_dev || (n_msgs && (!msg_off || !msg_len)) || (wire_bytes && !wire_dev)) return TFS_ERR_INVALID_ARG;
  *out = nullptr;
  return decode_scan_device(ctx, wire_dev, wire_bytes, msg_off, msg_len, n_msgs, msg_kind, flags, pool_dev, pool_capacity, out, nullptr);
}

tfs_status_t tfs_decoded_view(const tfs_decoded_t* dec, tfs_decoded_view_t* view) {
  if (!dec || !view) return TFS_ERR_INVALID_ARG;
  view->msgs = dec->msgs;
  view->tensors = dec->tensors;
  view->filters = dec->filters;
  view->strings = dec->strings;
  view->strings_bytes = dec->n_strings;
  view->pool_bytes = dec->totals[TOT_POOL];
  view->n_msgs = static_cast<uint32_t>(dec->n_msgs);
  view->n_tensors = static_cast<uint32_t>(dec->n_tensors);
  view->n_filters = static_cast<uint32_t>(dec->n_filters);
  view->n_failed = dec->n_failed;
  return TFS_OK;
}

tfs_status_t tfs_decode(tfs_ctx_t* ctx, tfs_decoded_t* dec, void* pool_dev, uint64_t pool_bytes) {
  if (!ctx || !dec) return TFS_ERR_INVALID_ARG;
  TFS_TRY(decode_device(ctx, dec, pool_dev, pool_bytes));
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return TFS_OK;
}

tfs_status_t tfs_decoded_free(tfs_ctx_t* ctx, tfs_decoded_t* dec) {
  (void)ctx;
  if (dec) down_release(dec->down);  // (the copies into it were waited for before the handle was handed out)
  delete dec;
  return TFS_OK;
}

}  // extern "C"
