// tfs_varint_dec.cuh — K4: packed-varint DECODE (terminator count, gather).
//
// Replaces prost's {int32,int64,uint32,uint64,bool}::merge_repeated -> decode_varint + Vec::push per element
// (reached from the reference at src/lib.rs:265-266).  Integer work bound by HBM bandwidth.
//
// A packed body is cut into 4 KiB tiles on 16-byte boundaries of the absolute address; one CTA per tile, one
// aligned 16-byte load per thread.  A terminator is a byte with its MSB clear; a value belongs to the tile that
// holds its terminator.
//   K4a k_varint_count    terminators per tile (+ detection of 10 continuation bytes in a row = over-long varint).
//   K4c k_varint_gather   value index = span base + tile base + block-exclusive count of terminators.
//                         Fast path (no value of the tile longer than 4 bytes, i.e. < 2^28: token ids, half
//                         patterns, bools): each thread keeps its 16 bytes and the 16 bytes before them in
//                         registers; for each of its 16 byte positions the four bytes ending there are cut out with
//                         one funnel shift, the 7-bit groups are packed with two mask/shift steps and shifted by
//                         the value's length (from the continuation mask).  No byte loads, no data-dependent loop.
//                         General path: bytes in shared memory, walk back over at most 9 continuation bytes.
//                         Both paths write the tile's values to shared memory first; the CTA then stores them to
//                         the pool with fully coalesced accesses (a direct store per thread would scatter 8-byte
//                         writes over 32 sectors per instruction — that was the bottleneck of the first version).
#pragma once
#include "tfs_device.cuh"

namespace tfs {
namespace {

// one warp per span: a 32-byte record per varint tile, so that a CTA starts with one broadcast load instead of a
// search over the span table
__global__ void __launch_bounds__(256) k_dtile_map(DecTables T, uint32_t n_spans, uint32_t n_vtiles) {
  const uint32_t si = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (si >= n_spans) return;
  const uint32_t t0 = T.span_vtile0[si];
  const uint32_t t1 = si + 1 < n_spans ? T.span_vtile0[si + 1] : n_vtiles;
  if (t0 >= t1) return;
  const DecSpan& s = T.spans[si];
  TileRef r;
  r.span = si;
  r.msg = s.msg;
  r.span_abs0 = reinterpret_cast<uintptr_t>(T.wire) + s.wire_off;
  r.span_end = r.span_abs0 + s.len;
  for (uint32_t t = t0 + lane_id(); t < t1; t += 32) {
    r.tile_abs = (r.span_abs0 & ~static_cast<uint64_t>(15)) + static_cast<uint64_t>(t - t0) * kDecTileBytes;
    T.dtiles[t] = r;
  }
}

__device__ __forceinline__ TileRef tile_ref(const DecTables& T, uint32_t tile) { return T.dtiles[tile]; }

// at decode time, one thread per varint tile: where its values go and as what
__global__ void __launch_bounds__(256) k_dtile_bind(DecTables T, uint32_t n_vtiles, uint8_t* pool, const uint64_t* __restrict__ gate) {
  const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
  if (tile >= n_vtiles || (gate && *gate != 0)) return;
  const DecSpan& s = T.spans[T.dtiles[tile].span];
  const tfs_tensor_out_t& t = T.tensors[s.tensor];
  const int32_t ty = t.dst_type[s.slot];
  const uint32_t esz = ty == TFS_T_I64 || ty == TFS_T_U64 ? 8u : ty == TFS_T_BOOL ? 1u : 4u;
  TileDst d;
  d.dst = pool + t.dst_off[s.slot] + (s.elem_base + T.tile_base[tile]) * esz;
  d.type = T.msgs[s.msg].status == TFS_OK ? ty : TFS_T_NONE;  // NONE: the message failed, nothing is written
  d.dtype = t.dtype;
  T.dtile_dst[tile] = d;
}

// continuation and validity masks of 16 loaded bytes (bit j = byte j); a chunk wholly outside the span has none
__device__ __forceinline__ void masks_of(const uint4& d, uint64_t chunk_abs, uint64_t span_abs0, uint64_t span_end, uint32_t* cont,
                                         uint32_t* valid) {
  // MSB of each byte -> one bit per byte: the four isolated bits are moved to bits 24..27 by one multiply
  // (2^24, 2^17, 2^10, 2^3 put bytes 0..3 there; no two partial products meet, so there are no carries)
  auto msb4 = [](uint32_t x) { return ((((x >> 7) & 0x01010101u) * 0x01020408u) >> 24) & 0xFu; };
  const uint32_t m = msb4(d.x) | (msb4(d.y) << 4) | (msb4(d.z) << 8) | (msb4(d.w) << 12);
  uint32_t v = 0xFFFFu;
  if (chunk_abs + 16 <= span_abs0 || chunk_abs >= span_end) {
    v = 0;
  } else {
    if (chunk_abs < span_abs0) v &= 0xFFFFu << (span_abs0 - chunk_abs);
    if (chunk_abs + 16 > span_end) v &= 0xFFFFu >> (chunk_abs + 16 - span_end);
  }
  *valid = v & 0xFFFFu;
  *cont = m & *valid;
}

// 16 bytes of the tile owned by this thread: continuation mask and validity mask (bit j = byte j)
__device__ __forceinline__ void chunk_masks(uint64_t chunk_abs, uint64_t span_abs0, uint64_t span_end, uint4* data,
                                            uint32_t* cont, uint32_t* valid) {
  if (chunk_abs + 16 <= span_abs0 || chunk_abs >= span_end) {
    *data = make_uint4(0, 0, 0, 0);
    *cont = 0;
    *valid = 0;
    return;
  }
  const uint4 d = ld_stream16(reinterpret_cast<const void*>(chunk_abs));
  *data = d;
  masks_of(d, chunk_abs, span_abs0, span_end, cont, valid);
}

// continuation mask of the 16 bytes before this thread's chunk: the previous lane has it; lane 0 re-loads
__device__ __forceinline__ void prev_chunk(const TileRef& r, uint64_t chunk_abs, const uint4& own, uint32_t own_cont, uint4* pdata,
                                           uint32_t* pcont) {
  uint4 p;
  p.x = __shfl_up_sync(0xFFFFFFFFu, own.x, 1);
  p.y = __shfl_up_sync(0xFFFFFFFFu, own.y, 1);
  p.z = __shfl_up_sync(0xFFFFFFFFu, own.z, 1);
  p.w = __shfl_up_sync(0xFFFFFFFFu, own.w, 1);
  uint32_t pc = __shfl_up_sync(0xFFFFFFFFu, own_cont, 1);
  if (lane_id() == 0) {
    uint32_t pv;
    chunk_masks(chunk_abs - 16, r.span_abs0, r.span_end, &p, &pc, &pv);
  }
  *pdata = p;
  *pcont = pc;
}

// The kernel is bound by instruction issue, not by bandwidth: every thread takes TWO adjacent 16-byte chunks (the per-thread
// bookkeeping - tile record, warp reduction, atomics - is then paid once per 32 bytes), and the common case is kept short: a
// chunk wholly inside its span counts its terminators with one popcount of the MSBs, and the exact over-long check (ten
// continuation bytes in a row) runs only in warps where some aligned 4-byte word of a chunk, or of the chunk before it, is
// all continuation bytes - a run of ten covers such a word.
constexpr int kCountThreads = kDecTileBytes / 32;  // two chunks per thread
__global__ void __launch_bounds__(kCountThreads) k_varint_count(DecTables T) {
  __shared__ uint32_t warp_tail[kCountThreads / 32];  // does the last chunk of each warp hold an all-continuation word?
  const uint32_t tile = blockIdx.x;
  const TileRef r = tile_ref(T, tile);
  constexpr uint32_t kMsb = 0x80808080u;
  uint4 d[2];
  bool own_word[2];
  uint32_t cnt = 0;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const uint64_t chunk_abs = r.tile_abs + threadIdx.x * 32ull + 16ull * c;
    const bool inside = chunk_abs >= r.span_abs0 && chunk_abs + 16 <= r.span_end;
    const bool touches = chunk_abs + 16 > r.span_abs0 && chunk_abs < r.span_end;
    d[c] = make_uint4(0, 0, 0, 0);
    if (touches) d[c] = ld_stream16(reinterpret_cast<const void*>(chunk_abs));
    const uint32_t hx = d[c].x & kMsb, hy = d[c].y & kMsb, hz = d[c].z & kMsb, hw = d[c].w & kMsb;
    // (bytes outside the span may make this true for an edge chunk: that only sends the warp to the exact check)
    own_word[c] = hx == kMsb || hy == kMsb || hz == kMsb || hw == kMsb;
    if (inside) {
      cnt += 16u - __popc((hx >> 7) | (hy >> 6) | (hz >> 5) | (hw >> 4));
    } else if (touches) {  // first / last chunk of the span
      uint32_t cont, valid;
      masks_of(d[c], chunk_abs, r.span_abs0, r.span_end, &cont, &valid);
      cnt += __popc(~cont & valid);
    }
  }
  if (lane_id() == 31) warp_tail[threadIdx.x >> 5] = own_word[1];
  __syncthreads();
  bool prev_word = __shfl_up_sync(0xFFFFFFFFu, own_word[1], 1);  // of the chunk in front of this thread's first one
  const uint64_t first_abs = r.tile_abs + threadIdx.x * 32ull;
  if (lane_id() == 0) {
    if (threadIdx.x != 0) {
      prev_word = warp_tail[(threadIdx.x >> 5) - 1] != 0;
    } else {  // the chunk before the tile
      prev_word = false;
      if (first_abs > r.span_abs0) {
        const uint4 p = ld_stream16(reinterpret_cast<const void*>(first_abs - 16));
        prev_word = (p.x & kMsb) == kMsb || (p.y & kMsb) == kMsb || (p.z & kMsb) == kMsb || (p.w & kMsb) == kMsb;
      }
    }
  }
  if (__any_sync(0xFFFFFFFFu, own_word[0] || own_word[1] || prev_word)) {
    uint32_t cont0, cont1, valid;
    masks_of(d[0], first_abs, r.span_abs0, r.span_end, &cont0, &valid);
    masks_of(d[1], first_abs + 16, r.span_abs0, r.span_end, &cont1, &valid);
    uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, cont1, 1);
    if (lane_id() == 0) {
      uint4 pd;
      uint32_t pv;
      chunk_masks(first_abs - 16, r.span_abs0, r.span_end, &pd, &prev, &pv);
    }
    auto ten = [](uint32_t x) {  // bit i..i+9 all set <=> ten continuation bytes in a row
      x &= x >> 1;
      x &= x >> 2;
      x &= x >> 4;
      x &= x >> 2;
      return x;
    };
    const bool bad0 = ten((cont0 << 16) | prev) != 0, bad1 = ten((cont1 << 16) | cont0) != 0;
    if (bad0 || bad1) {
      const uint64_t here = bad0 ? first_abs : first_abs + 16;
      const uint64_t at = here > r.span_abs0 ? here : r.span_abs0;
      const uint64_t off = at - reinterpret_cast<uintptr_t>(T.wire) - T.msg_off[r.msg];
      atomicMin(&T.err[r.msg], (static_cast<unsigned long long>(off) << 8) | TFS_ERR_INVALID_VARINT);
    }
  }
  cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
  if (lane_id() == 0 && cnt) atomicAdd(&T.tile_count[tile], cnt);  // tile_count[] is zeroed by the host
}

// ---- gather -----------------------------------------------------------------------------------------------
enum OutKind : int { O_U32 = 0, O_U64, O_BOOL, O_F16_F32, O_BF16_F32 };

template <int OUT>
struct OutTraits {
  using type = uint32_t;
};
template <>
struct OutTraits<O_U64> {
  using type = uint64_t;
};
template <>
struct OutTraits<O_BOOL> {
  using type = uint8_t;
};

template <int OUT>
__device__ __forceinline__ typename OutTraits<OUT>::type convert_out(uint64_t v) {
  if constexpr (OUT == O_U32) return static_cast<uint32_t>(v);  // `as i32` / `as u32` truncates
  else if constexpr (OUT == O_U64) return v;
  else if constexpr (OUT == O_BOOL) return static_cast<uint8_t>(v != 0);
  else if constexpr (OUT == O_F16_F32) return __float_as_uint(f16_bits_to_f32(static_cast<uint32_t>(v) & 0xFFFFu));
  else return __float_as_uint(bf16_bits_to_f32(static_cast<uint32_t>(v) & 0xFFFFu));
}

// the 32-bit word a value of at most 4 wire bytes is staged as (widened / narrowed to the output type by the flush)
template <int OUT>
__device__ __forceinline__ uint32_t convert_word(uint32_t v) {
  if constexpr (OUT == O_U32 || OUT == O_U64) return v;  // non-negative and < 2^28: the upper half of an i64 is zero
  else if constexpr (OUT == O_BOOL) return v != 0 ? 1u : 0u;
  else if constexpr (OUT == O_F16_F32) return __float_as_uint(f16_bits_to_f32(v & 0xFFFFu));
  else return __float_as_uint(bf16_bits_to_f32(v & 0xFFFFu));
}

// Output staging: 2048 values of up to 8 bytes.  A tile holds at most 4096 values (one per wire byte), so a dense
// tile is written out in two rounds; keeping the buffer at 16 KiB lets 8 CTAs share an SM.
constexpr uint32_t kRoundValues = 2048;
// Value i sits at slot i + i / 16: a thread's values are consecutive, so lanes write at strides that are often
// multiples of 16 slots (16 one-byte varints per 16-byte chunk is the dense case); the pad spreads them over
// the banks, and the flush still reads (almost) consecutive slots.
__device__ __forceinline__ uint32_t pad_slot(uint32_t i) { return i + (i >> 4); }
// The fast paths stage 32-bit words, 16 bytes of pad per 16 values: a thread with 16 one-byte values then owns four
// aligned 16-byte vectors (80-byte stride between threads: the eight lanes of a 128-bit store phase hit all 32 banks).
__device__ __forceinline__ uint32_t pad_word(uint32_t i) { return i + 4u * (i >> 4); }
constexpr uint32_t kGatherOutBytes = (kRoundValues + kRoundValues / 16 + 1) * 8;

// 32-bit shared-memory addresses held in registers, and loads / stores through them: inside the per-value loop of the fast
// path the compiler otherwise rebuilds the base of every shared array from the thread index on each iteration (8 of the 49
// instructions of the old loop body, profiles/README.md r02 gather capture)
__device__ __forceinline__ uint32_t smem_addr_of(const void* p) {
  uint32_t a;
  asm volatile("{ .reg .u64 t; cvta.to.shared.u64 t, %1; cvt.u32.u64 %0, t; }" : "=r"(a) : "l"(p));
  return a;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }

// One 16-byte chunk per thread, 256 threads per 4 KiB tile.  (Two chunks per thread, which pays for the count kernel, measured
// 7-12 % SLOWER here - a thread's values come out of a serial loop over its terminators, and twice as many per thread doubled
// that chain while the scan and the barriers it amortises are a small part of this kernel - and 8 KiB tiles with 512 threads
// 25-30 % slower: a barrier over 16 warps costs more than it saves.)
constexpr int kGatherThreads = kDecTileBytes / 16;
template <int OUT>
__device__ __forceinline__ void gather_tile(const TileRef& r, uint8_t* __restrict__ dst, uint8_t* sm_out, uint8_t* sm_in, uint32_t* scan_smem, uint32_t* edge) {
  using out_t = typename OutTraits<OUT>::type;
  out_t* out = reinterpret_cast<out_t*>(sm_out);
  uint32_t* out32 = reinterpret_cast<uint32_t*>(sm_out);
  const uint64_t chunk_abs = r.tile_abs + threadIdx.x * 16ull;
  uint4 d, pd;
  uint32_t cont, valid, pcont;
  chunk_masks(chunk_abs, r.span_abs0, r.span_end, &d, &cont, &valid);
  // The 16 bytes before this thread's chunk: the previous lane has them.  Lane 0 of warps 1..7 gets the previous warp's
  // last lane through shared memory (its last word and continuation mask are all the fast path needs; the general path
  // reads the staged tile), only thread 0 has to go back to global memory for the bytes in front of the tile.
  pd.x = pd.y = pd.z = 0;  // (only thread 0's copy of the whole chunk is ever used: it stages the bytes in front of the tile)
  pd.w = __shfl_up_sync(0xFFFFFFFFu, d.w, 1);
  pcont = __shfl_up_sync(0xFFFFFFFFu, cont, 1);
  if (lane_id() == 31) {
    edge[2 * warp_id()] = d.w;
    edge[2 * warp_id() + 1] = cont;
  }
  if (threadIdx.x == 0) {
    uint32_t pv;
    chunk_masks(chunk_abs - 16, r.span_abs0, r.span_end, &pd, &pcont, &pv);
  }
  const uint32_t term = ~cont & valid;
  uint32_t total;
  const uint32_t idx0 = block_exclusive_scan<uint32_t, kGatherThreads / 32>(__popc(term), scan_smem, &total);  // two barriers
  if (lane_id() == 0 && warp_id() != 0) {
    pd.w = edge[2 * warp_id() - 2];
    pcont = edge[2 * warp_id() - 1];
  }
  const uint32_t c32 = (cont << 16) | pcont;
  const uint32_t run4 = c32 & (c32 >> 1) & (c32 >> 2) & (c32 >> 3);  // bit i: continuation bytes at i..i+3
  const bool fast = __syncthreads_and((run4 & 0xFFFFF000u) == 0);       // no value ending in this tile is longer than 4 bytes
  // every byte of the span inside this tile is a terminator <=> every value is ONE byte (masks, bools, small ids)
  const uint64_t v_lo = r.span_abs0 > r.tile_abs ? r.span_abs0 : r.tile_abs;
  const uint64_t v_hi = r.span_end < r.tile_abs + kDecTileBytes ? r.span_end : r.tile_abs + kDecTileBytes;
  const bool all1 = total == static_cast<uint32_t>(v_hi - v_lo);
  out_t* gdst = reinterpret_cast<out_t*>(dst);
  if (!fast) {
    *reinterpret_cast<uint4*>(sm_in + 16 + threadIdx.x * 16) = d;
    if (threadIdx.x == 0) *reinterpret_cast<uint4*>(sm_in) = pd;
  }
  for (uint32_t round0 = 0; round0 < total; round0 += kRoundValues) {  // uniform trip count (1, rarely 2)
  if (round0) __syncthreads();  // the previous round's flush has read the staging buffer
  if (all1) {  // masks, bools, small ids: the byte IS the value
    const uint32_t w4[4] = {d.x, d.y, d.z, d.w};
    uint32_t slot = idx0 - round0;  // wraps when before this round
    if (total == kDecTileBytes) {   // a full tile of one-byte values: every thread owns 16 consecutive slots
      if (slot < kRoundValues) {
        uint4* v4 = reinterpret_cast<uint4*>(out32 + pad_word(slot));
#pragma unroll
        for (int k = 0; k < 4; ++k)
          v4[k] = make_uint4(convert_word<OUT>(w4[k] & 0xFFu), convert_word<OUT>((w4[k] >> 8) & 0xFFu),
                             convert_word<OUT>((w4[k] >> 16) & 0xFFu), convert_word<OUT>(w4[k] >> 24));
      }
    } else {
#pragma unroll
      for (int pos = 0; pos < 16; ++pos) {
        if ((valid >> pos) & 1u) {
          if (slot < kRoundValues) out32[pad_word(slot)] = convert_word<OUT>((w4[pos >> 2] >> (8 * (pos & 3))) & 0xFFu);
          ++slot;
        }
      }
    }
  } else if (fast && total <= kRoundValues) {
    // The common fast case (2-3 byte values: at most 2048 of them in a tile, one round).  As below, with what the profile of
    // the loop asked for: shared addresses kept in registers, the length of a value from the position of the terminator
    // before it (one subtraction) instead of from the continuation bits, values staged densely (no slot padding: a
    // thread's values are consecutive and the threads of a warp start 5-8 slots apart), no per-value round test.
    const uint32_t wa = smem_addr_of(sm_in) + threadIdx.x * 20u;
    sts_u32(wa, pd.w);
    sts_u32(wa + 4u, d.x);
    sts_u32(wa + 8u, d.y);
    sts_u32(wa + 12u, d.z);
    sts_u32(wa + 16u, d.w);
    // the value that ends at the first terminator of the chunk starts after the terminator before it: in the previous
    // chunk (its trailing continuation bytes, at most 3 here), or at the first byte of the span if that lies in this chunk
    int32_t jprev;
    {
      const uint32_t first_valid = static_cast<uint32_t>(__ffs(static_cast<int>(valid))) - 1u;  // (term != 0 => valid != 0)
      const uint32_t lead = static_cast<uint32_t>(__clz(static_cast<int>(~(pcont << 16))));     // continuation bytes that end the previous chunk
      jprev = static_cast<int32_t>(first_valid) - 1 - (first_valid == 0u ? static_cast<int32_t>(lead) : 0);
    }
    uint32_t oa = smem_addr_of(sm_out) + 4u * idx0;
    uint32_t t = term;
    while (t) {
      const int32_t j = __ffs(static_cast<int>(t)) - 1;
      t &= t - 1;
      const uint32_t len = static_cast<uint32_t>(j - jprev);
      jprev = j;
      const uint32_t q = static_cast<uint32_t>(j) + 1u;  // offset of the first of the 4 bytes that end at the terminator, in the 20-byte strip
      const uint32_t a = wa + (q & ~3u);
      const uint32_t x = __funnelshift_r(lds_u32(a), lds_u32(a + 4u), q << 3);  // (the shift is taken mod 32)
      uint32_t g = x & 0x7F7F7F7Fu;
      g = (g & 0x007F007Fu) | ((g & 0x7F007F00u) >> 1);
      g = (g & 0x00003FFFu) | ((g & 0x3FFF0000u) >> 2);
      sts_u32(oa, convert_word<OUT>(g >> (28u - 7u * len)));
      oa += 4u;
    }
  } else if (fast) {
    // The 20 bytes {last word before the chunk, the chunk} go to a thread-private strip of shared memory so that the
    // loop can run over the terminators that exist (5-7 per chunk for 2-3 byte values) instead of over all 16 byte
    // positions: the 4 bytes ending at terminator j are two word loads + one funnel shift.
    uint32_t* win = reinterpret_cast<uint32_t*>(sm_in) + threadIdx.x * 5u;
    if (round0 == 0) {
      win[0] = pd.w;
      win[1] = d.x;
      win[2] = d.y;
      win[3] = d.z;
      win[4] = d.w;
    }
    uint32_t slot = idx0 - round0;  // wraps when before this round
    uint32_t t = term;
    while (t) {
      const uint32_t j = static_cast<uint32_t>(__ffs(t)) - 1u;
      t &= t - 1;
      if (slot < kRoundValues) {
        const uint32_t q = j + 1u;  // offset of the first of the 4 bytes inside the 20-byte strip
        const uint32_t x = __funnelshift_r(win[q >> 2], win[(q >> 2) + 1u], 8u * (q & 3u));  // (q = 16: shift 0, the upper word is not used)
        const uint32_t t3 = (c32 >> (13u + j)) & 7u;  // continuation bits of the 3 bytes before the terminator
        const uint32_t len = 1u + static_cast<uint32_t>(__clz(static_cast<int>(~(t3 << 29))));
        uint32_t g = x & 0x7F7F7F7Fu;
        g = (g & 0x007F007Fu) | ((g & 0x7F007F00u) >> 1);
        g = (g & 0x00003FFFu) | ((g & 0x3FFF0000u) >> 2);
        out32[pad_word(slot)] = convert_word<OUT>(g >> (7u * (4u - len)));
      }
      ++slot;
    }
  } else {
    // general path (some value of the tile is longer than 4 bytes): the tile's bytes are in shared memory behind the 16
    // bytes that precede it.  Per terminator: its length from the run of continuation bits below it, the 12 bytes that
    // end at it as three registers (four aligned word loads + funnel shifts), the 7-bit groups packed 4 / 4 / 2 at a
    // time, and the groups of the previous value shifted out.
    __syncthreads();
    // (shared addresses in registers, as in the fast path: the compiler rebuilds them from the thread index otherwise)
    const uint32_t ta = smem_addr_of(sm_in) + 16u + threadIdx.x * 16u - 11u;  // byte address of "11 bytes before this chunk"
    const uint32_t out_a = smem_addr_of(sm_out);
    uint32_t t = term, idx = idx0;
    while (t) {
      const uint32_t j = static_cast<uint32_t>(__ffs(t)) - 1u;
      t &= t - 1;
      const uint32_t slot = idx++ - round0;
      if (slot >= kRoundValues) continue;  // belongs to another round
      const uint32_t run = static_cast<uint32_t>(__clz(static_cast<int>(~(c32 << (16u - j)))));  // continuation bytes right below it
      const uint32_t len = 1u + (run < 9u ? run : 9u);
      const uint32_t a = ta + j;  // first of the 12 bytes that end at the terminator
      const uint32_t wa4 = a & ~3u, sh = a << 3;  // (funnel shifts take the amount mod 32)
      const uint32_t W0 = lds_u32(wa4), W1 = lds_u32(wa4 + 4u), W2 = lds_u32(wa4 + 8u), W3 = lds_u32(wa4 + 12u);
      const uint32_t x0 = __funnelshift_r(W0, W1, sh), x1 = __funnelshift_r(W1, W2, sh), x2 = __funnelshift_r(W2, W3, sh);
      auto pack4 = [](uint32_t x) {
        uint32_t g = x & 0x7F7F7F7Fu;
        g = (g & 0x007F007Fu) | ((g & 0x7F007F00u) >> 1);
        return (g & 0x00003FFFu) | ((g & 0x3FFF0000u) >> 2);
      };
      // the value is the top `len` groups of the 70-bit string (x0's two oldest bytes, x1, x2): pick the window by length
      // so that nothing is shifted out of 64 bits before it is needed (only a 10-byte value may wrap, as in prost)
      const uint64_t A = pack4(x2), B = pack4(x1);
      uint64_t v;
      if (len <= 4u) {
        v = A >> (7u * (4u - len));
      } else if (len <= 8u) {
        v = ((A << 28) | B) >> (7u * (8u - len));
      } else {
        const uint32_t hi2 = x0 >> 16;  // the two oldest bytes
        const uint64_t c = (hi2 & 0x7Fu) | ((hi2 & 0x7F00u) >> 1);
        v = len == 9u ? ((A << 35) | (B << 7) | (c >> 7)) : ((A << 42) | (B << 14) | c);
      }
      const out_t ov = convert_out<OUT>(v);
      const uint32_t oa = out_a + pad_slot(slot) * static_cast<uint32_t>(sizeof(out_t));
      if constexpr (sizeof(out_t) == 8) asm volatile("st.shared.u64 [%0], %1;" ::"r"(oa), "l"(static_cast<unsigned long long>(ov)) : "memory");
      else if constexpr (sizeof(out_t) == 4) sts_u32(oa, static_cast<uint32_t>(ov));
      else asm volatile("st.shared.u8 [%0], %1;" ::"r"(oa), "r"(static_cast<uint32_t>(ov)) : "memory");
    }
  }
  __syncthreads();
  // coalesced flush of this round's values
  const uint32_t n_round = total - round0 < kRoundValues ? total - round0 : kRoundValues;
  if (fast && !all1 && total <= kRoundValues) {  // staged densely as 32-bit words
    for (uint32_t i = threadIdx.x; i < n_round; i += kGatherThreads) gdst[i] = static_cast<out_t>(out32[i]);
  } else if (fast) {  // (all1 implies fast) staged as 32-bit words, padded
    for (uint32_t i = threadIdx.x; i < n_round; i += kGatherThreads) gdst[round0 + i] = static_cast<out_t>(out32[pad_word(i)]);
  } else {
    for (uint32_t i = threadIdx.x; i < n_round; i += kGatherThreads) gdst[round0 + i] = out[pad_slot(i)];
  }
  }
}

__global__ void __launch_bounds__(kGatherThreads, 6) k_varint_gather(DecTables T, uint8_t* pool, const uint64_t* __restrict__ gate) {
  __shared__ __align__(16) uint8_t sm_out[kGatherOutBytes];
  // general path: 16 bytes before the tile, the tile, slack for word loads; fast path: a 20-byte strip per thread (+ 1 word)
  __shared__ __align__(16) uint8_t sm_in[kGatherThreads * 20 + 16];
  static_assert(kGatherThreads * 20 + 16 >= 16 + kDecTileBytes + 16, "the staged tile must fit too");
  __shared__ uint32_t scan_smem[kGatherThreads / 32 + 1];
  __shared__ uint32_t edge[2 * (kGatherThreads / 32)];  // last word and continuation mask of every warp's last lane
  if (gate && *gate != 0) return;  // the pool of a one-call decode turned out to be too small
  const uint32_t tile = blockIdx.x;
  const TileRef r = tile_ref(T, tile);       // two independent broadcast loads
  const TileDst td = T.dtile_dst[tile];
  (void)pool;
  uint8_t* dst = td.dst;
  const int32_t ty = td.type;
  switch (ty) {
    case TFS_T_I32: case TFS_T_U32: gather_tile<O_U32>(r, dst, sm_out, sm_in, scan_smem, edge); break;
    case TFS_T_I64: case TFS_T_U64: gather_tile<O_U64>(r, dst, sm_out, sm_in, scan_smem, edge); break;
    case TFS_T_BOOL: gather_tile<O_BOOL>(r, dst, sm_out, sm_in, scan_smem, edge); break;
    case TFS_T_F32:
      if (td.dtype == 19) gather_tile<O_F16_F32>(r, dst, sm_out, sm_in, scan_smem, edge);
      else gather_tile<O_BF16_F32>(r, dst, sm_out, sm_in, scan_smem, edge);
      break;
    default: break;
  }
}

}  // namespace
}  // namespace tfs
