// tfs_ragged.cuh — `repeated bytes` fields (TensorProto.string_val, tensor.proto:57) on the ENCODE side.
//
// prost writes one record per element, never packed: key byte, varint(len), the bytes
// (`bytes::encode_repeated`, [3P prost 0.6.0]).  The caller hands the elements over flattened (TFS_T_RAGGED:
// (n + 1) u64 offsets, then the bytes).  Three kernels:
//   K8a k_ragged_size   per tile of 2048 elements: the wire bytes of its records (and the validity of the offsets)
//   K8b k_ragged_scan   per field: exclusive scan of the tile sizes -> record area length (EncField::body_len)
//   K8c k_ragged_copy   OUTPUT-centric: one CTA per 64 KiB of a field's record area, whatever the element sizes
//                       (a tile of two-byte labels and a single 100 MB blob are balanced the same way); every
//                       thread assembles aligned 16-byte chunks of the destination, header bytes computed,
//                       payload bytes fetched (one unaligned 16-byte fetch when the chunk lies inside one element).
// HBM-bound byte work: no tensor cores.
#pragma once
#include "tfs_device.cuh"

namespace tfs {

constexpr int kRagThreads = 256;
static_assert(kRagThreads == kCopyThreads, "cta_copy_bytes is written for kCopyThreads threads");
constexpr int kRagTile = 2048;      // elements per sizing tile
constexpr int kRagOutTile = 65536;  // record-area bytes per copy CTA (amortises the ~10 dependent table reads each CTA starts with)

__device__ __forceinline__ uint64_t rag_rec_len(uint64_t len) { return 1u + varint_len(len) + len; }

__global__ void __launch_bounds__(kRagThreads) k_ragged_size(EncTables T) {
  __shared__ uint64_t red[kRagThreads / 32];
  const uint32_t t = blockIdx.x;
  const uint32_t r = find_segment(T.rfield_first_tile, T.n_rfields, t);
  const EncField& f = T.fields[T.rfield_index[r]];
  const uint64_t e0 = static_cast<uint64_t>(t - T.rfield_first_tile[r]) * kRagTile;
  const uint32_t n = static_cast<uint32_t>(f.n_elems - e0 < kRagTile ? f.n_elems - e0 : kRagTile);
  const uint64_t* __restrict__ offs = static_cast<const uint64_t*>(f.src);
  uint64_t sum = 0;
  bool bad = false;
  for (uint32_t i = threadIdx.x; i < n; i += kRagThreads) {
    const uint64_t a = offs[e0 + i], b = offs[e0 + i + 1];
    bad |= b < a;
    sum += rag_rec_len(b >= a ? b - a : 0);
  }
  if (e0 == 0 && threadIdx.x == 0) bad |= offs[0] != 0 || offs[f.n_elems] != f.aux;
  if (bad) atomicOr(reinterpret_cast<unsigned long long*>(&T.result[2]), 1ull);  // descriptor and memory disagree
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, d);
  if (lane_id() == 0) red[warp_id()] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint64_t s = 0;
#pragma unroll
    for (int w = 0; w < kRagThreads / 32; ++w) s += red[w];
    T.rtile_bytes[t] = s;
  }
}

// one warp per ragged field
__global__ void __launch_bounds__(256) k_ragged_scan(EncTables T) {
  const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= T.n_rfields) return;
  const uint32_t t0 = T.rfield_first_tile[r], t1 = T.rfield_first_tile[r + 1];
  uint64_t carry = 0;
  for (uint32_t base = t0; base < t1; base += 32) {
    const uint32_t t = base + lane_id();
    const uint64_t v = t < t1 ? T.rtile_bytes[t] : 0ull;
    const uint64_t incl = warp_inclusive_scan(v);
    if (t < t1) T.rtile_off[t] = carry + incl - v;
    carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
  }
  if (lane_id() == 0) T.fields[T.rfield_index[r]].body_len = carry;
}

// 16 bytes from an arbitrary address with two aligned vector loads (the second only when the bytes straddle it)
__device__ __forceinline__ uint4 load16_unaligned(const uint8_t* p) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const uint4 lo = *reinterpret_cast<const uint4*>(a & ~static_cast<uintptr_t>(15));
  const uint32_t sh = static_cast<uint32_t>(a & 15u);
  if (sh == 0) return lo;
  const uint4 hi = *reinterpret_cast<const uint4*>((a & ~static_cast<uintptr_t>(15)) + 16);
  const uint32_t k = sh >> 2, r = (sh & 3u) * 8u;
  uint32_t w0, w1, w2, w3, w4;
  switch (k) {
    case 0: w0 = lo.x; w1 = lo.y; w2 = lo.z; w3 = lo.w; w4 = hi.x; break;
    case 1: w0 = lo.y; w1 = lo.z; w2 = lo.w; w3 = hi.x; w4 = hi.y; break;
    case 2: w0 = lo.z; w1 = lo.w; w2 = hi.x; w3 = hi.y; w4 = hi.z; break;
    default: w0 = lo.w; w1 = hi.x; w2 = hi.y; w3 = hi.z; w4 = hi.w; break;
  }
  return make_uint4(__funnelshift_r(w0, w1, r), __funnelshift_r(w1, w2, r), __funnelshift_r(w2, w3, r), __funnelshift_r(w3, w4, r));
}

__global__ void __launch_bounds__(kRagThreads, 3) k_ragged_copy(EncTables T, uint8_t* __restrict__ dst) {
  __shared__ uint64_t s_rec[kRagTile + 1];  // start of every record of the current element tile, inside the record area
  __shared__ uint64_t s_scan[kRagThreads / 32 + 1];
  if (T.result[1] != 0 || T.result[2] != 0) return;  // dst too small, or offsets that do not match the descriptor
  const uint32_t ot = blockIdx.x;
  const uint32_t r = find_segment(T.rfield_first_out, T.n_rfields, ot);
  const EncField& f = T.fields[T.rfield_index[r]];
  const uint64_t a = static_cast<uint64_t>(ot - T.rfield_first_out[r]) * kRagOutTile;
  if (a >= f.body_len) return;  // the grid was sized from an upper bound
  const uint64_t b = a + kRagOutTile < f.body_len ? a + kRagOutTile : f.body_len;
  const uint64_t* __restrict__ offs = static_cast<const uint64_t*>(f.src);
  const uint8_t* __restrict__ bytes = static_cast<const uint8_t*>(f.src) + 8ull * (f.n_elems + 1);
  uint8_t* const out = dst + f.dst_off;
  const uintptr_t D = reinterpret_cast<uintptr_t>(out);
  const uint8_t key = static_cast<uint8_t>((static_cast<uint32_t>(f.kind) << 3) | 2u);
  const uint32_t t_first = T.rfield_first_tile[r], n_tiles = T.rfield_first_tile[r + 1] - t_first;
  uint32_t et = find_segment(T.rtile_off + t_first, n_tiles, a);  // element tile that holds byte `a`
  uint64_t cursor = a;
  while (cursor < b) {
    const uint64_t e0 = static_cast<uint64_t>(et) * kRagTile;
    const uint32_t ne = static_cast<uint32_t>(f.n_elems - e0 < kRagTile ? f.n_elems - e0 : kRagTile);
    const uint64_t tile_begin = T.rtile_off[t_first + et];
    // record starts of this element tile: coalesced loads, blocked scan (8 elements per thread)
    for (uint32_t i = threadIdx.x; i < kRagTile; i += kRagThreads) s_rec[i] = i < ne ? rag_rec_len(offs[e0 + i + 1] - offs[e0 + i]) : 0ull;
    __syncthreads();
    uint64_t mine[8], sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      mine[j] = s_rec[threadIdx.x * 8 + j];
      sum += mine[j];
    }
    uint64_t total;
    uint64_t run = tile_begin + block_exclusive_scan<uint64_t, kRagThreads / 32>(sum, s_scan, &total);  // syncs twice
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s_rec[threadIdx.x * 8 + j] = run;
      run += mine[j];
    }
    if (threadIdx.x == kRagThreads - 1) s_rec[kRagTile] = run;
    __syncthreads();
    const uint64_t tile_end = tile_begin + total;
    const uint64_t lo = cursor, hi = b < tile_end ? b : tile_end;
    {  // few, large elements in the range (encoded images, serialized Examples): element by element, the header bytes by
       // a handful of threads, the payload piece by a CTA-wide copy (any alignment on both sides)
      const uint32_t i0 = find_segment(s_rec, ne, lo), i1 = find_segment(s_rec, ne, hi - 1);  // uniform across the CTA
      if (i1 - i0 < 16u) {
        for (uint32_t idx = i0; idx <= i1; ++idx) {
          const uint64_t r0 = s_rec[idx], so = offs[e0 + idx];
          const uint64_t ln = offs[e0 + idx + 1] - so;
          const uint32_t h0 = 1u + varint_len(ln);
          if (threadIdx.x < h0) {
            const uint64_t p = r0 + threadIdx.x;
            if (p >= lo && p < hi)
              out[p] = threadIdx.x == 0 ? key
                                        : static_cast<uint8_t>(((ln >> (7u * (threadIdx.x - 1u))) & 0x7Fu) | (threadIdx.x + 1u < h0 ? 0x80u : 0u));
          }
          const uint64_t pa = lo > r0 + h0 ? lo : r0 + h0, pb = hi < r0 + h0 + ln ? hi : r0 + h0 + ln;
          if (pa < pb) cta_copy_bytes<4>(out + pa, bytes + so + (pa - r0 - h0), static_cast<uint32_t>(pb - pa));
        }
        cursor = hi;
        ++et;
        __syncthreads();
        continue;
      }
    }
    // aligned 16-byte chunks of the destination (absolute addresses) that intersect [lo, hi)
    const uint64_t A_lo = D + lo, A_hi = D + hi;
    const uint64_t chunk0 = A_lo & ~static_cast<uint64_t>(15);
    const uint32_t n_chunks = static_cast<uint32_t>((A_hi - chunk0 + 15) >> 4);
    for (uint32_t q = threadIdx.x; q < n_chunks; q += kRagThreads) {
      const uint64_t cs = chunk0 + 16ull * q;
      const uint64_t s = (cs > A_lo ? cs : A_lo) - D, e = (cs + 16 < A_hi ? cs + 16 : A_hi) - D;
      if (s >= e) continue;
      uint32_t idx = find_segment(s_rec, ne, s);
      uint64_t rec_start = s_rec[idx];
      uint64_t src_off = offs[e0 + idx];
      uint64_t len = offs[e0 + idx + 1] - src_off;
      uint32_t hdr = 1u + varint_len(len);
      if (e - s == 16 && s - rec_start >= hdr && e <= rec_start + hdr + len) {  // wholly inside one element's bytes
        *reinterpret_cast<uint4*>(out + s) = load16_unaligned(bytes + src_off + (s - rec_start - hdr));
        continue;
      }
      uint32_t w[4] = {0u, 0u, 0u, 0u};
      for (uint64_t p = s; p < e; ++p) {
        uint64_t rel = p - rec_start;
        if (rel >= hdr + len) {  // next record (records are never empty: at least key + length byte)
          ++idx;
          rec_start = s_rec[idx];
          src_off = offs[e0 + idx];
          len = offs[e0 + idx + 1] - src_off;
          hdr = 1u + varint_len(len);
          rel = 0;
        }
        uint32_t v;
        if (rel == 0) v = key;
        else if (rel < hdr) v = static_cast<uint32_t>((len >> (7u * (static_cast<uint32_t>(rel) - 1u))) & 0x7Fu) | (rel + 1 < hdr ? 0x80u : 0u);
        else v = bytes[src_off + rel - hdr];
        const uint32_t k = static_cast<uint32_t>(p - s);
        w[k >> 2] |= v << ((k & 3u) * 8u);
      }
      if (e - s == 16) {
        *reinterpret_cast<uint4*>(out + s) = make_uint4(w[0], w[1], w[2], w[3]);
      } else {
        for (uint32_t k = 0; k < static_cast<uint32_t>(e - s); ++k) out[s + k] = static_cast<uint8_t>(w[k >> 2] >> ((k & 3u) * 8u));
      }
    }
    cursor = hi;
    ++et;
    __syncthreads();  // s_rec is rebuilt for the next element tile
  }
}

}  // namespace tfs
