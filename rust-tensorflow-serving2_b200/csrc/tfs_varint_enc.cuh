// tfs_varint_enc.cuh — K1 (size) and K3 (scatter) of the packed-varint encoder.
//
// Replaces prost's {int32,int64,uint32,uint64}::{encoded_len_packed, encode_packed} and encode_varint for
// int_val / int64_val / half_val / uint32_val / uint64_val (tensor.proto:45,54,64,80,83; reached from the
// reference at src/lib.rs:265).  Integer work bound by HBM bandwidth; no tensor cores.
//
// Work unit: a tile of 2048 consecutive elements of one field, described by a 16-byte record (VTile) that a
// tiny kernel builds once per batch, so a CTA starts with ONE broadcast load instead of a search.
//   K1 k_varint_size     coalesced row loads, length from clz, warp + CTA sum -> bytes per tile.
//   K3 k_varint_scatter  the tile is staged in shared memory with cp.async (bank-swizzled), every thread then
//                        owns 8 CONSECUTIVE elements.  Fast path (a full tile whose values are all < 2^28, i.e.
//                        token ids, half patterns, small ints): the 7-bit groups are spread into bytes with
//                        three mask/shift steps, a thread's 8 varints are concatenated in a 64-bit register
//                        window and leave as aligned 32-bit shared-memory stores; the word shared with the
//                        neighbouring thread is merged through one warp shuffle.  No per-byte loop, no
//                        data-dependent branch.  Anything else (negatives, wide values, ragged last tile,
//                        unaligned source) takes the general path: ballot/popc row prefix + a byte loop.
//                        Either way the tile's bytes sit in shared memory at `address mod 16` of their final
//                        place and are flushed with 16-byte stores.
#pragma once
#include "tfs_device.cuh"

namespace tfs {
namespace {

__device__ __forceinline__ uint32_t vmode_esize(int32_t mode) {
  switch (mode) {
    case VM_I64: return 8;
    case VM_I32: case VM_U32: case VM_F32_F16: case VM_F32_BF16: return 4;
    case VM_I16: case VM_U16: return 2;
    default: return 1;
  }
}

// ---- tile records -------------------------------------------------------------------------------------
// one warp per varint field: its tiles' records
__global__ void __launch_bounds__(256) k_tile_map(EncTables T) {
  const uint32_t vf = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (vf >= T.n_vfields) return;
  const uint32_t fi = T.vfield_index[vf];
  const EncField& f = T.fields[fi];
  const uint32_t t0 = T.vfield_first_tile[vf];
  const uint32_t nt = T.vfield_first_tile[vf + 1] - t0;
  const uint32_t es = vmode_esize(f.mode);
  for (uint32_t i = lane_id(); i < nt; i += 32) {
    VTile r;
    r.src = static_cast<const uint8_t*>(f.src) + static_cast<uint64_t>(i) * kVarintTile * es;
    const uint64_t left = f.n_elems - static_cast<uint64_t>(i) * kVarintTile;
    r.n = left < kVarintTile ? static_cast<uint32_t>(left) : kVarintTile;
    r.mode = static_cast<uint16_t>(f.mode);
    r.esize = static_cast<uint16_t>(es);
    T.vtiles[t0 + i] = r;
    T.tile_field[t0 + i] = fi;
  }
}

// ---- lengths --------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t len_from_bits(uint32_t bits) { return ((bits + 6u) * 37u) >> 8; }  // ceil(bits / 7), bits in 1..64
__device__ __forceinline__ uint32_t varint_len_u32(uint32_t v) { return len_from_bits(32u - __clz(static_cast<int>(v | 1u))); }
__device__ __forceinline__ uint32_t varint_len_u64(uint32_t lo, uint32_t hi) {
  // branch-free: a data-dependent branch here makes the compiler serialise the loads of an unrolled loop
  const uint32_t chi = static_cast<uint32_t>(__clz(static_cast<int>(hi)));         // 32 when hi == 0
  const uint32_t clo = static_cast<uint32_t>(__clz(static_cast<int>(lo | 1u)));
  const uint32_t bits = hi ? 64u - chi : 32u - clo;  // a select, both sides are already computed
  return len_from_bits(bits);
}

// element `i` of a tile as the (lo, hi) halves of the u64 prost would varint-code (`as u64`: sign extension)
template <int MODE>
__device__ __forceinline__ void load_lohi(const uint8_t* __restrict__ src, uint32_t i, uint32_t* lo, uint32_t* hi) {
  if constexpr (MODE == VM_I64) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(src) + i);
    *lo = v.x;
    *hi = v.y;
  } else if constexpr (MODE == VM_I32) {
    const int32_t v = __ldg(reinterpret_cast<const int32_t*>(src) + i);
    *lo = static_cast<uint32_t>(v);
    *hi = v < 0 ? 0xFFFFFFFFu : 0u;
  } else if constexpr (MODE == VM_U32) {
    *lo = __ldg(reinterpret_cast<const uint32_t*>(src) + i);
    *hi = 0;
  } else if constexpr (MODE == VM_I16) {
    const int32_t v = __ldg(reinterpret_cast<const int16_t*>(src) + i);
    *lo = static_cast<uint32_t>(v);
    *hi = v < 0 ? 0xFFFFFFFFu : 0u;
  } else if constexpr (MODE == VM_U16) {
    *lo = __ldg(reinterpret_cast<const uint16_t*>(src) + i);
    *hi = 0;
  } else if constexpr (MODE == VM_I8) {
    const int32_t v = __ldg(reinterpret_cast<const int8_t*>(src) + i);
    *lo = static_cast<uint32_t>(v);
    *hi = v < 0 ? 0xFFFFFFFFu : 0u;
  } else if constexpr (MODE == VM_U8) {
    *lo = __ldg(src + i);
    *hi = 0;
  } else if constexpr (MODE == VM_F32_F16) {
    *lo = f32_to_f16_bits(__ldg(reinterpret_cast<const float*>(src) + i));
    *hi = 0;
  } else {
    *lo = f32_to_bf16_bits(__ldg(reinterpret_cast<const float*>(src) + i));
    *hi = 0;
  }
}

// ---- K1 ---------------------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ uint32_t size_rows(const VTile& r) {
  uint32_t s = 0;
  if (r.n == kVarintTile) {
    uint32_t lo[kVarintRows], hi[kVarintRows];
#pragma unroll
    for (int k = 0; k < kVarintRows; ++k) load_lohi<MODE>(r.src, k * kVarintThreads + threadIdx.x, &lo[k], &hi[k]);  // 8 loads in flight
#pragma unroll
    for (int k = 0; k < kVarintRows; ++k) s += varint_len_u64(lo[k], hi[k]);
  } else {
    for (uint32_t i = threadIdx.x; i < r.n; i += kVarintThreads) {
      uint32_t lo, hi;
      load_lohi<MODE>(r.src, i, &lo, &hi);
      s += varint_len_u64(lo, hi);
    }
  }
  return s;
}

__global__ void __launch_bounds__(kVarintThreads) k_varint_size(EncTables T) {
  __shared__ uint32_t warp_tot[kVarintThreads / 32];
  const uint32_t tile = blockIdx.x;
  const VTile r = T.vtiles[tile];
  uint32_t s = 0;
  switch (r.mode) {
    case VM_I64: s = size_rows<VM_I64>(r); break;
    case VM_I32: s = size_rows<VM_I32>(r); break;
    case VM_U32: s = size_rows<VM_U32>(r); break;
    case VM_I16: s = size_rows<VM_I16>(r); break;
    case VM_U16: s = size_rows<VM_U16>(r); break;
    case VM_I8: s = size_rows<VM_I8>(r); break;
    case VM_U8: s = size_rows<VM_U8>(r); break;
    case VM_F32_F16: s = size_rows<VM_F32_F16>(r); break;
    case VM_F32_BF16: s = size_rows<VM_F32_BF16>(r); break;
    default: break;
  }
  s = __reduce_add_sync(0xFFFFFFFFu, s);
  if (lane_id() == 0) warp_tot[warp_id()] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < kVarintThreads / 32; ++w) t += warp_tot[w];
    T.tile_bytes[tile] = t;
  }
}

// ---- K3 ---------------------------------------------------------------------------------------------------
constexpr uint32_t kStageIn = kVarintTile * 8;                       // staged input, at most 8 bytes per element
constexpr uint32_t kStageOut = 16 + kVarintTile * kVarintMaxLen + 16;  // output window incl. alignment skew

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// 16-byte chunk c of the staged tile lives at chunk (c ^ ((c >> 3) & 3)): threads that own consecutive groups of
// chunks then hit distinct banks when they read them back with 128-bit loads
__device__ __forceinline__ uint32_t swz(uint32_t c) { return c ^ ((c >> 3) & 3u); }

// 28-bit value -> its varint bytes in a 32-bit word (len bytes valid, the rest zero)
__device__ __forceinline__ uint32_t varint_word28(uint32_t v, uint32_t len) {
  uint32_t x = ((v & 0x0FFFC000u) << 2) | (v & 0x00003FFFu);
  x = ((x & 0x3F803F80u) << 1) | (x & 0x007F007Fu);
  return x | (0x00808080u >> (8u * (4u - len)));
}

// general path: row arrangement straight from global memory, ballot/popc prefix, byte loop
template <int MODE>
__device__ __forceinline__ uint32_t scatter_general(const VTile& r, uint8_t* stage, uint32_t skew, uint32_t* warp_tot) {
  uint32_t lo[kVarintRows], hi[kVarintRows], len[kVarintRows];
  const uint32_t base = warp_id() * (kVarintRows * 32u) + lane_id();
#pragma unroll
  for (int k = 0; k < kVarintRows; ++k) {  // all loads first
    const uint32_t i = base + k * 32u;
    lo[k] = hi[k] = 0;
    if (i < r.n) load_lohi<MODE>(r.src, i, &lo[k], &hi[k]);
  }
#pragma unroll
  for (int k = 0; k < kVarintRows; ++k) len[k] = (base + k * 32u < r.n) ? varint_len_u64(lo[k], hi[k]) : 0u;
  const uint32_t lt = (1u << lane_id()) - 1u;
  uint32_t off[kVarintRows];
  uint32_t run = 0;
#pragma unroll
  for (int k = 0; k < kVarintRows; ++k) {
    const uint32_t b0 = __ballot_sync(0xFFFFFFFFu, len[k] & 1u), b1 = __ballot_sync(0xFFFFFFFFu, len[k] & 2u);
    const uint32_t b2 = __ballot_sync(0xFFFFFFFFu, len[k] & 4u), b3 = __ballot_sync(0xFFFFFFFFu, len[k] & 8u);
    off[k] = run + __popc(b0 & lt) + 2u * __popc(b1 & lt) + 4u * __popc(b2 & lt) + 8u * __popc(b3 & lt);
    run += __popc(b0) + 2u * __popc(b1) + 4u * __popc(b2) + 8u * __popc(b3);
  }
  if (lane_id() == 0) warp_tot[warp_id()] = run;
  __syncthreads();
  uint32_t wbase = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kVarintThreads / 32; ++w) {
    const uint32_t t = warp_tot[w];
    if (w < static_cast<int>(warp_id())) wbase += t;
    total += t;
  }
#pragma unroll
  for (int k = 0; k < kVarintRows; ++k) {
    if (len[k]) {
      uint8_t* p = stage + skew + wbase + off[k];
      uint64_t x = (static_cast<uint64_t>(hi[k]) << 32) | lo[k];
      while (x >= 0x80) {
        *p++ = static_cast<uint8_t>((x & 0x7F) | 0x80);
        x >>= 7;
      }
      *p = static_cast<uint8_t>(x);
    }
  }
  return total;
}

// value j (0..7) of this thread out of the staged, swizzled tile; returns false if it does not fit the fast path
template <int MODE>
__device__ __forceinline__ bool staged_values(const uint8_t* in, uint32_t (&v)[8]) {
  const uint32_t t = threadIdx.x;
  bool ok = true;
  if constexpr (MODE == VM_I64) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const uint4 q = *reinterpret_cast<const uint4*>(in + 16u * swz(4u * t + p));
      v[2 * p] = q.x;
      v[2 * p + 1] = q.z;
      ok = ok && q.y == 0 && q.w == 0;
    }
  } else if constexpr (MODE == VM_I32 || MODE == VM_U32 || MODE == VM_F32_F16 || MODE == VM_F32_BF16) {
#pragma unroll
    uint32_t raw[8];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const uint4 q = *reinterpret_cast<const uint4*>(in + 16u * swz(2u * t + p));
      raw[4 * p] = q.x, raw[4 * p + 1] = q.y, raw[4 * p + 2] = q.z, raw[4 * p + 3] = q.w;
    }
    if constexpr (MODE == VM_F32_F16 || MODE == VM_F32_BF16) {
      // the hardware conversion for everything but NaN; a thread that holds a NaN (rare) redoes its eight values with the
      // payload-keeping form - the NaN arithmetic for every element was a tenth of this kernel's instructions
      bool nan = false;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float f = __uint_as_float(raw[j]);
        if constexpr (MODE == VM_F32_F16) v[j] = static_cast<uint32_t>(__half_as_ushort(__float2half_rn(f)));
        else v[j] = static_cast<uint32_t>(__bfloat16_as_ushort(__float2bfloat16_rn(f)));
        nan = nan || (raw[j] << 1) > 0xFF000000u;
      }
      if (nan) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if constexpr (MODE == VM_F32_F16) v[j] = f32_to_f16_bits(__uint_as_float(raw[j]));
          else v[j] = f32_to_bf16_bits(__uint_as_float(raw[j]));
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = raw[j];  // a negative i32 has bit 31 set and fails the range test below
    }
  } else {  // VM_U16: 8 elements = one chunk
    const uint4 q = *reinterpret_cast<const uint4*>(in + 16u * swz(t));
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v[2 * k] = w[k] & 0xFFFFu;
      v[2 * k + 1] = w[k] >> 16;
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) ok = ok && v[j] < (1u << 28);
  return ok;
}

constexpr uint32_t kUnskewed = 0x80000000u;  // flag in scatter_tile's return value: the bytes sit at stage + 16, not at stage + skew

// fast path: every thread concatenates the varints of its 8 consecutive elements in a register window and
// stores aligned 32-bit words
__device__ __forceinline__ uint32_t scatter_fast(const uint32_t (&v)[8], uint8_t* stage, uint32_t skew, uint32_t* warp_tot) {
  uint32_t len[8], tsum = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    len[j] = ((38u - static_cast<uint32_t>(__clz(static_cast<int>(v[j] | 1u)))) * 37u) >> 8;  // ceil(bits / 7) for values below 2^28
    tsum += len[j];
  }
  const uint32_t incl = warp_inclusive_scan(tsum);
  if (lane_id() == 31) warp_tot[warp_id()] = incl;
  __syncthreads();
  uint32_t wbase = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kVarintThreads / 32; ++w) {
    const uint32_t t = warp_tot[w];
    if (w < static_cast<int>(warp_id())) wbase += t;
    total += t;
  }
  if (total == static_cast<uint32_t>(kVarintTile)) {
    // every value of the tile is ONE byte (masks, segment ids, bools held as integers): the output is the low byte of
    // each value.  Staged without the destination skew, 8 bytes per thread in one store; the flush shifts (kUnskewed).
    uint2 w;
    w.x = v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24);
    w.y = v[4] | (v[5] << 8) | (v[6] << 16) | (v[7] << 24);
    *reinterpret_cast<uint2*>(stage + 16u + 8u * threadIdx.x) = w;
    return total | kUnskewed;
  }
  const uint32_t o = skew + wbase + (incl - tsum);  // byte offset of this thread's run in `stage`
  const uint32_t f0 = o & 3u;                        // bytes of the first word that belong to the thread before
  uint32_t* words = reinterpret_cast<uint32_t*>(stage) + (o >> 2);
  uint32_t acc_lo = 0, acc_hi = 0, nb = f0, emitted = 0, first_word = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t e = varint_word28(v[j], len[j]);
    const uint32_t s = 8u * nb;
    acc_lo |= e << s;
    acc_hi |= __funnelshift_l(e, 0u, s);  // the part of e that spills over bit 32 (0 when s == 0)
    nb += len[j];
    if (nb >= 4u) {  // a full word leaves (a thread's run is >= 8 bytes: at least two do)
      if (emitted == 0) first_word = acc_lo;
      else words[emitted] = acc_lo;
      ++emitted;
      acc_lo = acc_hi;
      acc_hi = 0;
      nb -= 4u;
    }  // (a select-only formulation of this block measured 10 % slower: more registers, spills)
  }
  // the word this thread shares with its predecessor: the predecessor's leftover bytes are exactly my f0 low
  // bytes (lane 0 and lane 31 meet another warp or the edge of the tile there: bytes are stored singly)
  const uint32_t prev_partial = __shfl_up_sync(0xFFFFFFFFu, acc_lo, 1);
  if (lane_id() != 0) {
    words[0] = first_word | prev_partial;
  } else {
    uint8_t* b = reinterpret_cast<uint8_t*>(words);
    for (uint32_t k = f0; k < 4u; ++k) b[k] = static_cast<uint8_t>(first_word >> (8u * k));
  }
  if (lane_id() == 31) {
    uint8_t* b = reinterpret_cast<uint8_t*>(words + emitted);
    for (uint32_t k = 0; k < nb; ++k) b[k] = static_cast<uint8_t>(acc_lo >> (8u * k));
  }
  return total;
}

template <int MODE>
__device__ __forceinline__ uint32_t scatter_tile(const VTile& r, uint8_t* smem_all, uint8_t** stage_used, uint32_t skew, uint32_t* warp_tot) {
  uint8_t* const stage_in = smem_all;
  uint8_t* stage = smem_all + kStageIn;  // fast path: behind the staged input
  *stage_used = stage;
  constexpr bool kFastMode = MODE == VM_I64 || MODE == VM_I32 || MODE == VM_U32 || MODE == VM_U16 || MODE == VM_F32_F16 || MODE == VM_F32_BF16;
  if constexpr (kFastMode) {
    // uniform over the CTA: a full tile from a 16-byte aligned source can be staged with cp.async
    if (r.n == kVarintTile && (reinterpret_cast<uintptr_t>(r.src) & 15u) == 0) {
      const uint32_t nchunks = kVarintTile * r.esize / 16u;
      for (uint32_t c = threadIdx.x; c < nchunks; c += kVarintThreads) cp_async16(stage_in + 16u * swz(c), r.src + 16u * c);
      cp_async_wait_all();
      __syncthreads();
      uint32_t v[8];
      const bool ok = staged_values<MODE>(stage_in, v);
      if (__syncthreads_and(ok)) return scatter_fast(v, stage, skew, warp_tot);
    }
  }
  __syncthreads();  // the staged input (if any) is dead: the general path reuses the whole buffer
  *stage_used = smem_all;
  return scatter_general<MODE>(r, smem_all, skew, warp_tot);
}

// Shared memory of one CTA: the fast path needs the staged input (16 KiB) plus at most 4 bytes per element of
// output; the general path needs no staged input but up to 10 bytes per element.  One buffer serves both.
constexpr uint32_t kFastOut = 16 + kVarintTile * 4 + 16;
constexpr uint32_t kScatterSmem = (kStageIn + kFastOut) > kStageOut ? (kStageIn + kFastOut) : kStageOut;

__global__ void __launch_bounds__(kVarintThreads, 6) k_varint_scatter(EncTables T, uint8_t* __restrict__ dst_base) {
  __shared__ __align__(16) uint8_t smem_all[kScatterSmem];
  __shared__ uint32_t warp_tot[kVarintThreads / 32];
  uint8_t* stage = smem_all;
  if (T.result[1] != 0) return;  // dst too small: nothing may be written
  const uint32_t tile = blockIdx.x;
  const VTile r = T.vtiles[tile];
  uint8_t* g = dst_base + T.fields[T.tile_field[tile]].dst_off + T.tile_off[tile];
  const uint32_t skew = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(g) & 15u);
  uint32_t total = 0;
  switch (r.mode) {
    case VM_I64: total = scatter_tile<VM_I64>(r, smem_all, &stage, skew, warp_tot); break;
    case VM_I32: total = scatter_tile<VM_I32>(r, smem_all, &stage, skew, warp_tot); break;
    case VM_U32: total = scatter_tile<VM_U32>(r, smem_all, &stage, skew, warp_tot); break;
    case VM_I16: total = scatter_tile<VM_I16>(r, smem_all, &stage, skew, warp_tot); break;
    case VM_U16: total = scatter_tile<VM_U16>(r, smem_all, &stage, skew, warp_tot); break;
    case VM_I8: total = scatter_tile<VM_I8>(r, smem_all, &stage, skew, warp_tot); break;
    case VM_U8: total = scatter_tile<VM_U8>(r, smem_all, &stage, skew, warp_tot); break;
    case VM_F32_F16: total = scatter_tile<VM_F32_F16>(r, smem_all, &stage, skew, warp_tot); break;
    case VM_F32_BF16: total = scatter_tile<VM_F32_BF16>(r, smem_all, &stage, skew, warp_tot); break;
    default: break;
  }
  __syncthreads();
  // flush: whole 16-byte words with vector stores, the ragged head / tail bytes one by one (a neighbouring
  // tile owns the other bytes of those words and writes only its own)
  const bool unskewed = (total & kUnskewed) != 0;
  total &= ~kUnskewed;
  uint8_t* ga = g - skew;
  const uint32_t end = skew + total;
  const uint32_t nchunks = (end + 15u) >> 4;
  if (!unskewed) {
    for (uint32_t c = threadIdx.x; c < nchunks; c += kVarintThreads) {
      const uint32_t lo = c << 4, hi = lo + 16u;
      if (lo >= skew && hi <= end) {
        *reinterpret_cast<uint4*>(ga + lo) = *reinterpret_cast<const uint4*>(stage + lo);
      } else {
        const uint32_t b0 = lo > skew ? lo : skew, b1 = hi < end ? hi : end;
        for (uint32_t b = b0; b < b1; ++b) ga[b] = stage[b];
      }
    }
  } else {  // a tile of one-byte values: stream byte i sits at stage + 16 + i; shift it to the destination's 16-byte grid
    for (uint32_t c = threadIdx.x; c < nchunks; c += kVarintThreads) {
      const uint32_t lo = c << 4, hi = lo + 16u;
      if (lo >= skew && hi <= end) {
        const uint32_t s0 = lo - skew, sh = s0 & 15u;
        const uint4 a = *reinterpret_cast<const uint4*>(stage + 16u + (s0 & ~15u));
        const uint4 b = *reinterpret_cast<const uint4*>(stage + 32u + (s0 & ~15u));
        *reinterpret_cast<uint4*>(ga + lo) = sh ? shift_window(a, b, sh) : a;
      } else {
        const uint32_t b0 = lo > skew ? lo : skew, b1 = hi < end ? hi : end;
        for (uint32_t b = b0; b < b1; ++b) ga[b] = stage[16u + b - skew];
      }
    }
  }
}

}  // namespace
}  // namespace tfs
