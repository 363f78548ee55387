// tfs_device.cuh — device-side building blocks (sm_100a).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "tfs_internal.h"

namespace tfs {

// ---- varint arithmetic (prost encoded_len_varint: max(1, ceil(bits / 7)); reference path src/lib.rs:265) ----
// (the framing arithmetic compiles for host and device alike; only device code calls it today)
__host__ __device__ __forceinline__ uint32_t varint_len(uint64_t v) {
#ifdef __CUDA_ARCH__
  return (70u - static_cast<uint32_t>(__clzll(static_cast<long long>(v | 1)))) / 7u;
#else
  return (70u - static_cast<uint32_t>(__builtin_clzll(v | 1))) / 7u;
#endif
}

__host__ __device__ __forceinline__ uint32_t key_len(int32_t field_number) { return field_number < 16 ? 1u : 2u; }

// ---- 16-bit float casts: round-to-nearest-even in hardware, NaN -> quiet NaN keeping sign + top payload ----
// (written as selects, not branches: a branch per element stops the compiler from batching the loads of an
// unrolled loop)
__device__ __forceinline__ uint32_t f32_to_f16_bits(float f) {
  const uint32_t x = __float_as_uint(f);
  const uint32_t nan_bits = ((x >> 16) & 0x8000u) | 0x7E00u | ((x >> 13) & 0x01FFu);
  const uint32_t cvt = static_cast<uint32_t>(__half_as_ushort(__float2half_rn(f)));
  return (x & 0x7FFFFFFFu) > 0x7F800000u ? nan_bits : cvt;
}
__device__ __forceinline__ uint32_t f32_to_bf16_bits(float f) {
  const uint32_t x = __float_as_uint(f);
  const uint32_t nan_bits = (x >> 16) | 0x0040u;
  const uint32_t cvt = static_cast<uint32_t>(__bfloat16_as_ushort(__float2bfloat16_rn(f)));
  return (x & 0x7FFFFFFFu) > 0x7F800000u ? nan_bits : cvt;
}
__device__ __forceinline__ float f16_bits_to_f32(uint32_t h) { return __half2float(__ushort_as_half(static_cast<unsigned short>(h))); }
__device__ __forceinline__ float bf16_bits_to_f32(uint32_t h) { return __uint_as_float(h << 16); }

// ---- element fetch for the varint encoder ---------------------------------------------------------
template <int MODE>
__device__ __forceinline__ uint64_t load_varint_value(const void* __restrict__ src, uint64_t i) {
  if constexpr (MODE == VM_I64) return static_cast<uint64_t>(__ldg(static_cast<const long long*>(src) + i));
  else if constexpr (MODE == VM_I32) return static_cast<uint64_t>(static_cast<int64_t>(__ldg(static_cast<const int32_t*>(src) + i)));
  else if constexpr (MODE == VM_U32) return static_cast<uint64_t>(__ldg(static_cast<const uint32_t*>(src) + i));
  else if constexpr (MODE == VM_I16) return static_cast<uint64_t>(static_cast<int64_t>(__ldg(static_cast<const int16_t*>(src) + i)));
  else if constexpr (MODE == VM_U16) return static_cast<uint64_t>(__ldg(static_cast<const uint16_t*>(src) + i));
  else if constexpr (MODE == VM_I8) return static_cast<uint64_t>(static_cast<int64_t>(__ldg(static_cast<const int8_t*>(src) + i)));
  else if constexpr (MODE == VM_U8) return static_cast<uint64_t>(__ldg(static_cast<const uint8_t*>(src) + i));
  else if constexpr (MODE == VM_F32_F16) return static_cast<uint64_t>(f32_to_f16_bits(__ldg(static_cast<const float*>(src) + i)));
  else return static_cast<uint64_t>(f32_to_bf16_bits(__ldg(static_cast<const float*>(src) + i)));
}

// ---- warp / block scans -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t warp_id() { return threadIdx.x >> 5; }

template <class T>
__device__ __forceinline__ T warp_inclusive_scan(T v) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T n = __shfl_up_sync(0xFFFFFFFFu, v, d);
    if (lane_id() >= static_cast<uint32_t>(d)) v += n;
  }
  return v;
}

// Exclusive scan of one value per thread over a block of NW warps. `smem` holds NW + 1 entries.
// Returns the exclusive prefix; *total receives the block sum. Contains two __syncthreads().
template <class T, int NW>
__device__ __forceinline__ T block_exclusive_scan(T v, T* smem, T* total) {
  T incl = warp_inclusive_scan(v);
  if (lane_id() == 31) smem[warp_id()] = incl;
  __syncthreads();
  if (threadIdx.x == 0) {
    T run = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      T t = smem[w];
      smem[w] = run;
      run += t;
    }
    smem[NW] = run;
  }
  __syncthreads();
  T base = smem[warp_id()];
  *total = smem[NW];
  return base + incl - v;
}

// Largest i in [0, n) with a[i] <= x (a is non-decreasing, a[0] <= x).
template <class T>
__device__ __forceinline__ uint32_t find_segment(const T* __restrict__ a, uint32_t n, T x) {
  uint32_t lo = 0, hi = n;  // invariant: a[lo] <= x, (hi == n or a[hi] > x)
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (a[mid] <= x) lo = mid;
    else hi = mid;
  }
  return lo;
}

// ---- byte sinks used by the message framing traversal ------------------------------------------------
struct CountSink {
  uint64_t pos = 0;
  __host__ __device__ __forceinline__ void put(uint8_t) { pos += 1; }
  __host__ __device__ __forceinline__ void put_bytes(const uint8_t*, uint32_t n) { pos += n; }
  __host__ __device__ __forceinline__ void skip(uint64_t n) { pos += n; }
  __host__ __device__ __forceinline__ void note(uint64_t*) {}  // only the writing pass records positions
};
struct ByteSink {
  uint8_t* base;
  uint64_t pos;
  __host__ __device__ __forceinline__ void put(uint8_t b) { base[pos++] = b; }
  __host__ __device__ __forceinline__ void put_bytes(const uint8_t* p, uint32_t n) {
    for (uint32_t i = 0; i < n; ++i) base[pos + i] = p[i];
    pos += n;
  }
  __host__ __device__ __forceinline__ void skip(uint64_t n) { pos += n; }
  __host__ __device__ __forceinline__ void note(uint64_t* slot) { *slot = pos; }
};
template <class Sink>
__host__ __device__ __forceinline__ void put_varint(Sink& s, uint64_t v) {
  while (v >= 0x80) {
    s.put(static_cast<uint8_t>((v & 0x7F) | 0x80));
    v >>= 7;
  }
  s.put(static_cast<uint8_t>(v));
}
template <class Sink>
__host__ __device__ __forceinline__ void put_key(Sink& s, uint32_t field_number, uint32_t wire_type) {
  put_varint(s, (static_cast<uint64_t>(field_number) << 3) | wire_type);
}

// ---- 16-byte global accesses that bypass L1 allocation (streaming data is touched once) -----------------
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- whole-CTA byte copy, any alignment on both sides (K5 generic path, K8 bulk path) ------------------------------
__device__ __forceinline__ uint4 shift_window(const uint4& a, const uint4& b, uint32_t delta) {
  // bytes [delta, delta+16) of the 32-byte little-endian window a:b, 0 < delta < 16
  const uint32_t sh = (delta & 3u) * 8u;
  uint4 r;
  switch (delta >> 2) {
    case 0:
      r.x = __funnelshift_r(a.x, a.y, sh); r.y = __funnelshift_r(a.y, a.z, sh);
      r.z = __funnelshift_r(a.z, a.w, sh); r.w = __funnelshift_r(a.w, b.x, sh);
      break;
    case 1:
      r.x = __funnelshift_r(a.y, a.z, sh); r.y = __funnelshift_r(a.z, a.w, sh);
      r.z = __funnelshift_r(a.w, b.x, sh); r.w = __funnelshift_r(b.x, b.y, sh);
      break;
    case 2:
      r.x = __funnelshift_r(a.z, a.w, sh); r.y = __funnelshift_r(a.w, b.x, sh);
      r.z = __funnelshift_r(b.x, b.y, sh); r.w = __funnelshift_r(b.y, b.z, sh);
      break;
    default:
      r.x = __funnelshift_r(a.w, b.x, sh); r.y = __funnelshift_r(b.x, b.y, sh);
      r.z = __funnelshift_r(b.y, b.z, sh); r.w = __funnelshift_r(b.z, b.w, sh);
      break;
  }
  return r;
}

__device__ __forceinline__ uint4 shfl_down_u4(const uint4& v) {
  uint4 r;
  r.x = __shfl_down_sync(0xFFFFFFFFu, v.x, 1);
  r.y = __shfl_down_sync(0xFFFFFFFFu, v.y, 1);
  r.z = __shfl_down_sync(0xFFFFFFFFu, v.z, 1);
  r.w = __shfl_down_sync(0xFFFFFFFFu, v.w, 1);
  return r;
}

constexpr int kVecPerThread = kCopyTileBytes / 16 / kCopyThreads;  // 8 x 16 B per thread and tile

// Copy `len` bytes src -> dst, any alignment on both sides, by the whole CTA (VPT 16-byte vectors in flight per thread).
template <int VPT = kVecPerThread>
__device__ __forceinline__ void cta_copy_bytes(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint32_t len) {
  constexpr int kVecPerThread = VPT;  // 16-byte vectors in flight per thread (shadows the default)
  const uint32_t tid = threadIdx.x;
  uint32_t head = (16u - static_cast<uint32_t>(reinterpret_cast<uintptr_t>(dst) & 15u)) & 15u;
  if (head > len) head = len;
  if (tid < head) dst[tid] = src[tid];
  const uint32_t nvec = (len - head) >> 4;
  uint8_t* d = dst + head;
  const uint8_t* s = src + head;
  const uint32_t delta = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(s) & 15u);
  if (delta == 0) {
    for (uint32_t base = 0; base < nvec; base += kVecPerThread * kCopyThreads) {
      uint4 v[kVecPerThread];
#pragma unroll
      for (int u = 0; u < kVecPerThread; ++u) {
        const uint32_t c = base + u * kCopyThreads + tid;
        if (c < nvec) v[u] = ld_stream16(s + 16ull * c);
      }
#pragma unroll
      for (int u = 0; u < kVecPerThread; ++u) {
        const uint32_t c = base + u * kCopyThreads + tid;
        if (c < nvec) st_stream16(d + 16ull * c, v[u]);
      }
    }
  } else {
    // output word c needs the aligned source words c and c+1: word c+1 is what the next lane loaded
    const uint8_t* sa = s - delta;  // 16-byte aligned; every word read below overlaps valid source bytes
    for (uint32_t base = 0; base < nvec; base += kVecPerThread * kCopyThreads) {
      uint4 a[kVecPerThread], b[kVecPerThread];
#pragma unroll
      for (int u = 0; u < kVecPerThread; ++u) {
        const uint32_t c = base + u * kCopyThreads + tid;
        // word nvec is read too: it is the upper neighbour of the last output word (it still overlaps
        // valid source bytes because delta > 0)
        a[u] = c <= nvec ? ld_stream16(sa + 16ull * c) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < kVecPerThread; ++u) {
        const uint32_t c = base + u * kCopyThreads + tid;
        b[u] = shfl_down_u4(a[u]);
        if (lane_id() == 31 && c < nvec) b[u] = ld_stream16(sa + 16ull * c + 16);
      }
#pragma unroll
      for (int u = 0; u < kVecPerThread; ++u) {
        const uint32_t c = base + u * kCopyThreads + tid;
        if (c < nvec) st_stream16(d + 16ull * c, shift_window(a[u], b[u], delta));
      }
    }
  }
  const uint32_t tail0 = head + (nvec << 4);
  if (tid < len - tail0) dst[tail0 + tid] = src[tail0 + tid];
}


}  // namespace tfs
