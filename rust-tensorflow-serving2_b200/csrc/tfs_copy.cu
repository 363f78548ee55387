// tfs_copy.cu — K5: bulk copy / cast over a table of segments (fixed-width packed fields, tensor_content,
// bool bodies, fused u8->f32 and f32<->f16/bf16 casts).
//
// Replaces, on the reference path (src/lib.rs:265): prost's per-element `put_f32_le` / `get_f32_le`
// loops of float/double::{encode_packed, merge_repeated}, bytes::{encode, merge}, and the client's own
// `*p as f32` + preprocessing_fn loop (src/lib.rs:237-242).
//
// HBM-bound byte work: no tensor cores.  Two kernels share one segment table and one tiling:
//   k_copy_tma  segments whose source and destination are congruent mod 16 (what TFS_F_ALIGN_PAYLOAD and the
//               64-byte aligned value pool give): a 3-stage x 32 KiB shared-memory ring driven by ONE thread per
//               CTA, one CTA per SM — cp.async.bulk global->shared (mbarrier complete_tx), then cp.async.bulk
//               shared->global.  The SM issues two instructions per 32 KiB; the TMA engine moves the bytes.
//   k_copy      everything else, one CTA per tile: 16-byte aligned streaming loads, source re-aligned in
//               registers (neighbour words come from a warp shuffle, funnel shifts do the byte shift), 16-byte
//               aligned streaming stores; element-wise fused casts.  Launched only when the host cannot rule
//               out such segments, and its CTAs leave at once when the batch turned out to have none.
// Byte segments are tiled on the 16-byte grid of the ABSOLUTE destination address (tile t of a segment covers
// [A0 + t*32 KiB, A0 + (t+1)*32 KiB) with A0 = dst & ~15), so every tile interior is vector / TMA aligned.
// Tiles are dealt round-robin to the CTAs: at any moment all SMs work inside one compact window of addresses,
// which measured 25 % faster on B200 than giving each CTA its own contiguous range (scripts/copy_sweep.cu).
#include <algorithm>

#include "tfs_device.cuh"

namespace tfs {

namespace {

constexpr uint32_t kTmaStages = 3;
constexpr uint32_t kTmaChunk = kCopyTileBytes;  // one bulk copy per tile
constexpr uint64_t kTmaMinBytes = 4096;         // smaller segments go through k_copy

__device__ __forceinline__ bool is_byte_op(int32_t op) { return op == CP_BYTES || op == CP_BOOL; }

__device__ __forceinline__ uint32_t out_size_of(int32_t op) {
  switch (op) {
    case CP_U8_F32: case CP_U8_F32_DIV: case CP_U8_F32_MUL: case CP_F16_F32: case CP_BF16_F32: return 4;
    case CP_F32_F16: case CP_F32_BF16: return 2;
    default: return 1;
  }
}
__device__ __forceinline__ uint32_t in_size_of(int32_t op) {
  switch (op) {
    case CP_F32_F16: case CP_F32_BF16: return 4;
    case CP_F16_F32: case CP_BF16_F32: return 2;
    default: return 1;
  }
}

// Output-element range [e0, e1) of tile `t` of segment `s` (bytes for byte ops).
__device__ __forceinline__ void tile_range(const CopySeg& s, uint64_t t, uint64_t* e0, uint64_t* e1) {
  if (is_byte_op(s.op)) {
    const uint64_t d = reinterpret_cast<uintptr_t>(s.dst);
    const uint64_t a0 = d & ~static_cast<uint64_t>(15);
    const uint64_t lo_abs = a0 + t * kCopyTileBytes, hi_abs = lo_abs + kCopyTileBytes;
    const uint64_t lo = lo_abs > d ? lo_abs - d : 0;
    const uint64_t hi = hi_abs > d ? hi_abs - d : 0;
    *e0 = lo < s.n ? lo : s.n;
    *e1 = hi < s.n ? hi : s.n;
  } else {
    const uint64_t per = kCopyTileBytes / out_size_of(s.op);
    const uint64_t lo = t * per, hi = lo + per;
    *e0 = lo < s.n ? lo : s.n;
    *e1 = hi < s.n ? hi : s.n;
  }
}

// Segment owning `tile`, searching forward from `si` (tiles visited by one CTA only grow).
__device__ __forceinline__ uint32_t seg_from(const CopySeg* __restrict__ segs, uint32_t n_segs, uint32_t si, uint64_t tile) {
  if (si + 1 >= n_segs || segs[si + 1].first_tile > tile) return si;
  uint32_t lo = si + 1, hi = n_segs;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (segs[mid].first_tile <= tile) lo = mid;
    else hi = mid;
  }
  return lo;
}

// Segment owning `tile`, found by one warp: every round the 32 lanes probe 32 evenly spaced table entries at once, so a
// table of up to 32^3 segments costs three dependent memory round trips instead of fifteen (a CTA that copies a few
// hundred bytes spends most of its life in this lookup).  All 32 lanes must call it; all return the same index.
__device__ __forceinline__ uint32_t seg_of_tile_warp(const CopySeg* __restrict__ segs, uint32_t n_segs, uint64_t tile, uint32_t from = 0) {
  uint32_t lo = from, hi = n_segs;  // invariant: segs[lo].first_tile <= tile, (hi == n_segs or segs[hi].first_tile > tile)
  while (hi - lo > 1) {
    const uint32_t span = hi - lo;
    const uint32_t step = (span + 31u) / 32u;
    const uint32_t idx = lo + lane_id() * step;  // lane 0 probes `lo` itself (true by the invariant)
    const bool le = idx < hi && segs[idx].first_tile <= tile;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, le);
    const uint32_t k = 31u - static_cast<uint32_t>(__clz(static_cast<int>(m)));  // last lane whose probe is <= tile (m has bit 0)
    const uint32_t nlo = lo + k * step;
    const uint32_t nhi = nlo + step < hi ? nlo + step : hi;
    lo = nlo;
    hi = nhi;
  }
  return lo;
}

// `x / p` and `x * p` as the reference's host computes them (src/lib.rs:237-242 on x86-64): the IEEE operation, and when the
// result is a NaN the bits SSE gives it - the NaN operand made quiet if `p` is one, the "real indefinite" 0xFFC00000 for
// 0 / 0 and 0 * inf - not the GPU's canonical 0x7FFFFFFF.  (x comes from a u8: never a NaN, never negative.)
__device__ __forceinline__ float nan_like_sse(float r, float p) {
  if (r != r) {
    const uint32_t pb = __float_as_uint(p);
    return __uint_as_float((pb & 0x7FFFFFFFu) > 0x7F800000u ? (pb | 0x00400000u) : 0xFFC00000u);
  }
  return r;
}
__device__ __forceinline__ float u8_div(float x, float p) { return nan_like_sse(__fdiv_rn(x, p), p); }
__device__ __forceinline__ float u8_mul(float x, float p) { return nan_like_sse(__fmul_rn(x, p), p); }

// ---------------------------------------------------------------------------------------------------
// generic path: one CTA per tile
// ---------------------------------------------------------------------------------------------------
template <class T>
__device__ __forceinline__ T load_unaligned(const uint8_t* p, bool aligned) {
  T v;
  if (aligned) {
    v = *reinterpret_cast<const T*>(p);
  } else {
    uint8_t* b = reinterpret_cast<uint8_t*>(&v);
#pragma unroll
    for (int i = 0; i < static_cast<int>(sizeof(T)); ++i) b[i] = p[i];
  }
  return v;
}
template <class T>
__device__ __forceinline__ void store_unaligned(uint8_t* p, T v, bool aligned) {
  if (aligned) {
    *reinterpret_cast<T*>(p) = v;
  } else {
    const uint8_t* b = reinterpret_cast<const uint8_t*>(&v);
#pragma unroll
    for (int i = 0; i < static_cast<int>(sizeof(T)); ++i) p[i] = b[i];
  }
}

// Element-wise cast of `n` output elements.  IN / OUT are the in-memory widths; F converts.
template <class IN, class OUT, class F>
__device__ __forceinline__ void cta_cast(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint32_t n, F f) {
  const bool in_al = (reinterpret_cast<uintptr_t>(src) & (sizeof(IN) - 1)) == 0;
  const bool out_al = (reinterpret_cast<uintptr_t>(dst) & (sizeof(OUT) - 1)) == 0;
  for (uint32_t i = threadIdx.x; i < n; i += kCopyThreads) {
    IN x = load_unaligned<IN>(src + static_cast<uint64_t>(i) * sizeof(IN), in_al);
    store_unaligned<OUT>(dst + static_cast<uint64_t>(i) * sizeof(OUT), f(x), out_al);
  }
}

__device__ __forceinline__ bool tma_eligible(const CopySeg& s) {
  return s.op == CP_BYTES && s.n >= kTmaMinBytes &&
         ((reinterpret_cast<uintptr_t>(s.src) ^ reinterpret_cast<uintptr_t>(s.dst)) & 15u) == 0;
}

// Widening / narrowing segments large enough for the staged cast kernel below (k_copy_cast); the source has to be aligned
// to its element size (anything else stays with the element-wise loop of k_copy).
constexpr uint64_t kCastMinElems = 1024;
__device__ __forceinline__ bool is_cast_op(int32_t op) { return op >= CP_U8_F32 && op <= CP_BF16_F32; }
__device__ __forceinline__ bool cast_eligible(const CopySeg& s) {
  return is_cast_op(s.op) && s.n >= kCastMinElems && (reinterpret_cast<uintptr_t>(s.src) & (in_size_of(s.op) - 1u)) == 0;
}

__global__ void __launch_bounds__(kCopyThreads) k_copy(const CopySeg* __restrict__ segs, uint32_t n_segs, uint64_t n_tiles,
                                                        const uint64_t* __restrict__ gate, uint64_t src_lo, uint64_t src_hi,
                                                        uint64_t tile_begin) {
  __shared__ uint32_t s_seg;
  if (gate && *gate != 0) return;  // an earlier kernel of the batch refused to go on (dst too small)
  const uint64_t tile = tile_begin + blockIdx.x;
  if (tile >= n_tiles) return;
  if (threadIdx.x < 32) {
    const uint32_t si = seg_of_tile_warp(segs, n_segs, tile);
    if (threadIdx.x == 0) s_seg = si;
  }
  __syncthreads();
  const CopySeg seg = segs[s_seg];
  if (seg.n == 0 || tma_eligible(seg) || cast_eligible(seg)) return;  // the TMA ring / the staged cast kernel owns this segment
  uint64_t e0, e1;
  tile_range(seg, tile - seg.first_tile, &e0, &e1);
  if (e0 >= e1) return;
  {  // a pipelined launch owns the tiles whose source bytes end inside its window
    const uint64_t src_end = reinterpret_cast<uintptr_t>(seg.src) + e1 * in_size_of(seg.op);
    if (src_end <= src_lo || src_end > src_hi) return;
  }
  const uint32_t n = static_cast<uint32_t>(e1 - e0);
  uint8_t* d = seg.dst + e0 * out_size_of(seg.op);
  const uint8_t* s = seg.src + e0 * in_size_of(seg.op);
  const float p = seg.param;
  switch (seg.op) {
    case CP_BYTES: cta_copy_bytes(d, s, n); break;
    case CP_BOOL: cta_cast<uint8_t, uint8_t>(d, s, n, [](uint8_t x) { return static_cast<uint8_t>(x != 0); }); break;
    case CP_U8_F32: cta_cast<uint8_t, float>(d, s, n, [](uint8_t x) { return static_cast<float>(x); }); break;
    case CP_U8_F32_DIV: cta_cast<uint8_t, float>(d, s, n, [p](uint8_t x) { return u8_div(static_cast<float>(x), p); }); break;
    case CP_U8_F32_MUL: cta_cast<uint8_t, float>(d, s, n, [p](uint8_t x) { return u8_mul(static_cast<float>(x), p); }); break;
    case CP_F32_F16: cta_cast<float, uint16_t>(d, s, n, [](float x) { return static_cast<uint16_t>(f32_to_f16_bits(x)); }); break;
    case CP_F32_BF16: cta_cast<float, uint16_t>(d, s, n, [](float x) { return static_cast<uint16_t>(f32_to_bf16_bits(x)); }); break;
    case CP_F16_F32: cta_cast<uint16_t, float>(d, s, n, [](uint16_t x) { return f16_bits_to_f32(x); }); break;
    case CP_BF16_F32: cta_cast<uint16_t, float>(d, s, n, [](uint16_t x) { return bf16_bits_to_f32(x); }); break;
    default: break;
  }
}

// ---------------------------------------------------------------------------------------------------
// small segments (<= kSmallSegBytes of output): they own no tiles; one WARP per segment of the table
// ---------------------------------------------------------------------------------------------------
template <class IN, class OUT, class F>
__device__ __forceinline__ void warp_cast(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint32_t n, F f) {
  const bool in_al = (reinterpret_cast<uintptr_t>(src) & (sizeof(IN) - 1)) == 0;
  const bool out_al = (reinterpret_cast<uintptr_t>(dst) & (sizeof(OUT) - 1)) == 0;
  for (uint32_t i = lane_id(); i < n; i += 32u) {
    IN x = load_unaligned<IN>(src + static_cast<uint64_t>(i) * sizeof(IN), in_al);
    store_unaligned<OUT>(dst + static_cast<uint64_t>(i) * sizeof(OUT), f(x), out_al);
  }
}

__device__ __forceinline__ bool small_seg(const CopySeg& s) { return s.n * out_size_of(s.op) <= kSmallSegBytes; }

__global__ void __launch_bounds__(kCopyThreads) k_copy_small(const CopySeg* __restrict__ segs, uint32_t n_segs, const uint64_t* __restrict__ gate) {
  if (gate && *gate != 0) return;
  const uint32_t si = (blockIdx.x * kCopyThreads + threadIdx.x) >> 5;
  if (si >= n_segs) return;
  const CopySeg seg = segs[si];
  if (seg.n == 0 || !small_seg(seg)) return;
  const uint32_t n = static_cast<uint32_t>(seg.n);
  uint8_t* d = seg.dst;
  const uint8_t* s = seg.src;
  const float p = seg.param;
  switch (seg.op) {
    case CP_BYTES:
      if (((reinterpret_cast<uintptr_t>(d) | reinterpret_cast<uintptr_t>(s)) & 3u) == 0) {  // words, then the tail
        for (uint32_t i = lane_id(); i < (n >> 2); i += 32u) reinterpret_cast<uint32_t*>(d)[i] = reinterpret_cast<const uint32_t*>(s)[i];
        for (uint32_t i = (n & ~3u) + lane_id(); i < n; i += 32u) d[i] = s[i];
      } else {
        for (uint32_t i = lane_id(); i < n; i += 32u) d[i] = s[i];
      }
      break;
    case CP_BOOL: warp_cast<uint8_t, uint8_t>(d, s, n, [](uint8_t x) { return static_cast<uint8_t>(x != 0); }); break;
    case CP_U8_F32: warp_cast<uint8_t, float>(d, s, n, [](uint8_t x) { return static_cast<float>(x); }); break;
    case CP_U8_F32_DIV: warp_cast<uint8_t, float>(d, s, n, [p](uint8_t x) { return u8_div(static_cast<float>(x), p); }); break;
    case CP_U8_F32_MUL: warp_cast<uint8_t, float>(d, s, n, [p](uint8_t x) { return u8_mul(static_cast<float>(x), p); }); break;
    case CP_F32_F16: warp_cast<float, uint16_t>(d, s, n, [](float x) { return static_cast<uint16_t>(f32_to_f16_bits(x)); }); break;
    case CP_F32_BF16: warp_cast<float, uint16_t>(d, s, n, [](float x) { return static_cast<uint16_t>(f32_to_bf16_bits(x)); }); break;
    case CP_F16_F32: warp_cast<uint16_t, float>(d, s, n, [](uint16_t x) { return f16_bits_to_f32(x); }); break;
    case CP_BF16_F32: warp_cast<uint16_t, float>(d, s, n, [](uint16_t x) { return bf16_bits_to_f32(x); }); break;
    default: break;
  }
}

// ---------------------------------------------------------------------------------------------------
// TMA path: cp.async.bulk global -> shared -> global, one driving thread per CTA, one CTA per SM
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// The payload streams through L2 exactly once in each direction: mark its lines evict-first so that the small
// tables and the code of the framing / walk kernels that run between two copies stay resident in the 126 MB L2.
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void tma_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}
__device__ __forceinline__ void tma_store(void* gdst, const void* smem_src, uint32_t bytes, uint64_t pol) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void tma_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// Walks the tiles blockIdx.x, blockIdx.x + G, ... and yields the 16-byte aligned interior of every tile that
// belongs to a TMA-eligible segment; the < 16-byte ragged ends of a segment are copied byte by byte on the way.
struct TileWalk {
  const CopySeg* segs;
  uint32_t n_segs, si;
  uint64_t tile, n_tiles, stride;
  uint64_t src_lo, src_hi;  // this launch owns the tiles whose source bytes END in (src_lo, src_hi]

  // COOP = false: one thread walks alone (few segments: the hint almost always hits).  COOP = true: called by all 32
  // lanes of the CTA's warp in lockstep (every lane computes the same result): the segment lookup is a warp-wide
  // search, the < 16 ragged bytes at a segment's ends are split over the lanes.
  template <bool COOP>
  __device__ bool next(const uint8_t** src, uint8_t** dst, uint32_t* bytes) {
    if (COOP) __syncwarp();  // the leader may come from a TMA-only block: reconverge before the warp-wide search
    const uint32_t l0 = COOP ? lane_id() : 0u, lstep = COOP ? 32u : 1u;
    for (; tile < n_tiles; tile += stride) {
      if (COOP) {
        if (si + 1 < n_segs && segs[si + 1].first_tile <= tile) si = seg_of_tile_warp(segs, n_segs, tile, si + 1);
      } else {
        si = seg_from(segs, n_segs, si, tile);
      }
      const CopySeg seg = segs[si];
      if (seg.n == 0 || !tma_eligible(seg)) continue;
      uint64_t a, b;
      tile_range(seg, tile - seg.first_tile, &a, &b);
      if (a >= b) continue;
      const uint64_t src_end = reinterpret_cast<uintptr_t>(seg.src) + b;
      if (src_end <= src_lo || src_end > src_hi) continue;
      const uint64_t d = reinterpret_cast<uintptr_t>(seg.dst);
      const uint32_t head = static_cast<uint32_t>((16u - ((d + a) & 15u)) & 15u);  // only in a segment's first tile
      if (head) {
        const uint64_t h1 = a + head < b ? a + head : b;
        for (uint64_t i = a + l0; i < h1; i += lstep) seg.dst[i] = seg.src[i];
        a = h1;
      }
      const uint32_t tail = static_cast<uint32_t>((d + b) & 15u);  // only in its last tile
      if (tail && b > a) {
        const uint64_t t0 = b - tail > a ? b - tail : a;
        for (uint64_t i = t0 + l0; i < b; i += lstep) seg.dst[i] = seg.src[i];
        b = t0;
      }
      if (a >= b) continue;
      *src = seg.src + a;
      *dst = seg.dst + a;
      *bytes = static_cast<uint32_t>(b - a);
      tile += stride;
      return true;
    }
    return false;
  }
};

// counts that are still being computed on the device when the kernel is queued (launch_copy_tma_deferred)
struct DeferredCounts {
  const uint64_t* go = nullptr;  // nullptr: the by-value arguments hold
  const uint64_t* n_segs = nullptr;
  const uint64_t* n_tiles = nullptr;
};

template <bool COOP>
__global__ void __launch_bounds__(32) k_copy_tma(const CopySeg* __restrict__ segs, uint32_t n_segs, uint64_t n_tiles,
                                                 const uint64_t* __restrict__ gate, uint64_t src_lo, uint64_t src_hi,
                                                 uint64_t tile_begin, DeferredCounts deferred = DeferredCounts{}) {
  extern __shared__ __align__(128) uint8_t ring[];  // kTmaStages x kTmaChunk
  __shared__ __align__(8) uint64_t full_bar[kTmaStages];
  // Lane 0 drives the TMA engine.  COOP = false (few segments, the C2 shape): the other lanes leave at once.  COOP = true
  // (hundreds of segments): they stay to help find the tiles (warp-wide segment search) and copy the ragged ends; all
  // lanes then run the same control flow on the same values.
  if (!COOP && threadIdx.x != 0) return;
  const bool leader = threadIdx.x == 0;
  if (gate && *gate != 0) return;
  if (deferred.go) {  // queued before the counts existed: they are in device memory now (written by the kernel in front)
    if (*deferred.go == 0) return;
    n_segs = static_cast<uint32_t>(*deferred.n_segs);
    n_tiles = *deferred.n_tiles;
  }
  if (tile_begin + blockIdx.x >= n_tiles) return;
  if (leader) {
#pragma unroll
    for (uint32_t s = 0; s < kTmaStages; ++s) mbar_init(&full_bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (COOP) __syncwarp();

  const uint64_t pol = l2_evict_first_policy();
  TileWalk walk{segs, n_segs, 0u, tile_begin + blockIdx.x, n_tiles, gridDim.x, src_lo, src_hi};
  uint8_t* dsts[kTmaStages];
  uint32_t sizes[kTmaStages];
  uint32_t issued = 0, stored = 0;
  auto issue = [&]() -> bool {
    const uint8_t* src;
    uint8_t* dst;
    uint32_t bytes;
    if (!walk.template next<COOP>(&src, &dst, &bytes)) return false;
    const uint32_t s = issued % kTmaStages;
    dsts[s] = dst;
    sizes[s] = bytes;
    if (leader) {
      mbar_expect_tx(&full_bar[s], bytes);
      tma_load(ring + s * kTmaChunk, src, bytes, &full_bar[s], pol);
    }
    ++issued;
    return true;
  };
  // prologue: kTmaStages - 1 loads in flight
  for (uint32_t k = 0; k + 1 < kTmaStages; ++k)
    if (!issue()) break;
  while (stored < issued) {
    const uint32_t s = stored % kTmaStages;
    if (leader) {
      mbar_wait(&full_bar[s], (stored / kTmaStages) & 1u);
      tma_store(dsts[s], ring + s * kTmaChunk, sizes[s], pol);
      tma_commit();
      // all but the newest store have finished reading shared memory: the stage of store (stored - 1) is free,
      // and it is exactly the stage the next load (index `issued`) maps to
      tma_wait_read<1>();
    }
    ++stored;
    if (COOP) __syncwarp();
    issue();
  }
  if (leader) tma_wait_all();
}

// ---------------------------------------------------------------------------------------------------
// staged casts: u8 -> f32 (x 1, / p, * p), f32 -> f16 / bf16, f16 / bf16 -> f32.  One CTA per 32 KiB tile of OUTPUT.
//   in   one cp.async.bulk (TMA) brings the 16-byte aligned window around the tile's source elements into shared memory
//        (mbarrier complete_tx);
//   cast all 8 warps convert from shared memory, 4 elements per step on aligned vectors, into an output stage whose
//        first element is padded so that the groups stay vector-aligned;
//   out  when the stage and the destination are congruent mod 16 (TFS_F_ALIGN_PAYLOAD, the 64-byte aligned value pool)
//        the interior leaves with one cp.async.bulk shared -> global; otherwise every thread shifts 16 staged bytes
//        onto the destination's 16-byte grid (two vector loads, funnel shifts) and stores them.
// u8 -> f32 with a scale goes through a 256-entry table built with the same IEEE operation (__fdiv_rn / __fmul_rn) the
// element-wise path uses: bit-identical results, no division per pixel.
// ---------------------------------------------------------------------------------------------------
constexpr uint32_t kCastPad = 64;  // slack around both stages (window alignment, group padding, flush over-read)

// u8 -> f32 with a scale, ARITH: computed, not looked up.  `x * p` is the operation itself; `x / p` is the product with the
// reciprocal corrected by the remainder (two fused multiply-adds), used only when the CTA has checked that it gives the bits
// of the IEEE division for all 256 inputs (k_copy_cast).  The table costs shared-memory bank conflicts: pixel values are
// random, so the 32 lanes of a look-up hit the same bank several at a time.
struct CastScale {
  float p, rcp;
};
template <int OP>
__device__ __forceinline__ float scale_u8(uint32_t b, const CastScale& s) {
  const float x = static_cast<float>(b);
  if constexpr (OP == CP_U8_F32_MUL) {
    return __fmul_rn(x, s.p);
  } else {
    const float q0 = __fmul_rn(x, s.rcp);
    return __fmaf_rn(__fmaf_rn(-q0, s.p, x), s.rcp, q0);
  }
}

template <int OP, bool ARITH = false>
__device__ __forceinline__ void cast_group(const uint8_t* in, uint8_t* out, const float* lut, const CastScale& sc) {
  // 4 elements: `in` / `out` are aligned to 4 input / output elements
  if constexpr (OP == CP_U8_F32 || OP == CP_U8_F32_DIV || OP == CP_U8_F32_MUL) {
    const uint32_t w = *reinterpret_cast<const uint32_t*>(in);
    float4 r;
    if constexpr (OP == CP_U8_F32) {
      r = make_float4(static_cast<float>(w & 0xFFu), static_cast<float>((w >> 8) & 0xFFu), static_cast<float>((w >> 16) & 0xFFu), static_cast<float>(w >> 24));
    } else if constexpr (ARITH) {
      r = make_float4(scale_u8<OP>(w & 0xFFu, sc), scale_u8<OP>((w >> 8) & 0xFFu, sc), scale_u8<OP>((w >> 16) & 0xFFu, sc), scale_u8<OP>(w >> 24, sc));
    } else {
      r = make_float4(lut[w & 0xFFu], lut[(w >> 8) & 0xFFu], lut[(w >> 16) & 0xFFu], lut[w >> 24]);
    }
    *reinterpret_cast<float4*>(out) = r;
  } else if constexpr (OP == CP_F32_F16 || OP == CP_F32_BF16) {
    const float4 v = *reinterpret_cast<const float4*>(in);
    uint32_t a, b, c, d;
    if constexpr (OP == CP_F32_F16) {
      a = f32_to_f16_bits(v.x); b = f32_to_f16_bits(v.y); c = f32_to_f16_bits(v.z); d = f32_to_f16_bits(v.w);
    } else {
      a = f32_to_bf16_bits(v.x); b = f32_to_bf16_bits(v.y); c = f32_to_bf16_bits(v.z); d = f32_to_bf16_bits(v.w);
    }
    *reinterpret_cast<uint2*>(out) = make_uint2(a | (b << 16), c | (d << 16));
  } else {
    const uint2 w = *reinterpret_cast<const uint2*>(in);
    float4 r;
    if constexpr (OP == CP_F16_F32) {
      r = make_float4(f16_bits_to_f32(w.x & 0xFFFFu), f16_bits_to_f32(w.x >> 16), f16_bits_to_f32(w.y & 0xFFFFu), f16_bits_to_f32(w.y >> 16));
    } else {
      r = make_float4(bf16_bits_to_f32(w.x & 0xFFFFu), bf16_bits_to_f32(w.x >> 16), bf16_bits_to_f32(w.y & 0xFFFFu), bf16_bits_to_f32(w.y >> 16));
    }
    *reinterpret_cast<float4*>(out) = r;
  }
}

template <int OP, bool ARITH = false>
__device__ __forceinline__ void cast_tile(const CopySeg& seg, uint64_t e0, uint32_t n, uint8_t* sm_in, uint8_t* sm_out, uint64_t* bar, const float* lut,
                                          const CastScale& sc = CastScale{0.f, 0.f}) {
  constexpr uint32_t IS = (OP == CP_F32_F16 || OP == CP_F32_BF16) ? 4u : (OP == CP_F16_F32 || OP == CP_BF16_F32) ? 2u : 1u;
  constexpr uint32_t OS = (OP == CP_F32_F16 || OP == CP_F32_BF16) ? 2u : 4u;
  const uint64_t s_abs = reinterpret_cast<uintptr_t>(seg.src) + e0 * IS;
  const uint64_t s_lo = s_abs & ~static_cast<uint64_t>(15);
  const uint32_t in_bytes = static_cast<uint32_t>(((s_abs + static_cast<uint64_t>(n) * IS + 15u) & ~static_cast<uint64_t>(15)) - s_lo);
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, in_bytes);
    tma_load(sm_in, reinterpret_cast<const void*>(s_lo), in_bytes, bar, l2_evict_first_policy());
  }
  // elements in front of the tile's first one inside the window; `lead` of them share its 4-element group
  const uint32_t skip = static_cast<uint32_t>(s_abs - s_lo) / IS;
  const uint32_t lead = skip & 3u;
  const uint32_t g0 = skip >> 2;                      // first group (of 4 input elements, window-aligned) that holds tile elements
  const uint32_t n_groups = (lead + n + 3u) >> 2;
  mbar_wait(bar, 0);
  // output stage: element j of the tile at sm_out + (lead + j) * OS, so that group g's four outputs start at 4 * OS * g
  for (uint32_t g = threadIdx.x; g < n_groups; g += kCopyThreads)
    cast_group<OP, ARITH>(sm_in + (g0 + g) * 4u * IS, sm_out + g * 4u * OS, lut, sc);
  __syncthreads();
  // flush the bytes [lead * OS, (lead + n) * OS) of the stage to dst
  uint8_t* d = seg.dst + e0 * OS;
  const uint32_t total = n * OS;
  const uint8_t* st = sm_out + lead * OS;  // byte k of the tile's output
  const uint32_t skew = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(d) & 15u);
  const uint32_t head = (16u - skew) & 15u;  // bytes up to the first 16-byte boundary of dst
  if (((lead * OS + head) & 15u) == 0 && total >= head + 16u) {
    // stage and destination congruent: one bulk store for the interior, the ragged ends byte by byte
    const uint32_t body = (total - head) & ~15u;
    if (threadIdx.x == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the stage was written by ordinary stores
      tma_store(d + head, st + head, body, l2_evict_first_policy());
      tma_commit();
    }
    for (uint32_t k = threadIdx.x; k < head; k += kCopyThreads) d[k] = st[k];
    for (uint32_t k = head + body + threadIdx.x; k < total; k += kCopyThreads) d[k] = st[k];
    if (threadIdx.x == 0) tma_wait_all();
  } else {
    uint8_t* da = d - skew;
    const uint32_t end = skew + total;
    const uint32_t nchunks = (end + 15u) >> 4;
    const uint32_t base = lead * OS;  // stage offset of output byte 0; st - sm_out
    for (uint32_t c = threadIdx.x; c < nchunks; c += kCopyThreads) {
      const uint32_t lo = c << 4, hi = lo + 16u;
      if (lo >= skew && hi <= end) {
        const uint32_t s0 = base + lo - skew, sh = s0 & 15u;
        const uint4 a = *reinterpret_cast<const uint4*>(sm_out + (s0 & ~15u));
        const uint4 b = *reinterpret_cast<const uint4*>(sm_out + (s0 & ~15u) + 16u);
        st_stream16(da + lo, sh ? shift_window(a, b, sh) : a);
      } else {
        const uint32_t b0 = lo > skew ? lo : skew, b1 = hi < end ? hi : end;
        for (uint32_t b = b0; b < b1; ++b) da[b] = st[b - skew];
      }
    }
  }
}

__global__ void __launch_bounds__(kCopyThreads) k_copy_cast(const CopySeg* __restrict__ segs, uint32_t n_segs, uint64_t n_tiles,
                                                             const uint64_t* __restrict__ gate, uint64_t src_lo, uint64_t src_hi,
                                                             uint64_t tile_begin, uint32_t in_stage_bytes) {
  extern __shared__ __align__(128) uint8_t cast_smem[];  // input stage | output stage
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t s_seg;
  __shared__ float lut[256];
  if (gate && *gate != 0) return;
  const uint64_t tile = tile_begin + blockIdx.x;
  if (tile >= n_tiles) return;
  if (threadIdx.x < 32) {
    const uint32_t si = seg_of_tile_warp(segs, n_segs, tile);
    if (threadIdx.x == 0) s_seg = si;
  }
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  const CopySeg seg = segs[s_seg];
  if (seg.n == 0 || !cast_eligible(seg)) return;
  uint64_t e0, e1;
  tile_range(seg, tile - seg.first_tile, &e0, &e1);
  if (e0 >= e1) return;
  {  // a pipelined launch owns the tiles whose source bytes end inside its window
    const uint64_t src_end = reinterpret_cast<uintptr_t>(seg.src) + e1 * in_size_of(seg.op);
    if (src_end <= src_lo || src_end > src_hi) return;
  }
  const CastScale sc{seg.param, __frcp_rn(seg.param)};
  bool arith = false;
  if (seg.op == CP_U8_F32_DIV) {  // the table, and whether the two-FMA form reproduces it bit for bit (it does for 255.f)
    const float x = static_cast<float>(threadIdx.x);
    const float exact = u8_div(x, seg.param);
    lut[threadIdx.x] = exact;
    arith = __syncthreads_and(__float_as_uint(exact) == __float_as_uint(scale_u8<CP_U8_F32_DIV>(threadIdx.x, sc))) != 0;
  } else if (seg.op == CP_U8_F32_MUL) {  // a finite factor cannot make a NaN: the multiplication itself; otherwise the table
    arith = (__float_as_uint(seg.param) & 0x7F800000u) != 0x7F800000u;
    if (!arith) {
      lut[threadIdx.x] = u8_mul(static_cast<float>(threadIdx.x), seg.param);
      __syncthreads();
    }
  }
  uint8_t* sm_in = cast_smem;
  uint8_t* sm_out = cast_smem + in_stage_bytes;
  const uint32_t n = static_cast<uint32_t>(e1 - e0);
  switch (seg.op) {
    case CP_U8_F32: cast_tile<CP_U8_F32>(seg, e0, n, sm_in, sm_out, &bar, lut); break;
    case CP_U8_F32_DIV:
      if (arith) cast_tile<CP_U8_F32_DIV, true>(seg, e0, n, sm_in, sm_out, &bar, lut, sc);
      else cast_tile<CP_U8_F32_DIV>(seg, e0, n, sm_in, sm_out, &bar, lut);
      break;
    case CP_U8_F32_MUL:
      if (arith) cast_tile<CP_U8_F32_MUL, true>(seg, e0, n, sm_in, sm_out, &bar, lut, sc);
      else cast_tile<CP_U8_F32_MUL>(seg, e0, n, sm_in, sm_out, &bar, lut);
      break;
    case CP_F32_F16: cast_tile<CP_F32_F16>(seg, e0, n, sm_in, sm_out, &bar, lut); break;
    case CP_F32_BF16: cast_tile<CP_F32_BF16>(seg, e0, n, sm_in, sm_out, &bar, lut); break;
    case CP_F16_F32: cast_tile<CP_F16_F32>(seg, e0, n, sm_in, sm_out, &bar, lut); break;
    case CP_BF16_F32: cast_tile<CP_BF16_F32>(seg, e0, n, sm_in, sm_out, &bar, lut); break;
    default: break;
  }
}

}  // namespace

tfs_status_t launch_copy(tfs_ctx* ctx, const CopySeg* segs_dev, uint32_t n_segs, uint64_t n_tiles, const uint64_t* gate,
                         bool need_generic, const std::vector<HostSeg>* host_segs, bool need_tma, int cast_kind, bool any_small) {
  if (n_segs == 0) return TFS_OK;
  auto launch_small = [&]() -> tfs_status_t {
    if (!any_small) return TFS_OK;
    ProfScope prof(ctx, PK_COPY);
    k_copy_small<<<static_cast<uint32_t>((static_cast<uint64_t>(n_segs) * 32u + kCopyThreads - 1) / kCopyThreads), kCopyThreads, 0, ctx->stream>>>(segs_dev, n_segs, gate);
    ctx->launches++;
    return TFS_OK;
  };
  if (n_tiles == 0) {  // nothing but small segments
    if (!ctx->pipe.empty()) {
      TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipe.back().ready, 0));
      ctx->pipe.clear();
    }
    TFS_TRY(launch_small());
    TFS_CUDA(ctx, cudaGetLastError());
    return TFS_OK;
  }
  if (n_tiles > 0x7FFFFFFFull) return set_err(ctx, TFS_ERR_INVALID_ARG, "batch too large (copy tiles)");
  const uint32_t ring_bytes = kTmaStages * kTmaChunk;
  if (!ctx->copy_attr_set) {  // a per-device attribute: once per context, not once per process
    TFS_CUDA(ctx, cudaFuncSetAttribute(k_copy_tma<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(ring_bytes)));
    TFS_CUDA(ctx, cudaFuncSetAttribute(k_copy_tma<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(ring_bytes)));
    TFS_CUDA(ctx, cudaFuncSetAttribute(k_copy_cast, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(3 * kCopyTileBytes + 2 * kCastPad)));
    ctx->copy_attr_set = true;
  }
  // staged casts: the input stage holds the source bytes of one 32 KiB output tile: 8 / 16 KiB for the widening ops,
  // 64 KiB for f32 -> 16-bit
  const uint32_t cast_in = (cast_kind == 2 ? 2u * kCopyTileBytes : kCopyTileBytes / 2u) + kCastPad;
  const uint32_t cast_smem_bytes = cast_in + kCopyTileBytes + kCastPad;
  // tiles [tb, te), of which the kernels take those whose source bytes end in (src_lo, src_hi]
  auto launch = [&](uint64_t src_lo, uint64_t src_hi, uint64_t tb, uint64_t te) {
    if (te <= tb) return;
    if (need_generic) {  // segments that are not congruent mod 16, are tiny, or need arithmetic: one CTA per tile
      ProfScope prof(ctx, PK_COPY);
      k_copy<<<static_cast<uint32_t>(te - tb), kCopyThreads, 0, ctx->stream>>>(segs_dev, n_segs, te, gate, src_lo, src_hi, tb);
      ctx->launches++;
    }
    if (cast_kind) {  // widening / narrowing segments: one CTA per tile, staged through shared memory
      ProfScope prof(ctx, PK_COPY_CAST);
      k_copy_cast<<<static_cast<uint32_t>(te - tb), kCopyThreads, cast_smem_bytes, ctx->stream>>>(segs_dev, n_segs, te, gate, src_lo, src_hi, tb, cast_in);
      ctx->launches++;
    }
    if (need_tma) {  // the TMA ring: one CTA (96 KiB of shared memory, one driving thread) per SM
      const uint64_t cap = static_cast<uint64_t>(ctx->sm_count);
      ProfScope prof(ctx, PK_COPY_TMA);
      const uint32_t grid = static_cast<uint32_t>(te - tb < cap ? te - tb : cap);
      if (n_segs > 64) k_copy_tma<true><<<grid, 32, ring_bytes, ctx->stream>>>(segs_dev, n_segs, te, gate, src_lo, src_hi, tb);
      else k_copy_tma<false><<<grid, 32, ring_bytes, ctx->stream>>>(segs_dev, n_segs, te, gate, src_lo, src_hi, tb);
      ctx->launches++;
    }
  };
  if (ctx->pipe.empty()) {
    TFS_TRY(launch_small());
    launch(0, ~0ull, 0, n_tiles);
  } else if (!host_segs) {
    TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipe.back().ready, 0));
    ctx->pipe.clear();
    TFS_TRY(launch_small());
    launch(0, ~0ull, 0, n_tiles);
  } else {
    // the sources are still arriving over PCIe, chunk by chunk: one launch per chunk, each behind its chunk's event,
    // over the tiles that can end inside the chunk (two tiles of slack: the kernels apply the exact test)
    uint64_t lo = 0;
    for (size_t i = 0; i < ctx->pipe.size(); ++i) {
      const uint64_t hi = i + 1 == ctx->pipe.size() ? ~0ull : ctx->pipe[i].src_hi;
      uint64_t tb = ~0ull, te = 0;
      for (const HostSeg& h : *host_segs) {
        if (h.src_bytes == 0 || h.n_tiles == 0) continue;
        const uint64_t s0 = h.src, s1 = h.src + h.src_bytes;
        if (s1 <= lo || s0 > hi) continue;  // no source byte of this segment ends inside the window
        const uint64_t from = lo > s0 ? (lo - s0) / h.src_tile_max : 0;
        const uint64_t upto = hi >= s1 ? h.n_tiles : std::min<uint64_t>(h.n_tiles, (hi - s0) / h.src_tile_min + 2);
        tb = std::min(tb, h.first_tile + (from > 1 ? from - 1 : 0));
        te = std::max(te, h.first_tile + upto);
      }
      if (te > tb) {
        TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipe[i].ready, 0));
        launch(lo, hi, tb, std::min(te, n_tiles));
      }
      lo = hi;
    }
    if (any_small) TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipe.back().ready, 0));  // (their sources may sit in any chunk)
    ctx->pipe.clear();
    TFS_TRY(launch_small());
  }
  TFS_CUDA(ctx, cudaGetLastError());
  return TFS_OK;
}

tfs_status_t launch_copy_tma_deferred(tfs_ctx* ctx, const CopySeg* segs_dev, const uint64_t* go, const uint64_t* n_segs_dev,
                                      const uint64_t* n_tiles_dev, bool many_segs) {
  const uint32_t ring_bytes = kTmaStages * kTmaChunk;
  if (!ctx->copy_attr_set) {
    TFS_CUDA(ctx, cudaFuncSetAttribute(k_copy_tma<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(ring_bytes)));
    TFS_CUDA(ctx, cudaFuncSetAttribute(k_copy_tma<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(ring_bytes)));
    TFS_CUDA(ctx, cudaFuncSetAttribute(k_copy_cast, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(3 * kCopyTileBytes + 2 * kCastPad)));
    ctx->copy_attr_set = true;
  }
  DeferredCounts d;
  d.go = go;
  d.n_segs = n_segs_dev;
  d.n_tiles = n_tiles_dev;
  {
    ProfScope prof(ctx, PK_COPY_TMA);
    const uint32_t grid = static_cast<uint32_t>(ctx->sm_count);
    if (many_segs) k_copy_tma<true><<<grid, 32, ring_bytes, ctx->stream>>>(segs_dev, 0u, 0ull, nullptr, 0ull, ~0ull, 0ull, d);
    else k_copy_tma<false><<<grid, 32, ring_bytes, ctx->stream>>>(segs_dev, 0u, 0ull, nullptr, 0ull, ~0ull, 0ull, d);
  }
  ctx->launches++;
  TFS_CUDA(ctx, cudaGetLastError());
  return TFS_OK;
}

}  // namespace tfs
