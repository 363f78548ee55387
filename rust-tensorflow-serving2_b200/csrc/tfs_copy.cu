// tfs_copy.cu — K5: bulk copy / cast over a table of segments (fixed-width packed fields, tensor_content,
// bool bodies, fused u8->f32 and f32<->f16/bf16 casts).
//
// Replaces, on the reference path (src/lib.rs:265): prost's per-element `put_f32_le` / `get_f32_le`
// loops of float/double::{encode_packed, merge_repeated}, bytes::{encode, merge}, and the client's own
// `*p as f32` + preprocessing_fn loop (src/lib.rs:237-242).
//
// HBM-bound byte work: no tensor cores.  One CTA processes one 32 KiB output tile; the destination is
// always written with 16-byte aligned vector stores, the source is read with 16-byte aligned vector
// loads and re-aligned in registers with funnel shifts (wire payloads start at arbitrary byte offsets).
#include "tfs_device.cuh"

namespace tfs {

namespace {

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

// Copy `len` bytes src -> dst (any alignment on both sides), cooperatively by the whole CTA.
__device__ __forceinline__ void cta_copy_bytes(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint32_t len) {
  const uint32_t tid = threadIdx.x;
  uint32_t head = (16u - static_cast<uint32_t>(reinterpret_cast<uintptr_t>(dst) & 15u)) & 15u;
  if (head > len) head = len;
  if (tid < head) dst[tid] = src[tid];
  const uint32_t body = len - head;
  const uint32_t nvec = body >> 4;
  uint8_t* d = dst + head;
  const uint8_t* s = src + head;
  const uint32_t delta = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(s) & 15u);
  if (delta == 0) {
    uint32_t c = tid;
    for (; c + 3 * kCopyThreads < nvec; c += 4 * kCopyThreads) {
      uint4 v0 = ld_stream16(s + 16ull * c);
      uint4 v1 = ld_stream16(s + 16ull * (c + kCopyThreads));
      uint4 v2 = ld_stream16(s + 16ull * (c + 2 * kCopyThreads));
      uint4 v3 = ld_stream16(s + 16ull * (c + 3 * kCopyThreads));
      st_stream16(d + 16ull * c, v0);
      st_stream16(d + 16ull * (c + kCopyThreads), v1);
      st_stream16(d + 16ull * (c + 2 * kCopyThreads), v2);
      st_stream16(d + 16ull * (c + 3 * kCopyThreads), v3);
    }
    for (; c < nvec; c += kCopyThreads) st_stream16(d + 16ull * c, ld_stream16(s + 16ull * c));
  } else {
    const uint8_t* sa = s - delta;  // 16-byte aligned; every block read below overlaps valid source bytes
    uint32_t c = tid;
    for (; c + kCopyThreads < nvec; c += 2 * kCopyThreads) {
      uint4 a0 = ld_stream16(sa + 16ull * c);
      uint4 b0 = ld_stream16(sa + 16ull * c + 16);
      uint4 a1 = ld_stream16(sa + 16ull * (c + kCopyThreads));
      uint4 b1 = ld_stream16(sa + 16ull * (c + kCopyThreads) + 16);
      st_stream16(d + 16ull * c, shift_window(a0, b0, delta));
      st_stream16(d + 16ull * (c + kCopyThreads), shift_window(a1, b1, delta));
    }
    for (; c < nvec; c += kCopyThreads) {
      uint4 a = ld_stream16(sa + 16ull * c);
      uint4 b = ld_stream16(sa + 16ull * c + 16);
      st_stream16(d + 16ull * c, shift_window(a, b, delta));
    }
  }
  const uint32_t tail0 = head + (nvec << 4);
  if (tid < len - tail0) dst[tail0 + tid] = src[tail0 + tid];
}

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

__global__ void __launch_bounds__(kCopyThreads) k_copy(const CopySeg* __restrict__ segs, uint32_t n_segs,
                                                        uint64_t n_tiles, const uint64_t* __restrict__ n_tiles_dev,
                                                        const uint64_t* __restrict__ gate) {
  __shared__ uint32_t s_seg;
  if (gate && *gate != 0) return;  // an earlier kernel of the batch refused to go on (dst too small)
  if (n_tiles_dev) n_tiles = *n_tiles_dev;
  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    __syncthreads();  // s_seg reuse across iterations
    if (threadIdx.x == 0) {
      uint32_t lo = 0, hi = n_segs;
      while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (segs[mid].first_tile <= tile) lo = mid;
        else hi = mid;
      }
      s_seg = lo;
    }
    __syncthreads();
    const CopySeg seg = segs[s_seg];
    const uint32_t osz = out_size_of(seg.op), isz = in_size_of(seg.op);
    const uint64_t elems_per_tile = kCopyTileBytes / osz;
    const uint64_t e0 = (tile - seg.first_tile) * elems_per_tile;
    if (e0 >= seg.n) continue;
    const uint32_t n = static_cast<uint32_t>(min(elems_per_tile, seg.n - e0));
    uint8_t* d = seg.dst + e0 * osz;
    const uint8_t* s = seg.src + e0 * isz;
    const float p = seg.param;
    switch (seg.op) {
      case CP_BYTES: cta_copy_bytes(d, s, n); break;
      case CP_BOOL: cta_cast<uint8_t, uint8_t>(d, s, n, [](uint8_t x) { return static_cast<uint8_t>(x != 0); }); break;
      case CP_U8_F32: cta_cast<uint8_t, float>(d, s, n, [](uint8_t x) { return static_cast<float>(x); }); break;
      case CP_U8_F32_DIV: cta_cast<uint8_t, float>(d, s, n, [p](uint8_t x) { return __fdiv_rn(static_cast<float>(x), p); }); break;
      case CP_U8_F32_MUL: cta_cast<uint8_t, float>(d, s, n, [p](uint8_t x) { return __fmul_rn(static_cast<float>(x), p); }); break;
      case CP_F32_F16: cta_cast<float, uint16_t>(d, s, n, [](float x) { return static_cast<uint16_t>(f32_to_f16_bits(x)); }); break;
      case CP_F32_BF16: cta_cast<float, uint16_t>(d, s, n, [](float x) { return static_cast<uint16_t>(f32_to_bf16_bits(x)); }); break;
      case CP_F16_F32: cta_cast<uint16_t, float>(d, s, n, [](uint16_t x) { return f16_bits_to_f32(x); }); break;
      case CP_BF16_F32: cta_cast<uint16_t, float>(d, s, n, [](uint16_t x) { return bf16_bits_to_f32(x); }); break;
      default: break;
    }
  }
}

}  // namespace

tfs_status_t launch_copy(tfs_ctx* ctx, const CopySeg* segs_dev, uint32_t n_segs, uint64_t n_tiles,
                         const uint64_t* n_tiles_dev, const uint64_t* gate) {
  if (n_segs == 0 || (n_tiles == 0 && !n_tiles_dev)) return TFS_OK;
  uint64_t cap = static_cast<uint64_t>(ctx->sm_count) * 8;  // 8 resident 256-thread CTAs per SM
  uint32_t grid = static_cast<uint32_t>(n_tiles_dev ? cap : (n_tiles < cap ? n_tiles : cap));
  {
    ProfScope prof(ctx, PK_COPY);
    k_copy<<<grid, kCopyThreads, 0, ctx->stream>>>(segs_dev, n_segs, n_tiles, n_tiles_dev, gate);
  }
  ctx->launches++;
  TFS_CUDA(ctx, cudaGetLastError());
  return TFS_OK;
}

}  // namespace tfs
