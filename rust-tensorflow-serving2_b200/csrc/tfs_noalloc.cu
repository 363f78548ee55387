// tfs_noalloc.cu — tfs_pack_int64_noalloc: packed-varint encoding of an int64 array (the body of
// `int64_val`, tensor.proto:64) for callers that may not allocate GPU memory.
//
// Who that is.  The regular entry points (tfs_encode & co.) keep their tile tables in buffers the context
// grows with cudaMalloc when a bigger batch arrives.  Two kinds of caller cannot live with that:
//   * a codec called from inside tonic's poll loop on a serving path: cudaMalloc synchronises the whole
//     device, so one late growth stalls the kernels of every other request in flight on that GPU;
//   * a client that shares the GPU with a TensorFlow Serving process which has already claimed almost all
//     of the device memory: an allocation made after start-up simply fails.
// For them this one function works out of a fixed block of device memory that is part of the library image
// itself: it is reserved when the library is loaded (on every GPU that loads it) and released at exit.
//
// The price of a fixed block:
//   * at most TFS_NOALLOC_MAX_ELEMS elements per call (16 Mi values = 128 MiB of int64); larger inputs are
//     refused with TFS_ERR_TOO_LARGE and nothing is launched;
//   * every call uses the SAME block.  Two calls whose kernels can overlap on the GPU (different streams
//     with no dependency between them, or two host threads) race on it and corrupt each other's output.
//     Issue the calls on one stream, or order them yourself.  Nothing is carried from one call to the
//     next: each call first resets the status words it relies on and overwrites every table entry it
//     later reads, so stale contents of an earlier (even larger) call are harmless.
//
// First use on a device: the function looks up the device address of the block with
// cudaGetSymbolAddress the first time it is called while that device is current, and remembers it.
// Make that first call outside of any stream capture (it is a plain, synchronous runtime query).
// The function itself calls no cudaMalloc / cudaFree and no library routine that allocates.
#include <mutex>

#include "tfs_device.cuh"

namespace tfs {
namespace {

constexpr uint32_t kNaTile = kVarintTile;  // 2048 elements per tile, 8 warps x 8 rows x 32 lanes
constexpr uint32_t kNaMaxTiles = static_cast<uint32_t>(TFS_NOALLOC_MAX_ELEMS / kNaTile);

// Layout of the fixed block: [0,256) status words, then the per-tile byte counts, then the per-tile offsets.
struct NaHeader {
  unsigned long long total;     // bytes of the packed body
  unsigned long long overflow;  // 1 when the body does not fit dst_capacity
};
constexpr size_t kNaHeaderBytes = 256;
constexpr size_t kNaCountsBytes = sizeof(uint32_t) * kNaMaxTiles;
constexpr size_t kNaOffsetsBytes = sizeof(unsigned long long) * kNaMaxTiles;
constexpr size_t kNaScratchBytes = kNaHeaderBytes + kNaCountsBytes + kNaOffsetsBytes;  // 98 560 bytes

// The block itself.  In host code this name is only a handle for the runtime, not an address the GPU
// understands; kernels receive the address obtained with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaScratchBytes];

__device__ __forceinline__ NaHeader* na_header(unsigned char* s) { return reinterpret_cast<NaHeader*>(s); }
__device__ __forceinline__ uint32_t* na_counts(unsigned char* s) { return reinterpret_cast<uint32_t*>(s + kNaHeaderBytes); }
__device__ __forceinline__ unsigned long long* na_offsets(unsigned char* s) {
  return reinterpret_cast<unsigned long long*>(s + kNaHeaderBytes + kNaCountsBytes);
}

__device__ __forceinline__ void na_load_rows(const long long* __restrict__ src, uint64_t n, uint64_t elem0,
                                             uint64_t (&v)[kVarintRows], uint32_t (&len)[kVarintRows]) {
#pragma unroll
  for (int r = 0; r < kVarintRows; ++r) {
    const uint64_t e = elem0 + static_cast<uint64_t>(r) * 32u + lane_id();
    const bool ok = e < n;
    v[r] = ok ? static_cast<uint64_t>(__ldg(src + e)) : 0ull;  // `as u64` of an i64: the bits are the value
    len[r] = ok ? varint_len(v[r]) : 0u;
  }
}

// pass 1: bytes each tile will produce; block 0 also resets the status words of this call
__global__ void __launch_bounds__(kVarintThreads) k_na_size(const long long* __restrict__ src, uint64_t n, unsigned char* scratch) {
  __shared__ uint32_t warp_tot[kVarintThreads / 32];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    na_header(scratch)->total = 0;
    na_header(scratch)->overflow = 0;
  }
  uint64_t v[kVarintRows];
  uint32_t len[kVarintRows];
  na_load_rows(src, n, static_cast<uint64_t>(blockIdx.x) * kNaTile + warp_id() * (kVarintRows * 32u), v, len);
  uint32_t s = 0;
#pragma unroll
  for (int r = 0; r < kVarintRows; ++r) s += len[r];
  s = __reduce_add_sync(0xFFFFFFFFu, s);
  if (lane_id() == 0) warp_tot[warp_id()] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < kVarintThreads / 32; ++w) t += warp_tot[w];
    na_counts(scratch)[blockIdx.x] = t;
  }
}

// pass 2: exclusive scan of the tile byte counts (one CTA), body length out, capacity check
__global__ void __launch_bounds__(1024) k_na_scan(unsigned char* scratch, uint32_t n_tiles, uint64_t dst_capacity,
                                                   uint64_t* __restrict__ body_len_out) {
  __shared__ unsigned long long smem[33];
  unsigned long long carry = 0;
  for (uint32_t base = 0; base < n_tiles; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const unsigned long long x = i < n_tiles ? na_counts(scratch)[i] : 0ull;
    unsigned long long total;
    const unsigned long long ex = block_exclusive_scan<unsigned long long, 32>(x, smem, &total);
    if (i < n_tiles) na_offsets(scratch)[i] = carry + ex;
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    na_header(scratch)->total = carry;
    na_header(scratch)->overflow = carry > dst_capacity ? 1ull : 0ull;
    *body_len_out = carry;
  }
}

// pass 3: write the varint bytes of each tile at its offset (staged in shared memory, flushed with
// 16-byte stores; the ragged first / last bytes of a tile are stored one by one)
__global__ void __launch_bounds__(kVarintThreads) k_na_scatter(const long long* __restrict__ src, uint64_t n,
                                                               uint8_t* __restrict__ dst, unsigned char* scratch) {
  __shared__ __align__(16) uint8_t stage[16 + kNaTile * kVarintMaxLen + 16];
  __shared__ uint32_t warp_tot[kVarintThreads / 32];
  if (na_header(scratch)->overflow != 0) return;  // does not fit: write nothing
  uint64_t v[kVarintRows];
  uint32_t len[kVarintRows];
  na_load_rows(src, n, static_cast<uint64_t>(blockIdx.x) * kNaTile + warp_id() * (kVarintRows * 32u), v, len);
  const uint32_t lt = (1u << lane_id()) - 1u;
  uint32_t off[kVarintRows];
  uint32_t run = 0;
#pragma unroll
  for (int r = 0; r < kVarintRows; ++r) {
    const uint32_t b0 = __ballot_sync(0xFFFFFFFFu, len[r] & 1u), b1 = __ballot_sync(0xFFFFFFFFu, len[r] & 2u);
    const uint32_t b2 = __ballot_sync(0xFFFFFFFFu, len[r] & 4u), b3 = __ballot_sync(0xFFFFFFFFu, len[r] & 8u);
    off[r] = run + __popc(b0 & lt) + 2u * __popc(b1 & lt) + 4u * __popc(b2 & lt) + 8u * __popc(b3 & lt);
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
  uint8_t* g = dst + na_offsets(scratch)[blockIdx.x];
  const uint32_t skew = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(g) & 15u);
#pragma unroll
  for (int r = 0; r < kVarintRows; ++r) {
    if (len[r]) {
      uint8_t* p = stage + skew + wbase + off[r];
      uint64_t x = v[r];
      while (x >= 0x80) {
        *p++ = static_cast<uint8_t>((x & 0x7F) | 0x80);
        x >>= 7;
      }
      *p = static_cast<uint8_t>(x);
    }
  }
  __syncthreads();
  uint8_t* ga = g - skew;
  const uint32_t end = skew + total;
  const uint32_t nchunks = (end + 15u) >> 4;
  for (uint32_t c = threadIdx.x; c < nchunks; c += kVarintThreads) {
    const uint32_t lo = c << 4, hi = lo + 16u;
    if (lo >= skew && hi <= end) {
      *reinterpret_cast<uint4*>(ga + lo) = *reinterpret_cast<const uint4*>(stage + lo);
    } else {
      const uint32_t b0 = lo > skew ? lo : skew, b1 = hi < end ? hi : end;
      for (uint32_t b = b0; b < b1; ++b) ga[b] = stage[b];
    }
  }
}

// set-up, once per device: the device address of the fixed block
constexpr int kNaMaxDevices = 64;
std::mutex g_na_mutex;
unsigned char* g_na_ptr[kNaMaxDevices] = {nullptr};

tfs_status_t na_scratch_for_current_device(unsigned char** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kNaMaxDevices) return TFS_ERR_CUDA;
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // with `dev` current
    if (e != cudaSuccess || p == nullptr) return TFS_ERR_CUDA;
    g_na_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_na_ptr[dev];
  return TFS_OK;
}

}  // namespace
}  // namespace tfs

using namespace tfs;

extern "C" tfs_status_t tfs_pack_int64_noalloc(const int64_t* src_dev, uint64_t n_elems, void* dst_dev, uint64_t dst_capacity,
                                               uint64_t* body_len_dev, void* cuda_stream) {
  if (!body_len_dev || (n_elems && !src_dev) || (dst_capacity && !dst_dev)) return TFS_ERR_INVALID_ARG;
  if (n_elems > TFS_NOALLOC_MAX_ELEMS) return TFS_ERR_TOO_LARGE;  // refuse before anything is launched
  unsigned char* scratch = nullptr;
  const tfs_status_t st = na_scratch_for_current_device(&scratch);
  if (st != TFS_OK) return st;
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  const uint32_t n_tiles = static_cast<uint32_t>((n_elems + kNaTile - 1) / kNaTile);
  const long long* src = reinterpret_cast<const long long*>(src_dev);
  if (n_tiles) k_na_size<<<n_tiles, kVarintThreads, 0, stream>>>(src, n_elems, scratch);
  k_na_scan<<<1, 1024, 0, stream>>>(scratch, n_tiles, dst_capacity, body_len_dev);
  if (n_tiles) k_na_scatter<<<n_tiles, kVarintThreads, 0, stream>>>(src, n_elems, static_cast<uint8_t*>(dst_dev), scratch);
  return cudaGetLastError() == cudaSuccess ? TFS_OK : TFS_ERR_CUDA;
}
