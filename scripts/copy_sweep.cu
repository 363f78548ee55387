// copy_sweep.cu — development microbenchmark (not part of the library): what is the ceiling of a device-to-device
// copy of one C2 batch (16 x 38.5 MB) on this B200, and which kernel shape reaches it?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o copy_sweep scripts/copy_sweep.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#define CK(x)                                                                  \
  do {                                                                         \
    cudaError_t e = (x);                                                       \
    if (e != cudaSuccess) {                                                    \
      std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      std::exit(1);                                                            \
    }                                                                          \
  } while (0)

__device__ __forceinline__ uint4 ld16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st16(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// plain vector copy, UNROLL x 16 B per thread per iteration, grid-stride
template <int UNROLL>
__global__ void __launch_bounds__(256) k_vec(const uint4* __restrict__ src, uint4* __restrict__ dst, uint64_t n16) {
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + (UNROLL - 1) * stride < n16; i += UNROLL * stride) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) v[u] = ld16(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) st16(dst + i + u * stride, v[u]);
  }
  for (; i < n16; i += stride) st16(dst + i, ld16(src + i));
}

// blocked: each CTA copies one contiguous range
template <int UNROLL>
__global__ void __launch_bounds__(256) k_vec_blocked(const uint4* __restrict__ src, uint4* __restrict__ dst, uint64_t n16) {
  const uint64_t b0 = n16 * blockIdx.x / gridDim.x, b1 = n16 * (blockIdx.x + 1) / gridDim.x;
  uint64_t i = b0 + threadIdx.x;
  for (; i + (UNROLL - 1) * 256 < b1; i += UNROLL * 256) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) v[u] = ld16(src + i + u * 256);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) st16(dst + i + u * 256, v[u]);
  }
  for (; i < b1; i += 256) st16(dst + i, ld16(src + i));
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred P1;\nLAB_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra DONE;\nbra LAB_WAIT;\nDONE:\n}" ::"r"(
          smem_u32(bar)),
      "r"(parity)
      : "memory");
}
template <bool HINT>
__device__ __forceinline__ void tma_load(void* s, const void* g, uint32_t bytes, uint64_t* bar, uint64_t pol) {
  if (HINT)
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(s)),
                 "l"(g), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
  else
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(s)), "l"(g), "r"(bytes),
                 "r"(smem_u32(bar)) : "memory");
}
template <bool HINT>
__device__ __forceinline__ void tma_store(void* g, const void* s, uint32_t bytes, uint64_t pol) {
  if (HINT)
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(g), "r"(smem_u32(s)), "r"(bytes), "l"(pol) : "memory");
  else
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(s)), "r"(bytes) : "memory");
}

// TMA ring: one thread, STAGES x CHUNK bytes of shared memory, contiguous range per CTA
template <int STAGES, int CHUNK, bool HINT>
__global__ void __launch_bounds__(32) k_tma(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint64_t bytes) {
  extern __shared__ __align__(128) uint8_t ring[];
  __shared__ __align__(8) uint64_t bar[STAGES];
  if (threadIdx.x != 0) return;
  const uint64_t nch = (bytes + CHUNK - 1) / CHUNK;
  const uint64_t c0 = nch * blockIdx.x / gridDim.x, c1 = nch * (blockIdx.x + 1) / gridDim.x;
  if (c0 >= c1) return;
  for (int s = 0; s < STAGES; ++s) mbar_init(&bar[s], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  uint64_t pol_ld = 0, pol_st = 0;
  if (HINT) {
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_ld));
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_st));
  }
  uint64_t issued = c0, stored = c0;
  auto size_of = [&](uint64_t c) { return static_cast<uint32_t>((c + 1) * CHUNK <= bytes ? CHUNK : bytes - c * CHUNK); };
  auto issue = [&]() {
    if (issued >= c1) return;
    const uint32_t s = static_cast<uint32_t>((issued - c0) % STAGES);
    const uint32_t n = size_of(issued);
    mbar_expect_tx(&bar[s], n);
    tma_load<HINT>(ring + s * CHUNK, src + issued * CHUNK, n, &bar[s], pol_ld);
    ++issued;
  };
  for (int k = 0; k + 1 < STAGES; ++k) issue();
  while (stored < c1) {
    const uint64_t j = stored - c0;
    const uint32_t s = static_cast<uint32_t>(j % STAGES);
    mbar_wait(&bar[s], static_cast<uint32_t>(j / STAGES) & 1u);
    tma_store<HINT>(dst + stored * CHUNK, ring + s * CHUNK, size_of(stored), pol_st);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    ++stored;
    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    issue();
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// TMA ring, chunks dealt round-robin to the CTAs (chunk c -> CTA c % grid): all CTAs sweep one compact window
template <int STAGES, int CHUNK>
__global__ void __launch_bounds__(32) k_tma_strided(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint64_t bytes) {
  extern __shared__ __align__(128) uint8_t ring[];
  __shared__ __align__(8) uint64_t bar[STAGES];
  if (threadIdx.x != 0) return;
  const uint64_t nch = (bytes + CHUNK - 1) / CHUNK;
  const uint64_t G = gridDim.x;
  if (blockIdx.x >= nch) return;
  const uint64_t mine = (nch - blockIdx.x + G - 1) / G;  // chunks of this CTA: blockIdx.x + k*G
  for (int s = 0; s < STAGES; ++s) mbar_init(&bar[s], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  uint64_t issued = 0, stored = 0;
  auto chunk_of = [&](uint64_t k) { return blockIdx.x + k * G; };
  auto size_of = [&](uint64_t c) { return static_cast<uint32_t>((c + 1) * CHUNK <= bytes ? CHUNK : bytes - c * CHUNK); };
  auto issue = [&]() {
    if (issued >= mine) return;
    const uint32_t s = static_cast<uint32_t>(issued % STAGES);
    const uint64_t c = chunk_of(issued);
    const uint32_t n = size_of(c);
    mbar_expect_tx(&bar[s], n);
    tma_load<false>(ring + s * CHUNK, src + c * CHUNK, n, &bar[s], 0);
    ++issued;
  };
  for (int k = 0; k + 1 < STAGES; ++k) issue();
  while (stored < mine) {
    const uint32_t s = static_cast<uint32_t>(stored % STAGES);
    mbar_wait(&bar[s], static_cast<uint32_t>(stored / STAGES) & 1u);
    const uint64_t c = chunk_of(stored);
    tma_store<false>(dst + c * CHUNK, ring + s * CHUNK, size_of(c), 0);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    ++stored;
    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    issue();
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// one pass, one CTA per TILE bytes (non-persistent)
template <int TILE>
__global__ void __launch_bounds__(256) k_vec_tile(const uint4* __restrict__ src, uint4* __restrict__ dst, uint64_t n16) {
  constexpr int PER = TILE / 16 / 256;
  const uint64_t base = static_cast<uint64_t>(blockIdx.x) * (TILE / 16) + threadIdx.x;
  uint4 v[PER];
#pragma unroll
  for (int u = 0; u < PER; ++u)
    if (base + u * 256 < n16) v[u] = ld16(src + base + u * 256);
#pragma unroll
  for (int u = 0; u < PER; ++u)
    if (base + u * 256 < n16) st16(dst + base + u * 256, v[u]);
}

template <class F>
float time_it(F f, int reps, uint8_t* flush, size_t flush_bytes) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  float best = 1e30f, sum = 0;
  for (int r = 0; r < reps + 2; ++r) {
    if (flush) CK(cudaMemsetAsync(flush, r, flush_bytes));
    CK(cudaEventRecord(a));
    f();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (r >= 2) {
      best = ms < best ? ms : best;
      sum += ms;
    }
  }
  CK(cudaGetLastError());
  std::printf("best %.4f ms avg %.4f ms", best, sum / reps);
  return best;
}

// host <-> device over PCIe: copy engine vs SM loads/stores vs the TMA ring on pinned (mapped) host memory
int host_mode() {
  const uint64_t bytes = 256ull << 20;
  uint8_t *hsrc, *hdst, *dsrc, *ddst;
  CK(cudaHostAlloc(&hsrc, bytes, cudaHostAllocDefault));
  CK(cudaHostAlloc(&hdst, bytes, cudaHostAllocDefault));
  CK(cudaMalloc(&dsrc, bytes));
  CK(cudaMalloc(&ddst, bytes));
  for (uint64_t i = 0; i < bytes; i += 4096) hsrc[i] = 7;
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const uint64_t n16 = bytes / 16;
  auto rep1 = [&](const char* name, float ms, double factor) { std::printf("  -> %-52s %8.2f GB/s\n", name, factor * bytes / (ms * 1e-3) / 1e9); };
  std::printf("host mode: %.0f MB pinned; GB/s counts payload bytes once per direction\n", bytes / 1e6);
  rep1("cudaMemcpyAsync H2D", time_it([&] { CK(cudaMemcpyAsync(ddst, hsrc, bytes, cudaMemcpyHostToDevice)); }, 5, nullptr, 0), 1);
  rep1("cudaMemcpyAsync D2H", time_it([&] { CK(cudaMemcpyAsync(hdst, dsrc, bytes, cudaMemcpyDeviceToHost)); }, 5, nullptr, 0), 1);
  cudaStream_t s2;
  CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
  cudaEvent_t done2;
  CK(cudaEventCreate(&done2));
  rep1("cudaMemcpyAsync H2D + D2H concurrently (sum)", time_it([&] {
         CK(cudaMemcpyAsync(hdst, dsrc, bytes, cudaMemcpyDeviceToHost, s2));
         CK(cudaEventRecord(done2, s2));
         CK(cudaMemcpyAsync(ddst, hsrc, bytes, cudaMemcpyHostToDevice));
         CK(cudaStreamWaitEvent(0, done2));
       }, 5, nullptr, 0), 2);
  rep1("k_vec<8> host->host grid 16/SM (sum both dirs)", time_it([&] { k_vec<8><<<sms * 16, 256>>>((const uint4*)hsrc, (uint4*)hdst, n16); }, 5, nullptr, 0), 2);
  for (int per_sm : {1, 2, 4, 8}) {
    char nm[96];
    std::snprintf(nm, sizeof nm, "k_vec<4> host->device grid %d/SM", per_sm);
    rep1(nm, time_it([&] { k_vec<4><<<sms * per_sm, 256>>>((const uint4*)hsrc, (uint4*)ddst, n16); }, 5, nullptr, 0), 1);
    std::snprintf(nm, sizeof nm, "k_vec<4> device->host grid %d/SM", per_sm);
    rep1(nm, time_it([&] { k_vec<4><<<sms * per_sm, 256>>>((const uint4*)dsrc, (uint4*)hdst, n16); }, 5, nullptr, 0), 1);
    std::snprintf(nm, sizeof nm, "k_vec<4> host->host grid %d/SM (sum both dirs)", per_sm);
    rep1(nm, time_it([&] { k_vec<4><<<sms * per_sm, 256>>>((const uint4*)hsrc, (uint4*)hdst, n16); }, 5, nullptr, 0), 2);
  }
  rep1("k_vec_tile<32K> host->host one pass (sum)", time_it([&] { k_vec_tile<32768><<<(unsigned)(bytes / 32768), 256>>>((const uint4*)hsrc, (uint4*)hdst, n16); }, 5, nullptr, 0), 2);
  CK(cudaFuncSetAttribute(k_tma_strided<3, 32768>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 32768));
  rep1("k_tma_strided<3,32K> host->device 1/SM", time_it([&] { k_tma_strided<3, 32768><<<sms, 32, 3 * 32768>>>(hsrc, ddst, bytes); }, 5, nullptr, 0), 1);
  rep1("k_tma_strided<3,32K> device->host 1/SM", time_it([&] { k_tma_strided<3, 32768><<<sms, 32, 3 * 32768>>>(dsrc, hdst, bytes); }, 5, nullptr, 0), 1);
  rep1("k_tma_strided<3,32K> host->host 1/SM (sum)", time_it([&] { k_tma_strided<3, 32768><<<sms, 32, 3 * 32768>>>(hsrc, hdst, bytes); }, 5, nullptr, 0), 2);
  // hybrids: one direction by the copy engine, the other by SM stores / loads at the same time
  cudaStream_t s3;
  CK(cudaStreamCreateWithFlags(&s3, cudaStreamNonBlocking));
  cudaEvent_t go, done3;
  CK(cudaEventCreate(&go));
  CK(cudaEventCreate(&done3));
  auto both = [&](auto ce, auto sm) {
    return time_it([&] {
      CK(cudaEventRecord(go, 0));
      CK(cudaStreamWaitEvent(s3, go));
      ce(s3);
      CK(cudaEventRecord(done3, s3));
      sm();
      CK(cudaStreamWaitEvent(0, done3));
    }, 5, nullptr, 0);
  };
  rep1("CE H2D + CE D2H, two non-blocking streams (sum)", both([&](cudaStream_t st) { CK(cudaMemcpyAsync(ddst, hsrc, bytes, cudaMemcpyHostToDevice, st)); },
                                                              [&] { CK(cudaMemcpyAsync(hdst, dsrc, bytes, cudaMemcpyDeviceToHost, s2)); CK(cudaEventRecord(done2, s2)); CK(cudaStreamWaitEvent(0, done2)); }), 2);
  rep1("CE H2D + k_vec<4> device->host 2/SM (sum)", both([&](cudaStream_t st) { CK(cudaMemcpyAsync(ddst, hsrc, bytes, cudaMemcpyHostToDevice, st)); },
                                                        [&] { k_vec<4><<<sms * 2, 256>>>((const uint4*)dsrc, (uint4*)hdst, n16); }), 2);
  rep1("CE H2D + k_tma_strided device->host 1/SM (sum)", both([&](cudaStream_t st) { CK(cudaMemcpyAsync(ddst, hsrc, bytes, cudaMemcpyHostToDevice, st)); },
                                                             [&] { k_tma_strided<3, 32768><<<sms, 32, 3 * 32768>>>(dsrc, hdst, bytes); }), 2);
  rep1("CE D2H + k_vec<4> host->device 2/SM (sum)", both([&](cudaStream_t st) { CK(cudaMemcpyAsync(hdst, dsrc, bytes, cudaMemcpyDeviceToHost, st)); },
                                                        [&] { k_vec<4><<<sms * 2, 256>>>((const uint4*)hsrc, (uint4*)ddst, n16); }), 2);
  rep1("CE D2H + k_tma_strided host->device 1/SM (sum)", both([&](cudaStream_t st) { CK(cudaMemcpyAsync(hdst, dsrc, bytes, cudaMemcpyDeviceToHost, st)); },
                                                             [&] { k_tma_strided<3, 32768><<<sms, 32, 3 * 32768>>>(hsrc, ddst, bytes); }), 2);
  CK(cudaDeviceSynchronize());
  for (uint64_t i = 0; i < bytes; i += 4096)
    if (hdst[i] != 7) {
      std::printf("HOST COPY WRONG at %llu\n", (unsigned long long)i);
      return 1;
    }
  std::printf("host ok\n");
  return 0;
}

int main(int argc, char** argv) {
  if (argc > 1 && std::string(argv[1]) == "host") return host_mode();
  const uint64_t bytes = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 16ull * 38535168ull;
  uint8_t *src, *dst, *flush;
  const size_t flush_bytes = 256u << 20;
  CK(cudaMalloc(&src, bytes + 256));
  CK(cudaMalloc(&dst, bytes + 256));
  CK(cudaMalloc(&flush, flush_bytes));
  CK(cudaMemset(src, 7, bytes));
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  std::printf("bytes %llu (%.1f MB), SMs %d; GB/s counts read+write bytes\n", (unsigned long long)bytes, bytes / 1e6, sms);
  auto report = [&](const char* name, float ms) { std::printf("  -> %-44s %8.1f GB/s\n", name, 2.0 * bytes / (ms * 1e-3) / 1e9); };
  const int reps = 10;
  report("cudaMemcpyAsync D2D", time_it([&] { CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice)); }, reps, flush, flush_bytes));
  const uint64_t n16 = bytes / 16;
  for (int per_sm : {4, 8, 16}) {
    char nm[96];
    std::snprintf(nm, sizeof nm, "k_vec<4> grid %d/SM", per_sm);
    report(nm, time_it([&] { k_vec<4><<<sms * per_sm, 256>>>((const uint4*)src, (uint4*)dst, n16); }, reps, flush, flush_bytes));
    std::snprintf(nm, sizeof nm, "k_vec<8> grid %d/SM", per_sm);
    report(nm, time_it([&] { k_vec<8><<<sms * per_sm, 256>>>((const uint4*)src, (uint4*)dst, n16); }, reps, flush, flush_bytes));
    std::snprintf(nm, sizeof nm, "k_vec_blocked<4> grid %d/SM", per_sm);
    report(nm, time_it([&] { k_vec_blocked<4><<<sms * per_sm, 256>>>((const uint4*)src, (uint4*)dst, n16); }, reps, flush, flush_bytes));
  }
  {
    const uint32_t g = static_cast<uint32_t>((n16 + 256 * 4 - 1) / (256 * 4));
    report("k_vec<4> one pass (grid = n/1024)", time_it([&] { k_vec<4><<<g, 256>>>((const uint4*)src, (uint4*)dst, n16); }, reps, flush, flush_bytes));
  }
#define TMA_CASE(ST, CH, HINT, PER_SM)                                                                              \
  {                                                                                                                 \
    CK(cudaFuncSetAttribute(k_tma<ST, CH, HINT>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST * CH));            \
    char nm[96];                                                                                                    \
    std::snprintf(nm, sizeof nm, "k_tma stages %d chunk %d hint %d grid %d/SM", ST, CH, (int)HINT, PER_SM);         \
    report(nm, time_it([&] { k_tma<ST, CH, HINT><<<sms * PER_SM, 32, ST * CH>>>(src, dst, bytes); }, reps, flush, flush_bytes)); \
  }
  TMA_CASE(4, 16384, false, 3)
  TMA_CASE(4, 16384, true, 3)
  TMA_CASE(4, 16384, false, 2)
  TMA_CASE(4, 16384, false, 1)
  TMA_CASE(8, 8192, false, 3)
  TMA_CASE(8, 8192, false, 2)
  TMA_CASE(6, 32768, false, 1)
  TMA_CASE(3, 32768, false, 2)
  TMA_CASE(4, 8192, false, 6)
  TMA_CASE(4, 4096, false, 8)
  TMA_CASE(8, 4096, false, 6)
  TMA_CASE(2, 32768, false, 3)
#define TMAS_CASE(ST, CH, PER_SM)                                                                                   \
  {                                                                                                                 \
    CK(cudaFuncSetAttribute(k_tma_strided<ST, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST * CH));          \
    char nm[96];                                                                                                    \
    std::snprintf(nm, sizeof nm, "k_tma_strided stages %d chunk %d grid %d/SM", ST, CH, PER_SM);                    \
    report(nm, time_it([&] { k_tma_strided<ST, CH><<<sms * PER_SM, 32, ST * CH>>>(src, dst, bytes); }, reps, flush, flush_bytes)); \
  }
  TMAS_CASE(4, 16384, 1)
  TMAS_CASE(4, 16384, 2)
  TMAS_CASE(4, 16384, 3)
  TMAS_CASE(6, 16384, 1)
  TMAS_CASE(8, 16384, 1)
  TMAS_CASE(3, 32768, 1)
  TMAS_CASE(4, 32768, 1)
  TMAS_CASE(6, 32768, 1)
  TMAS_CASE(3, 32768, 2)
  TMAS_CASE(8, 8192, 1)
  TMAS_CASE(8, 8192, 2)
  TMAS_CASE(3, 65536, 1)
  report("k_vec_tile<16K> one pass", time_it([&] { k_vec_tile<16384><<<(unsigned)((bytes + 16383) / 16384), 256>>>((const uint4*)src, (uint4*)dst, n16); }, reps, flush, flush_bytes));
  report("k_vec_tile<32K> one pass", time_it([&] { k_vec_tile<32768><<<(unsigned)((bytes + 32767) / 32768), 256>>>((const uint4*)src, (uint4*)dst, n16); }, reps, flush, flush_bytes));
  report("k_vec_tile<8K> one pass", time_it([&] { k_vec_tile<8192><<<(unsigned)((bytes + 8191) / 8192), 256>>>((const uint4*)src, (uint4*)dst, n16); }, reps, flush, flush_bytes));
  // sanity: the last copy is correct
  std::vector<uint8_t> h(1024);
  CK(cudaMemcpy(h.data(), dst + bytes - 1024, 1024, cudaMemcpyDeviceToHost));
  for (auto b : h)
    if (b != 7) {
      std::printf("COPY WRONG\n");
      return 1;
    }
  std::printf("ok\n");
  return 0;
}
