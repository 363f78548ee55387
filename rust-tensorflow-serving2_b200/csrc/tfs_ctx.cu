// tfs_ctx.cu — context, memory plumbing and the small host-only helpers of include/tfs_gpu.h.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>

#include "tfs_internal.h"

namespace tfs {

tfs_status_t set_err(tfs_ctx* ctx, tfs_status_t st, const std::string& what) {
  if (ctx) ctx->err = what;
  return st;
}

tfs_status_t cuda_fail(tfs_ctx* ctx, cudaError_t e, const char* where) {
  if (ctx) ctx->err = std::string(where) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
  return TFS_ERR_CUDA;
}

// Grow-only device buffer. Growing synchronises the ctx stream first: kernels in flight may still be
// reading the old allocation.
tfs_status_t dev_reserve(tfs_ctx* ctx, DevBuf* b, size_t bytes) {
  if (bytes <= b->cap) return TFS_OK;
  size_t want = std::max(bytes + bytes / 2, static_cast<size_t>(1) << 16);
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (b->p) TFS_CUDA(ctx, cudaFree(b->p));
  b->p = nullptr;
  b->cap = 0;
  TFS_CUDA(ctx, cudaMalloc(&b->p, want));
  b->cap = want;
  return TFS_OK;
}

tfs_status_t pinned_reserve(tfs_ctx* ctx, void** p, size_t* cap, size_t bytes) {
  if (bytes <= *cap) return TFS_OK;
  size_t want = std::max(bytes + bytes / 2, static_cast<size_t>(1) << 16);
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (*p) TFS_CUDA(ctx, cudaFreeHost(*p));
  *p = nullptr;
  *cap = 0;
  TFS_CUDA(ctx, cudaHostAlloc(p, want, cudaHostAllocDefault));
  *cap = want;
  return TFS_OK;
}

namespace {
std::mutex g_down_mutex;
std::vector<PinnedBlock> g_down_free;            // at most kDownKeep blocks, kDownKeepBytes in total
constexpr size_t kDownKeep = 8;
constexpr size_t kDownKeepBytes = size_t(2) << 30;
}  // namespace

tfs_status_t down_acquire(tfs_ctx* ctx, size_t bytes, PinnedBlock* out) {
  {
    std::lock_guard<std::mutex> lock(g_down_mutex);
    int best = -1;
    for (size_t i = 0; i < g_down_free.size(); ++i)
      if (g_down_free[i].cap >= bytes && (best < 0 || g_down_free[i].cap < g_down_free[best].cap)) best = static_cast<int>(i);
    if (best >= 0) {
      *out = g_down_free[best];
      g_down_free.erase(g_down_free.begin() + best);
      return TFS_OK;
    }
  }
  const size_t want = std::max(bytes + bytes / 2, static_cast<size_t>(1) << 16);
  void* p = nullptr;
  TFS_CUDA(ctx, cudaHostAlloc(&p, want, cudaHostAllocPortable));
  out->p = p;
  out->cap = want;
  return TFS_OK;
}

void down_release(PinnedBlock b) {
  if (!b.p) return;
  PinnedBlock drop;
  {
    std::lock_guard<std::mutex> lock(g_down_mutex);
    g_down_free.push_back(b);
    size_t total = 0;
    for (const PinnedBlock& x : g_down_free) total += x.cap;
    if (g_down_free.size() > kDownKeep || total > kDownKeepBytes) {  // let the smallest go (a large one is the expensive one to get back)
      size_t k = 0;
      for (size_t i = 1; i < g_down_free.size(); ++i)
        if (g_down_free[i].cap < g_down_free[k].cap) k = i;
      if (total > kDownKeepBytes && g_down_free.size() <= kDownKeep) {  // over the byte budget: the largest goes instead
        for (size_t i = 0; i < g_down_free.size(); ++i)
          if (g_down_free[i].cap > g_down_free[k].cap) k = i;
      }
      drop = g_down_free[k];
      g_down_free.erase(g_down_free.begin() + k);
    }
  }
  if (drop.p) cudaFreeHost(drop.p);
}

ProfScope::ProfScope(tfs_ctx* c, int kernel_id) : ctx(c), id(kernel_id) {
  if (!ctx->prof_on) return;
  auto take = [&]() {
    cudaEvent_t e = nullptr;
    if (!ctx->prof_pool.empty()) {
      e = ctx->prof_pool.back();
      ctx->prof_pool.pop_back();
    } else if (cudaEventCreate(&e) != cudaSuccess) {
      e = nullptr;
    }
    return e;
  };
  start = take();
  stop = take();
  if (start && stop) cudaEventRecord(start, ctx->stream);
}
ProfScope::~ProfScope() {
  if (!start || !stop) return;
  cudaEventRecord(stop, ctx->stream);
  ctx->prof_open[id].emplace_back(start, stop);
}

}  // namespace tfs

using namespace tfs;

extern "C" {

tfs_status_t tfs_prof_enable(tfs_ctx_t* ctx, int32_t on) {
  if (!ctx) return TFS_ERR_INVALID_ARG;
  ctx->prof_on = on != 0;
  return TFS_OK;
}

// Fold every finished (start, stop) pair into the totals and report them; synchronises the stream.
tfs_status_t tfs_prof_read(tfs_ctx_t* ctx, int32_t kernel_id, uint64_t* launches_out, double* ms_out, int32_t reset) {
  if (!ctx || kernel_id < 0 || kernel_id >= PK_COUNT_) return TFS_ERR_INVALID_ARG;
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (auto& pr : ctx->prof_open[kernel_id]) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) {
      ctx->prof_ms[kernel_id] += ms;
      ctx->prof_launches[kernel_id]++;
    }
    ctx->prof_pool.push_back(pr.first);
    ctx->prof_pool.push_back(pr.second);
  }
  ctx->prof_open[kernel_id].clear();
  if (launches_out) *launches_out = ctx->prof_launches[kernel_id];
  if (ms_out) *ms_out = ctx->prof_ms[kernel_id];
  if (reset) {
    ctx->prof_launches[kernel_id] = 0;
    ctx->prof_ms[kernel_id] = 0;
  }
  return TFS_OK;
}

int32_t tfs_abi_version(void) { return TFS_ABI_VERSION; }

const char* tfs_strerror(tfs_status_t status) {
  switch (status) {
    case TFS_OK: return "ok";
    case TFS_ERR_INVALID_VARINT: return "invalid varint";
    case TFS_ERR_BUFFER_UNDERFLOW: return "buffer underflow";
    case TFS_ERR_LENGTH_EXCEEDED: return "delimited length exceeded";
    case TFS_ERR_WIRE_TYPE: return "invalid wire type";
    case TFS_ERR_WIRE_TYPE_VALUE: return "invalid wire type value";
    case TFS_ERR_INVALID_KEY: return "invalid key value";
    case TFS_ERR_INVALID_TAG: return "invalid tag value: 0";
    case TFS_ERR_RECURSION: return "recursion limit reached";
    case TFS_ERR_END_GROUP: return "unexpected end group tag";
    case TFS_ERR_UTF8: return "invalid string value: data is not UTF-8 encoded";
    case TFS_ERR_GRPC_FLAG: return "grpc frame: compression flag set";
    case TFS_ERR_UNSUPPORTED: return "message uses a field this library leaves to prost";
    case TFS_ERR_INVALID_ARG: return "invalid argument";
    case TFS_ERR_NOMEM: return "out of memory or capacity too small";
    case TFS_ERR_CUDA: return "CUDA error";
    case TFS_ERR_BATCH: return "some messages of the batch failed";
    case TFS_ERR_TOO_LARGE: return "input exceeds TFS_NOALLOC_MAX_ELEMS, the fixed working block of tfs_pack_int64_noalloc";
    default: return "unknown status";
  }
}

tfs_status_t tfs_ctx_create(int32_t device, uint64_t pinned_bytes, tfs_ctx_t** out) {
  if (!out) return TFS_ERR_INVALID_ARG;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || device < 0 || device >= n) {
    // no CPU fallback: without a device there is no codec
    std::fprintf(stderr, "tfs_ctx_create: no usable CUDA device %d (%s)\n", device,
                 e != cudaSuccess ? cudaGetErrorString(e) : "index out of range");
    return TFS_ERR_CUDA;
  }
  tfs_ctx* ctx = new tfs_ctx();
  ctx->device = device;
  ctx->trace = std::getenv("TFS_TRACE") != nullptr;
  if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking)) != cudaSuccess) {
    std::fprintf(stderr, "tfs_ctx_create: %s\n", cudaGetErrorString(e));
    delete ctx;
    return TFS_ERR_CUDA;
  }
  ctx->stream = ctx->own_stream;
  cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
  if (cudaMalloc(reinterpret_cast<void**>(&ctx->d_ctrl), 256) != cudaSuccess || cudaMemset(ctx->d_ctrl, 0, 256) != cudaSuccess) {
    std::fprintf(stderr, "tfs_ctx_create: no device memory for the control words\n");
    tfs_ctx_destroy(ctx);
    return TFS_ERR_CUDA;
  }
  if (pinned_bytes) {  // the staging pool: pinned once, here (exactly the requested size)
    tfs_status_t st = TFS_OK;
    if (cudaHostAlloc(&ctx->h_slab, pinned_bytes, cudaHostAllocDefault) != cudaSuccess) {
      cudaGetLastError();
      st = TFS_ERR_NOMEM;
    } else {
      ctx->h_slab_cap = pinned_bytes;
    }
    if (st != TFS_OK) {
      tfs_ctx_destroy(ctx);
      return st;
    }
  }
  *out = ctx;
  return TFS_OK;
}

tfs_status_t tfs_ctx_destroy(tfs_ctx_t* ctx) {
  if (!ctx) return TFS_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (DevBuf* b : {&ctx->d_tables, &ctx->d_work, &ctx->d_stage_in, &ctx->d_stage_out, &ctx->d_dec1, &ctx->d_dec2, &ctx->d_dec3})
    if (b->p) cudaFree(b->p);
  if (ctx->d_ctrl) cudaFree(ctx->d_ctrl);
  if (ctx->h_up) cudaFreeHost(ctx->h_up);
  if (ctx->h_down) cudaFreeHost(ctx->h_down);
  if (ctx->h_rec) cudaFreeHost(ctx->h_rec);
  ctx->stage_crew.reset();  // joins the helper threads
  if (ctx->enc_layout_ev) cudaEventDestroy(ctx->enc_layout_ev);
  if (ctx->dec_tables_ev) cudaEventDestroy(ctx->dec_tables_ev);
  if (ctx->h_ring) cudaFreeHost(ctx->h_ring);
  for (cudaEvent_t e : ctx->ring_ev)
    if (e) cudaEventDestroy(e);
  if (ctx->h_slab) cudaFreeHost(ctx->h_slab);
  for (auto& e : ctx->pool_extra) cudaFreeHost(e.first);
  if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
  for (cudaEvent_t e : ctx->pipe_events) cudaEventDestroy(e);
  for (cudaEvent_t e : ctx->prof_pool) cudaEventDestroy(e);
  for (auto& open : ctx->prof_open)
    for (auto& pr : open) {
      cudaEventDestroy(pr.first);
      cudaEventDestroy(pr.second);
    }
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
  return TFS_OK;
}

const char* tfs_last_error_detail(const tfs_ctx_t* ctx) { return ctx ? ctx->err.c_str() : ""; }

tfs_status_t tfs_last_error_location(const tfs_ctx_t* ctx, uint32_t* msg_index, uint64_t* byte_offset) {
  if (!ctx || !ctx->err_has_location) return TFS_ERR_INVALID_ARG;
  if (msg_index) *msg_index = ctx->err_msg_index;
  if (byte_offset) *byte_offset = ctx->err_byte_offset;
  return TFS_OK;
}

tfs_status_t tfs_ctx_set_stream(tfs_ctx_t* ctx, void* cuda_stream) {
  if (!ctx) return TFS_ERR_INVALID_ARG;
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  return TFS_OK;
}

tfs_status_t tfs_sync(tfs_ctx_t* ctx) {
  if (!ctx) return TFS_ERR_INVALID_ARG;
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return TFS_OK;
}

uint64_t tfs_launch_count(const tfs_ctx_t* ctx) { return ctx ? ctx->launches : 0; }

tfs_status_t tfs_host_alloc(tfs_ctx_t* ctx, uint64_t bytes, void** out) {
  if (!ctx || !out) return TFS_ERR_INVALID_ARG;
  TFS_CUDA(ctx, cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
  return TFS_OK;
}
tfs_status_t tfs_host_free(tfs_ctx_t* ctx, void* p) {
  if (!ctx) return TFS_ERR_INVALID_ARG;
  if (p) TFS_CUDA(ctx, cudaFreeHost(p));
  return TFS_OK;
}
// ---- pinned staging pool -------------------------------------------------------------------------------------
// slab id: bit 63 = one-off allocation (index in the low bits is its serial), else (class << 48) | (offset / 4096)
namespace {
constexpr uint64_t kPoolGrain = 4096;
constexpr uint64_t kPoolExtraBit = 1ull << 63;
inline uint32_t pool_class_of(uint64_t bytes) {  // smallest c with (4096 << c) >= bytes
  uint32_t c = 0;
  while ((kPoolGrain << c) < bytes) ++c;
  return c;
}
}  // namespace

tfs_status_t tfs_stage_acquire(tfs_ctx_t* ctx, uint64_t bytes, uint32_t align_skew, tfs_slab_t* out) {
  if (!ctx || !out) return TFS_ERR_INVALID_ARG;
  const uint64_t skew = align_skew & 15u;
  const uint64_t need = bytes + skew;  // blocks start on 4 KiB boundaries: the skew is applied inside the block
  const uint32_t cls = pool_class_of(need ? need : 1);
  const uint64_t block = kPoolGrain << cls;
  std::lock_guard<std::mutex> lock(ctx->pool_mutex);
  if (ctx->pool_free.size() <= cls) ctx->pool_free.resize(cls + 1);
  uint64_t off = ~0ull;
  if (!ctx->pool_free[cls].empty()) {
    off = ctx->pool_free[cls].back();
    ctx->pool_free[cls].pop_back();
  } else if (ctx->h_slab && ctx->pool_bump + block <= ctx->h_slab_cap) {
    off = ctx->pool_bump;
    ctx->pool_bump += block;
  }
  if (off != ~0ull) {
    out->ptr = static_cast<uint8_t*>(ctx->h_slab) + off + skew;
    out->capacity = block - skew;
    out->id = (static_cast<uint64_t>(cls) << 48) | (off / kPoolGrain);
    ctx->pool_in_use += block;
    return TFS_OK;
  }
  // the pool is exhausted (or was never sized): a one-off pinned allocation, freed again on release
  void* p = nullptr;
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  TFS_CUDA(ctx, cudaHostAlloc(&p, need + 16, cudaHostAllocDefault));
  const uint64_t id = kPoolExtraBit | ctx->pool_next_id++;
  ctx->pool_extra.emplace_back(p, id);
  ctx->pool_fallbacks++;
  out->ptr = static_cast<uint8_t*>(p) + skew;  // (cudaHostAlloc returns page-aligned memory)
  out->capacity = need + 16 - skew;
  out->id = id;
  return TFS_OK;
}

tfs_status_t tfs_stage_release(tfs_ctx_t* ctx, const tfs_slab_t* slab) {
  if (!ctx || !slab) return TFS_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(ctx->pool_mutex);
  if (slab->id & kPoolExtraBit) {
    for (size_t i = 0; i < ctx->pool_extra.size(); ++i)
      if (ctx->pool_extra[i].second == slab->id) {
        void* p = ctx->pool_extra[i].first;
        ctx->pool_extra.erase(ctx->pool_extra.begin() + static_cast<long>(i));
        TFS_CUDA(ctx, cudaFreeHost(p));
        return TFS_OK;
      }
    return set_err(ctx, TFS_ERR_INVALID_ARG, "tfs_stage_release: unknown slab");
  }
  const uint32_t cls = static_cast<uint32_t>(slab->id >> 48);
  const uint64_t off = (slab->id & ((1ull << 48) - 1)) * kPoolGrain;
  if (cls >= 40 || off + (kPoolGrain << cls) > ctx->pool_bump) return set_err(ctx, TFS_ERR_INVALID_ARG, "tfs_stage_release: not a slab of this pool");
  if (ctx->pool_free.size() <= cls) ctx->pool_free.resize(cls + 1);
  ctx->pool_free[cls].push_back(off);
  ctx->pool_in_use -= kPoolGrain << cls;
  return TFS_OK;
}

tfs_status_t tfs_stage_stats(tfs_ctx_t* ctx, uint64_t* in_use, uint64_t* pool_bytes, uint64_t* fallbacks) {
  if (!ctx) return TFS_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(ctx->pool_mutex);
  if (in_use) *in_use = ctx->pool_in_use;
  if (pool_bytes) *pool_bytes = ctx->h_slab_cap;
  if (fallbacks) *fallbacks = ctx->pool_fallbacks;
  return TFS_OK;
}

tfs_status_t tfs_dev_alloc(tfs_ctx_t* ctx, uint64_t bytes, void** out) {
  if (!ctx || !out) return TFS_ERR_INVALID_ARG;
  TFS_CUDA(ctx, cudaMalloc(out, bytes ? bytes : 1));
  return TFS_OK;
}
tfs_status_t tfs_dev_free(tfs_ctx_t* ctx, void* p) {
  if (!ctx) return TFS_ERR_INVALID_ARG;
  if (p) {
    TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    TFS_CUDA(ctx, cudaFree(p));
  }
  return TFS_OK;
}
tfs_status_t tfs_copy_h2d(tfs_ctx_t* ctx, void* dst_dev, const void* src_host, uint64_t bytes) {
  if (!ctx || (bytes && (!dst_dev || !src_host))) return TFS_ERR_INVALID_ARG;
  if (bytes) TFS_CUDA(ctx, cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, ctx->stream));
  return TFS_OK;
}
tfs_status_t tfs_copy_d2h(tfs_ctx_t* ctx, void* dst_host, const void* src_dev, uint64_t bytes) {
  if (!ctx || (bytes && (!dst_host || !src_dev))) return TFS_ERR_INVALID_ARG;
  if (bytes) TFS_CUDA(ctx, cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  return TFS_OK;
}

// Greedy largest-first balancing of independent messages over shards: no collective,
// every GPU later runs the same single-GPU pipeline on its shard.
tfs_status_t tfs_shard_plan(const uint64_t* msg_bytes, uint32_t n_msgs, uint32_t n_shards, uint32_t* shard_out) {
  if (n_shards == 0 || (n_msgs && (!msg_bytes || !shard_out))) return TFS_ERR_INVALID_ARG;
  std::vector<uint32_t> order(n_msgs);
  std::iota(order.begin(), order.end(), 0u);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return msg_bytes[a] > msg_bytes[b]; });
  std::vector<uint64_t> load(n_shards, 0);
  for (uint32_t i : order) {
    uint32_t best = 0;
    for (uint32_t s = 1; s < n_shards; ++s)
      if (load[s] < load[best]) best = s;
    shard_out[i] = best;
    load[best] += msg_bytes[i] + 1;  // +1: empty messages still spread round-robin
  }
  return TFS_OK;
}

int32_t tfs_slot_of_kind(int32_t kind) {
  static const int32_t kinds[TFS_N_SLOTS] = {4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 16, 17, 14, 15};
  for (int32_t i = 0; i < TFS_N_SLOTS; ++i)
    if (kinds[i] == kind) return i;
  return -1;
}
int32_t tfs_kind_of_slot(int32_t slot) {
  static const int32_t kinds[TFS_N_SLOTS] = {4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 16, 17, 14, 15};
  return slot >= 0 && slot < TFS_N_SLOTS ? kinds[slot] : 0;
}
uint32_t tfs_type_size(int32_t t) {
  switch (t) {
    case TFS_T_U8: case TFS_T_I8: case TFS_T_BOOL: return 1;
    case TFS_T_U16: case TFS_T_I16: case TFS_T_F16: case TFS_T_BF16: return 2;
    case TFS_T_I32: case TFS_T_U32: case TFS_T_F32: return 4;
    case TFS_T_I64: case TFS_T_U64: case TFS_T_F64: return 8;
    default: return 0;
  }
}

}  // extern "C"
