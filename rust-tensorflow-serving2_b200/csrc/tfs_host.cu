// tfs_host.cu — the host-memory entry points: what the Rust shim calls from inside tonic's body stream,
// where the message payloads (Vec<f32>, ...) and the BytesMut both live in host memory.  Host<->device traffic
// is part of these calls (the `e2e` number of bench.py).
//
// Ways across PCIe, chosen per buffer:
//   * PINNED host memory (tfs_host_alloc / cudaHostAlloc / cudaHostRegister — the shim's staging pool) is
//     addressable by the GPU.  The kernels always STORE their output straight into a pinned destination.  A pinned
//     source is read in place when it is small; above 8 MB it is brought to HBM by the copy engine in 8 / 16 MB chunks on
//     a second stream while the copy kernel follows chunk by chunk (one launch per chunk behind the chunk's event,
//     limited to the tiles that can end inside it).  Measured on this B200 box (scripts/copy_sweep.cu host), summed
//     over both directions: copy engine in + kernel stores out 92 GB/s, copy engine both ways 95, kernels reading AND
//     writing host memory 78, kernel loads in + copy engine out 63-69 (SM loads from host memory are the weak leg).
//     Arrays that are varint-coded or ragged are read more than once, so they always go to HBM first.
//   * pageable host memory (what prost's Vecs and tonic's BytesMut are) cannot be addressed by the GPU.  Buffers of
//     1 MB and more cross a pinned ring of four 16 MB slots: helper threads copy pageable <-> pinned in parallel (with
//     non-temporal stores) while the copy engine drains / fills the other slots.  Measured on this box
//     (scripts/memcpy_sweep.py, scripts/pageable_e2e.py): the driver's own pageable cudaMemcpy 11 GB/s in, 21 out, pinned
//     DMA 55; C2 requests pageable -> pageable, encode or decode: 6.9 GB/s of wire through the driver's path, 16 / 20 / 22
//     on 4 / 6 / 8 staging threads (plain memcpy instead of streaming stores: 13 / 18 / 14).  Smaller buffers go through
//     cudaMemcpyAsync as they are.
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <system_error>
#include <thread>

#include "tfs_internal.h"

struct tfs_decoded;

namespace tfs {
tfs_status_t encode_device(tfs_ctx* ctx, const tfs_batch_t* batch, void* dst_dev, uint64_t dst_capacity, uint32_t flags,
                           uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out, bool size_only);
tfs_status_t decode_scan_device(tfs_ctx* ctx, const void* wire_dev, uint64_t wire_bytes, const uint64_t* msg_off,
                                const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, void* pool_dev,
                                uint64_t pool_capacity, tfs_decoded** out, const void* walk_wire_dev = nullptr);
tfs_status_t decode_device(tfs_ctx* ctx, tfs_decoded* dec, void* pool_dev, uint64_t pool_bytes);
uint64_t decoded_pool_bytes(const tfs_decoded* dec);
void decoded_mark_host_staged(tfs_decoded* dec, uint64_t stage_generation);
bool decoded_stage_is_current(const tfs_decoded* dec, uint64_t stage_generation);

namespace {

// Device-usable address of a host pointer when the memory is pinned / registered, nullptr otherwise.
void* device_alias(const void* p) {
  if (!p) return nullptr;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();  // unregistered host memory reports an error on some drivers: not pinned
    return nullptr;
  }
  if ((a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged) && a.devicePointer) return a.devicePointer;
  return nullptr;
}

bool is_varint_kind(int32_t kind) {
  return kind == TFS_KIND_INT_VAL || kind == TFS_KIND_INT64_VAL || kind == TFS_KIND_HALF_VAL || kind == TFS_KIND_UINT32_VAL ||
         kind == TFS_KIND_UINT64_VAL;
}

constexpr uint64_t kPipeChunk = 8ull << 20;    // bytes per copy-engine chunk ...
constexpr uint64_t kPipeChunkBig = 16ull << 20;  // ... of a transfer of 32 MB and more (measured on 32 x 38.5 MB: 4 MB chunks 33.1 ms per
                                                //     encode, 8 MB 29.2, 16 MB 28.5, 32 MB 28.4: fewer launches behind fewer events)
constexpr uint64_t kPipeMinBytes = 8ull << 20;  // smaller transfers are not worth the extra launches

// Queue `bytes` from pinned host memory to `dst_dev` on the copy stream, in chunks; every chunk gets an event and an
// entry in ctx->pipe, which the next launch_copy turns into one gated kernel launch per chunk.  (Measured on this box,
// scripts/copy_sweep.cu host: copy engine in + kernel stores out reach 92 GB/s summed, kernels reading AND writing host
// memory 78, kernel loads in + copy engine out only 63-69.)
tfs_status_t pipelined_h2d(tfs_ctx* ctx, uint8_t* dst_dev, const uint8_t* src_pinned, uint64_t bytes) {
  if (!ctx->copy_stream) TFS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  auto event_at = [&](size_t i, cudaEvent_t* e) -> tfs_status_t {
    while (ctx->pipe_events.size() <= i) {
      cudaEvent_t ne;
      TFS_CUDA(ctx, cudaEventCreateWithFlags(&ne, cudaEventDisableTiming));
      ctx->pipe_events.push_back(ne);
    }
    *e = ctx->pipe_events[i];
    return TFS_OK;
  };
  if (ctx->pipe.empty()) {  // the staging buffer may still be read by kernels of the previous call
    cudaEvent_t fence;
    TFS_TRY(event_at(0, &fence));
    TFS_CUDA(ctx, cudaEventRecord(fence, ctx->stream));
    TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, fence, 0));
  }
  const uint64_t chunk = bytes >= (32ull << 20) ? kPipeChunkBig : kPipeChunk;
  for (uint64_t off = 0; off < bytes; off += chunk) {
    const uint64_t n = bytes - off < chunk ? bytes - off : chunk;
    TFS_CUDA(ctx, cudaMemcpyAsync(dst_dev + off, src_pinned + off, n, cudaMemcpyHostToDevice, ctx->copy_stream));
    cudaEvent_t e;
    TFS_TRY(event_at(ctx->pipe.size() + 1, &e));  // slot 0 is the fence
    TFS_CUDA(ctx, cudaEventRecord(e, ctx->copy_stream));
    ctx->pipe.push_back({e, reinterpret_cast<uintptr_t>(dst_dev) + off + n});
  }
  return TFS_OK;
}

// whatever happened, the next call must not inherit chunk events; and nothing may still be in flight on the copy stream
void pipe_reset(tfs_ctx* ctx) {
  if (!ctx->pipe.empty() && ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
  ctx->pipe.clear();
}

// ---- pageable buffers: parallel memcpy through a pinned ring ----------------------------------------------------------
// A part of a staging copy: megabytes that are written once and next read by the copy engine (or by the caller, later).
// Non-temporal stores: the destination lines are not read first and do not push the source out of the cache.
void stream_copy(uint8_t* dst, const uint8_t* src, size_t n) {
#if defined(__SSE2__)
  if (n >= 4096) {
    const size_t head = (16 - (reinterpret_cast<uintptr_t>(dst) & 15u)) & 15u;
    if (head) std::memcpy(dst, src, head);
    dst += head, src += head, n -= head;
    const size_t blocks = n / 64;
    for (size_t i = 0; i < blocks; ++i) {
      const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src) + 0), b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src) + 1),
                    c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src) + 2), d = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src) + 3);
      _mm_stream_si128(reinterpret_cast<__m128i*>(dst) + 0, a);
      _mm_stream_si128(reinterpret_cast<__m128i*>(dst) + 1, b);
      _mm_stream_si128(reinterpret_cast<__m128i*>(dst) + 2, c);
      _mm_stream_si128(reinterpret_cast<__m128i*>(dst) + 3, d);
      src += 64, dst += 64;
    }
    _mm_sfence();
    n -= blocks * 64;
  }
#endif
  if (n) std::memcpy(dst, src, n);
}

class StageCrew {
 public:
  explicit StageCrew(unsigned helpers) {
    threads_.reserve(helpers);
    for (unsigned i = 0; i < helpers; ++i) {
      try {
        threads_.emplace_back([this, i] { run(i + 1); });
      } catch (const std::system_error&) {  // no more threads to be had: work with the ones that started
        break;
      }
    }
    parts_ = static_cast<unsigned>(threads_.size()) + 1;
  }
  ~StageCrew() {
    {
      std::lock_guard<std::mutex> lock(m_);
      stop_ = true;
    }
    work_.notify_all();
    for (std::thread& t : threads_) t.join();
  }
  // `bytes` from src to dst, split over the helpers and the calling thread; returns when every part is done
  void copy(uint8_t* dst, const uint8_t* src, size_t bytes) {
    const unsigned parts = parts_;
    if (bytes < (256u << 10) || parts == 1) {
      std::memcpy(dst, src, bytes);
      return;
    }
    {
      std::lock_guard<std::mutex> lock(m_);
      dst_ = dst, src_ = src, bytes_ = bytes;
      left_ = parts - 1;
      ++generation_;
    }
    work_.notify_all();
    part(0, parts);
    std::unique_lock<std::mutex> lock(m_);
    done_.wait(lock, [this] { return left_ == 0; });
  }

 private:
  void part(unsigned id, unsigned parts) {
    const size_t lo = (bytes_ * id / parts) & ~size_t(63), hi = id + 1 == parts ? bytes_ : (bytes_ * (id + 1) / parts) & ~size_t(63);
    if (hi > lo) stream_copy(dst_ + lo, src_ + lo, hi - lo);
  }
  void run(unsigned id) {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> lock(m_);
        work_.wait(lock, [&] { return stop_ || generation_ != seen; });
        if (stop_) return;
        seen = generation_;
      }
      part(id, parts_);
      bool last;
      {
        std::lock_guard<std::mutex> lock(m_);
        last = --left_ == 0;
      }
      if (last) done_.notify_one();
    }
  }
  std::vector<std::thread> threads_;
  unsigned parts_ = 1;  // helpers + the calling thread (fixed once the constructor is done)
  std::mutex m_;
  std::condition_variable work_, done_;
  uint64_t generation_ = 0;
  unsigned left_ = 0;
  bool stop_ = false;
  uint8_t* dst_ = nullptr;
  const uint8_t* src_ = nullptr;
  size_t bytes_ = 0;
};

constexpr uint64_t kRingSlot = 16ull << 20;     // one slot is being copied by the crew while the copy engine works on the others
constexpr uint32_t kRingSlots = 4;              // (tfs_ctx::ring_ev / ring_busy have this many entries)
constexpr uint64_t kStageMinBytes = 1ull << 20;  // smaller pageable buffers are left to cudaMemcpyAsync

tfs_status_t stage_ready(tfs_ctx* ctx, StageCrew** crew) {
  if (!ctx->copy_stream) TFS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  if (!ctx->h_ring) {
    TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    TFS_CUDA(ctx, cudaHostAlloc(&ctx->h_ring, kRingSlots * kRingSlot, cudaHostAllocDefault));
    for (cudaEvent_t& e : ctx->ring_ev) TFS_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  if (!ctx->stage_crew) {
    unsigned helpers = std::min(7u, std::max(2u, std::thread::hardware_concurrency() / 2) - 1u);  // 8 threads on this box's 16 cores
    if (const char* e = std::getenv("TFS_STAGE_THREADS")) helpers = static_cast<unsigned>(std::max(1, std::atoi(e))) - 1u;
    ctx->stage_crew = std::make_shared<StageCrew>(helpers);
  }
  *crew = static_cast<StageCrew*>(ctx->stage_crew.get());
  return TFS_OK;
}

cudaEvent_t* pipe_event_slot(tfs_ctx* ctx, size_t i) {
  while (ctx->pipe_events.size() <= i) {
    cudaEvent_t ne;
    if (cudaEventCreateWithFlags(&ne, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    ctx->pipe_events.push_back(ne);
  }
  return &ctx->pipe_events[i];
}

// Pageable host memory -> `dst_dev`.  With `pieces` every 8 MB chunk gets an entry in ctx->pipe (the copy kernels follow
// chunk by chunk, as behind pipelined_h2d); without, ctx->stream waits for the last chunk.
tfs_status_t staged_h2d(tfs_ctx* ctx, uint8_t* dst_dev, const uint8_t* src, uint64_t bytes, bool pieces) {
  StageCrew* crew = nullptr;
  TFS_TRY(stage_ready(ctx, &crew));
  {  // the device staging buffer may still be read by kernels of the previous call
    cudaEvent_t* fence = pipe_event_slot(ctx, 0);
    if (!fence) return set_err(ctx, TFS_ERR_CUDA, "cudaEventCreate");
    TFS_CUDA(ctx, cudaEventRecord(*fence, ctx->stream));
    TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, *fence, 0));
  }
  uint8_t* ring = static_cast<uint8_t*>(ctx->h_ring);
  uint32_t slot = 0;
  for (uint64_t off = 0; off < bytes; off += kRingSlot) {
    slot = ctx->ring_next++ % kRingSlots;
    if (ctx->ring_busy[slot]) TFS_CUDA(ctx, cudaEventSynchronize(ctx->ring_ev[slot]));  // its last transfer has drained
    const uint64_t n = std::min<uint64_t>(kRingSlot, bytes - off);
    crew->copy(ring + slot * kRingSlot, src + off, n);
    for (uint64_t o2 = 0; o2 < n; o2 += kPipeChunk) {
      const uint64_t m = std::min<uint64_t>(kPipeChunk, n - o2);
      TFS_CUDA(ctx, cudaMemcpyAsync(dst_dev + off + o2, ring + slot * kRingSlot + o2, m, cudaMemcpyHostToDevice, ctx->copy_stream));
      if (pieces) {
        cudaEvent_t* e = pipe_event_slot(ctx, ctx->pipe.size() + 1);  // slot 0 is the fence
        if (!e) return set_err(ctx, TFS_ERR_CUDA, "cudaEventCreate");
        TFS_CUDA(ctx, cudaEventRecord(*e, ctx->copy_stream));
        ctx->pipe.push_back({*e, reinterpret_cast<uintptr_t>(dst_dev) + off + o2 + m});
      }
    }
    TFS_CUDA(ctx, cudaEventRecord(ctx->ring_ev[slot], ctx->copy_stream));
    ctx->ring_busy[slot] = true;
  }
  if (!pieces && bytes) TFS_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ring_ev[slot], 0));
  return TFS_OK;
}

// `src_dev` (written by work already queued on ctx->stream) -> pageable host memory; returns with the bytes in place
tfs_status_t staged_d2h(tfs_ctx* ctx, uint8_t* dst, const uint8_t* src_dev, uint64_t bytes) {
  StageCrew* crew = nullptr;
  TFS_TRY(stage_ready(ctx, &crew));
  uint8_t* ring = static_cast<uint8_t*>(ctx->h_ring);
  for (uint32_t s = 0; s < kRingSlots; ++s)  // (transfers of the upload phase out of a slot have been consumed by the kernels; be exact)
    if (ctx->ring_busy[s]) {
      TFS_CUDA(ctx, cudaEventSynchronize(ctx->ring_ev[s]));
      ctx->ring_busy[s] = false;
    }
  const uint64_t n_slots = (bytes + kRingSlot - 1) / kRingSlot;
  auto issue = [&](uint64_t k) -> tfs_status_t {
    const uint64_t off = k * kRingSlot, n = std::min<uint64_t>(kRingSlot, bytes - off);
    TFS_CUDA(ctx, cudaMemcpyAsync(ring + (k % kRingSlots) * kRingSlot, src_dev + off, n, cudaMemcpyDeviceToHost, ctx->stream));
    TFS_CUDA(ctx, cudaEventRecord(ctx->ring_ev[k % kRingSlots], ctx->stream));
    return TFS_OK;
  };
  for (uint64_t k = 0; k < n_slots && k < kRingSlots; ++k) TFS_TRY(issue(k));
  for (uint64_t k = 0; k < n_slots; ++k) {
    const uint64_t off = k * kRingSlot, n = std::min<uint64_t>(kRingSlot, bytes - off);
    TFS_CUDA(ctx, cudaEventSynchronize(ctx->ring_ev[k % kRingSlots]));
    crew->copy(dst + off, ring + (k % kRingSlots) * kRingSlot, n);
    if (k + kRingSlots < n_slots) TFS_TRY(issue(k + kRingSlots));
  }
  return TFS_OK;
}

// wire bytes: pinned -> read in place; pageable -> copied to the device slab
tfs_status_t wire_for_device(tfs_ctx* ctx, const void* wire_host, uint64_t wire_bytes, const void** out) {
  void* alias = device_alias(wire_host);
  if (alias || wire_bytes == 0) {
    *out = alias;
    return TFS_OK;
  }
  TFS_TRY(dev_reserve(ctx, &ctx->d_stage_in, wire_bytes + 256));
  if (wire_bytes >= kStageMinBytes)
    TFS_TRY(staged_h2d(ctx, static_cast<uint8_t*>(ctx->d_stage_in.p), static_cast<const uint8_t*>(wire_host), wire_bytes, false));
  else
    TFS_CUDA(ctx, cudaMemcpyAsync(ctx->d_stage_in.p, wire_host, wire_bytes, cudaMemcpyHostToDevice, ctx->stream));
  *out = ctx->d_stage_in.p;
  return TFS_OK;
}

}  // namespace
}  // namespace tfs

using namespace tfs;

extern "C" {

tfs_status_t tfs_encode_host(tfs_ctx_t* ctx, const tfs_batch_t* batch, void* dst_host, uint64_t dst_capacity, uint32_t flags,
                             uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out) {
  if (!ctx || !batch || (!dst_host && dst_capacity)) return TFS_ERR_INVALID_ARG;
  if ((batch->n_fields && !batch->fields)) return TFS_ERR_INVALID_ARG;
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->stage_generation++;  // the staging buffers are reused: a pending tfs_decode_host handle becomes stale
  HostTrace tr(ctx, "encode_host");
  // 1. payload arrays: pinned bulk arrays are read in place by the kernels; the rest is copied to a device slab
  std::vector<tfs_field_desc_t> fields(batch->fields, batch->fields + batch->n_fields);
  std::vector<uint64_t> at(batch->n_fields, ~0ull);
  std::vector<uint8_t> piped(batch->n_fields, 0);
  uint64_t in_bytes = 0;
  pipe_reset(ctx);
  const bool dst_pinned = device_alias(dst_host) != nullptr;
  auto src_bytes = [](const tfs_field_desc_t& f) -> uint64_t {
    if (f.src_type == TFS_T_RAGGED) return f.n_elems ? 8ull * (f.n_elems + 1) + f.aux : 0ull;  // offsets, then the bytes
    return f.n_elems * tfs_type_size(f.src_type);
  };
  for (uint32_t i = 0; i < batch->n_fields; ++i) {
    const uint32_t w = fields[i].src_type == TFS_T_RAGGED ? 1u : tfs_type_size(fields[i].src_type);
    const uint64_t b = src_bytes(fields[i]);
    if (b && (!fields[i].data || w == 0)) return set_err(ctx, TFS_ERR_INVALID_ARG, "payload field without data or type");
    if (!b) continue;
    // read more than once by the kernels (sizing + writing): varint-coded arrays and ragged blocks go to HBM first
    void* alias = (is_varint_kind(fields[i].kind) || fields[i].src_type == TFS_T_RAGGED) ? nullptr : device_alias(fields[i].data);
    // large pinned bulk arrays headed for a pinned destination: copy engine in (chunked), kernel stores out, overlapped
    if (alias && dst_pinned && b >= kPipeMinBytes) {
      piped[i] = 1;
      alias = nullptr;
    }
    if (alias) {
      fields[i].data = alias;
    } else {
      at[i] = in_bytes;
      in_bytes = align_up(in_bytes + b, 256);
    }
  }
  if (in_bytes) TFS_TRY(dev_reserve(ctx, &ctx->d_stage_in, in_bytes + 256));
  uint8_t* din = static_cast<uint8_t*>(ctx->d_stage_in.p);
  for (uint32_t i = 0; i < batch->n_fields; ++i) {
    if (at[i] == ~0ull) continue;
    const uint64_t b = src_bytes(fields[i]);
    if (piped[i]) {
      tfs_status_t st = pipelined_h2d(ctx, din + at[i], static_cast<const uint8_t*>(fields[i].data), b);
      if (st != TFS_OK) {
        pipe_reset(ctx);
        return st;
      }
    } else if (b >= kStageMinBytes && !device_alias(fields[i].data)) {
      // pageable (a prost Vec): through the pinned ring on the helper threads.  A copy-class array headed for a pinned
      // destination lets the copy kernel follow chunk by chunk; arrays the kernels read more than once just wait.
      const bool pieces = dst_pinned && !is_varint_kind(fields[i].kind) && fields[i].src_type != TFS_T_RAGGED;
      tfs_status_t st = staged_h2d(ctx, din + at[i], static_cast<const uint8_t*>(fields[i].data), b, pieces);
      if (st != TFS_OK) {
        pipe_reset(ctx);
        return st;
      }
    } else {
      TFS_CUDA(ctx, cudaMemcpyAsync(din + at[i], fields[i].data, b, cudaMemcpyHostToDevice, ctx->stream));
    }
    fields[i].data = din + at[i];
  }
  tr.mark("sources");
  flags &= ~static_cast<uint32_t>(TFS_F_ASYNC);  // a host destination is complete when the call returns
  tfs_batch_t dev_batch = *batch;
  dev_batch.fields = fields.data();
  // 2. encode; the wire bytes go straight into a pinned destination, else through a device slab + D2H
  uint64_t total = 0;
  void* dst_alias = device_alias(dst_host);
  if (dst_alias) {
    const tfs_status_t st = encode_device(ctx, &dev_batch, dst_alias, dst_capacity, flags, msg_off_out, msg_len_out, &total, false);
    pipe_reset(ctx);
    if (st != TFS_OK) return st;
  } else {
    TFS_TRY(dev_reserve(ctx, &ctx->d_stage_out, dst_capacity + 256));
    const tfs_status_t st = encode_device(ctx, &dev_batch, ctx->d_stage_out.p, dst_capacity, flags, msg_off_out, msg_len_out, &total, false);
    pipe_reset(ctx);
    if (st != TFS_OK) return st;
    tr.mark("encode");
    if (total >= kStageMinBytes)
      TFS_TRY(staged_d2h(ctx, static_cast<uint8_t*>(dst_host), static_cast<const uint8_t*>(ctx->d_stage_out.p), total));
    else if (total)
      TFS_CUDA(ctx, cudaMemcpyAsync(dst_host, ctx->d_stage_out.p, total, cudaMemcpyDeviceToHost, ctx->stream));
    TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    tr.mark("wire_out");
  }
  if (total_out) *total_out = total;
  return TFS_OK;
}

tfs_status_t tfs_decode_host(tfs_ctx_t* ctx, const void* wire_host, uint64_t wire_bytes, const uint64_t* msg_off,
                             const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, tfs_decoded_t** out) {
  if (!ctx || !out || (n_msgs && (!msg_off || !msg_len)) || (wire_bytes && !wire_host)) return TFS_ERR_INVALID_ARG;
  *out = nullptr;
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint64_t gen = ++ctx->stage_generation;
  const void* wire_dev = nullptr;
  TFS_TRY(wire_for_device(ctx, wire_host, wire_bytes, &wire_dev));
  tfs_status_t st = decode_scan_device(ctx, wire_dev, wire_bytes, msg_off, msg_len, n_msgs, msg_kind, flags, nullptr, 0, out);
  if (*out) decoded_mark_host_staged(*out, gen);
  return st;
}

tfs_status_t tfs_decode_host_finish(tfs_ctx_t* ctx, tfs_decoded_t* dec, void* pool_host, uint64_t pool_bytes) {
  if (!ctx || !dec) return TFS_ERR_INVALID_ARG;
  if (!decoded_stage_is_current(dec, ctx->stage_generation))
    return set_err(ctx, TFS_ERR_INVALID_ARG, "stale handle: another *_host call reused the staging buffers since tfs_decode_host");
  const uint64_t need = decoded_pool_bytes(dec);
  if (pool_bytes < need || (need && !pool_host)) return set_err(ctx, TFS_ERR_NOMEM, "value pool is smaller than tfs_decoded_view().pool_bytes");
  void* alias = device_alias(pool_host);
  if (alias) {
    TFS_TRY(decode_device(ctx, dec, alias, need));
  } else {
    TFS_TRY(dev_reserve(ctx, &ctx->d_stage_out, need + 256));
    TFS_TRY(decode_device(ctx, dec, ctx->d_stage_out.p, need));
    if (need >= kStageMinBytes)
      TFS_TRY(staged_d2h(ctx, static_cast<uint8_t*>(pool_host), static_cast<const uint8_t*>(ctx->d_stage_out.p), need));
    else if (need)
      TFS_CUDA(ctx, cudaMemcpyAsync(pool_host, ctx->d_stage_out.p, need, cudaMemcpyDeviceToHost, ctx->stream));
  }
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return TFS_OK;
}

tfs_status_t tfs_decode_host_into(tfs_ctx_t* ctx, const void* wire_host, uint64_t wire_bytes, const uint64_t* msg_off,
                                  const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, void* pool_host,
                                  uint64_t pool_capacity, tfs_decoded_t** out) {
  if (!ctx || !out || !pool_host || (n_msgs && (!msg_off || !msg_len)) || (wire_bytes && !wire_host)) return TFS_ERR_INVALID_ARG;
  *out = nullptr;
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint64_t gen = ++ctx->stage_generation;
  pipe_reset(ctx);
  void* alias = device_alias(pool_host);
  void* wire_alias = device_alias(wire_host);
  tfs_status_t st;
  if (alias && wire_alias && wire_bytes >= kPipeMinBytes) {
    // both buffers pinned, a large batch: the walk reads the few header bytes straight from host memory while the copy
    // engine brings the wire to HBM in chunks; the copy kernel follows chunk by chunk and stores into the host pool
    TFS_TRY(dev_reserve(ctx, &ctx->d_stage_in, wire_bytes + 256));
    st = pipelined_h2d(ctx, static_cast<uint8_t*>(ctx->d_stage_in.p), static_cast<const uint8_t*>(wire_host), wire_bytes);
    if (st == TFS_OK)
      st = decode_scan_device(ctx, ctx->d_stage_in.p, wire_bytes, msg_off, msg_len, n_msgs, msg_kind, flags, alias, pool_capacity, out, wire_alias);
    if (*out) cudaStreamSynchronize(ctx->stream);
    pipe_reset(ctx);
    if (*out) decoded_mark_host_staged(*out, gen);
    return st;
  }
  const void* wire_dev = nullptr;
  TFS_TRY(wire_for_device(ctx, wire_host, wire_bytes, &wire_dev));
  if (alias) {
    st = decode_scan_device(ctx, wire_dev, wire_bytes, msg_off, msg_len, n_msgs, msg_kind, flags, alias, pool_capacity, out);
    // the value kernels may still be writing the host pool (the device-pool form is asynchronous): wait for them
    if (*out) TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    TFS_TRY(dev_reserve(ctx, &ctx->d_stage_out, pool_capacity + 256));
    st = decode_scan_device(ctx, wire_dev, wire_bytes, msg_off, msg_len, n_msgs, msg_kind, flags, ctx->d_stage_out.p, pool_capacity, out);
    if (*out && (st == TFS_OK || st == TFS_ERR_BATCH)) {
      const uint64_t need = decoded_pool_bytes(*out);
      if (need >= kStageMinBytes)
        TFS_TRY(staged_d2h(ctx, static_cast<uint8_t*>(pool_host), static_cast<const uint8_t*>(ctx->d_stage_out.p), need));
      else if (need)
        TFS_CUDA(ctx, cudaMemcpyAsync(pool_host, ctx->d_stage_out.p, need, cudaMemcpyDeviceToHost, ctx->stream));
      TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
  }
  if (*out) decoded_mark_host_staged(*out, gen);
  return st;
}

}  // extern "C"
