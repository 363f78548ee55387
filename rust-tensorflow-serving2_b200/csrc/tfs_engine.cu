// tfs_engine.cu — worker threads over one or several GPUs, asynchronous submission, request coalescing
// (host code only; the kernels are reached through the ordinary host entry points).
//
// What it stands in for on the reference's side: the client is `Clone` and its clones are used from concurrent tokio
// tasks (src/lib.rs:148-156, examples/async.rs:29-46); each task builds ONE request and sends it (src/lib.rs:229-265).
// A GPU wants batches and a tfs_ctx_t is neither re-entrant nor asynchronous, so:
//   * a lane = one tfs_ctx_t (own stream, own staging pool) + one host thread that runs the jobs queued for it with the
//     synchronous entry points; two lanes per GPU let the upload of one job run under the download of the previous one;
//   * lanes on different GPUs make one client process drive several GPUs; requests are independent, so there is no
//     collective and no peer access, every GPU reads and writes only its own memory and the pinned host buffers;
//   * the coalescer gathers the single-message calls of many threads into one batch per window and hands every caller
//     a pointer into the pinned slab the batch was encoded into (the transport sends from there: no copy in between).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <thread>

#include "tfs_internal.h"

struct tfs_decoded;

namespace {

struct Job {
  std::function<tfs_status_t(tfs_ctx*)> run;
  uint64_t bytes = 0;
  tfs_status_t status = TFS_OK;
  std::string detail;
  bool done = false;
};

struct Lane {
  tfs_ctx* ctx = nullptr;
  std::thread worker;
  std::mutex m;
  std::condition_variable cv;
  std::deque<std::shared_ptr<Job>> queue;
  std::atomic<uint64_t> queued_bytes{0};
  bool stop = false;
};

}  // namespace

struct tfs_engine {
  std::vector<std::unique_ptr<Lane>> lanes;
  std::mutex m;  // tickets
  std::condition_variable cv;
  std::map<uint64_t, std::shared_ptr<Job>> tickets;
  std::map<tfs_decoded*, uint32_t> decoded_owner;  // which lane made a decode handle
  uint64_t next_ticket = 1;
};

namespace {

void lane_main(tfs_engine* eng, Lane* lane) {
  cudaSetDevice(lane->ctx->device);
  for (;;) {
    std::shared_ptr<Job> job;
    {
      std::unique_lock<std::mutex> lock(lane->m);
      lane->cv.wait(lock, [&] { return lane->stop || !lane->queue.empty(); });
      if (lane->queue.empty()) return;  // stop requested and nothing left
      job = lane->queue.front();
      lane->queue.pop_front();
    }
    const tfs_status_t st = job->run(lane->ctx);
    lane->queued_bytes -= job->bytes;
    {
      std::lock_guard<std::mutex> lock(eng->m);
      job->status = st;
      if (st != TFS_OK) job->detail = lane->ctx->err;
      job->done = true;
    }
    eng->cv.notify_all();
  }
}

uint32_t pick_lane(tfs_engine* eng, int32_t lane) {
  if (lane >= 0) return static_cast<uint32_t>(lane) % static_cast<uint32_t>(eng->lanes.size());
  uint32_t best = 0;
  for (uint32_t i = 1; i < eng->lanes.size(); ++i)
    if (eng->lanes[i]->queued_bytes < eng->lanes[best]->queued_bytes) best = i;
  return best;
}

tfs_ticket_t submit(tfs_engine* eng, uint32_t lane_index, std::shared_ptr<Job> job) {
  Lane* lane = eng->lanes[lane_index].get();
  uint64_t t;
  {
    std::lock_guard<std::mutex> lock(eng->m);
    t = eng->next_ticket++;
    eng->tickets[t] = job;
  }
  lane->queued_bytes += job->bytes;
  {
    std::lock_guard<std::mutex> lock(lane->m);
    lane->queue.push_back(std::move(job));
  }
  lane->cv.notify_one();
  return t;
}

uint64_t batch_bytes(const tfs_batch_t* b) {
  uint64_t n = 0;
  for (uint32_t i = 0; i < b->n_fields; ++i) n += b->fields[i].n_elems * (tfs_type_size(b->fields[i].src_type) ? tfs_type_size(b->fields[i].src_type) : 1) + b->fields[i].aux;
  return n;
}

}  // namespace

extern "C" {

tfs_status_t tfs_shard_ranges(const uint64_t* msg_bytes, uint32_t n_msgs, uint32_t n_shards, uint32_t* first_out) {
  if (n_shards == 0 || !first_out || (n_msgs && !msg_bytes)) return TFS_ERR_INVALID_ARG;
  uint64_t total = 0;
  for (uint32_t i = 0; i < n_msgs; ++i) total += msg_bytes[i] + 1;  // +1: empty messages still spread
  uint32_t i = 0;
  uint64_t done = 0;
  for (uint32_t s = 0; s < n_shards; ++s) {
    first_out[s] = i;
    const uint64_t goal = total * (s + 1) / n_shards;  // bytes that should lie in front of run s + 1
    while (i < n_msgs && done + (msg_bytes[i] + 1) / 2 < goal) done += msg_bytes[i++] + 1;  // a message goes where most of it lies
  }
  first_out[n_shards] = n_msgs;
  // whatever is left belongs to the last run (its goal is the total: nothing is left unless rounding kept a message back)
  return TFS_OK;
}

tfs_status_t tfs_engine_create(const int32_t* devices, uint32_t n_devices, uint32_t lanes_per_device, uint64_t pinned_bytes_per_lane,
                               tfs_engine_t** out) {
  if (!out || !devices || n_devices == 0 || lanes_per_device == 0 || lanes_per_device > 8) return TFS_ERR_INVALID_ARG;
  *out = nullptr;
  std::unique_ptr<tfs_engine> eng(new tfs_engine());
  for (uint32_t d = 0; d < n_devices; ++d)
    for (uint32_t l = 0; l < lanes_per_device; ++l) {
      std::unique_ptr<Lane> lane(new Lane());
      const tfs_status_t st = tfs_ctx_create(devices[d], pinned_bytes_per_lane, &lane->ctx);
      if (st != TFS_OK) {
        for (auto& made : eng->lanes) tfs_ctx_destroy(made->ctx);
        return st;
      }
      eng->lanes.push_back(std::move(lane));
    }
  for (auto& lane : eng->lanes) lane->worker = std::thread(lane_main, eng.get(), lane.get());
  *out = eng.release();
  return TFS_OK;
}

tfs_status_t tfs_engine_destroy(tfs_engine_t* eng) {
  if (!eng) return TFS_OK;
  for (auto& lane : eng->lanes) {
    {
      std::lock_guard<std::mutex> lock(lane->m);
      lane->stop = true;
    }
    lane->cv.notify_all();
  }
  for (auto& lane : eng->lanes)
    if (lane->worker.joinable()) lane->worker.join();  // (queued jobs are finished first)
  for (auto& lane : eng->lanes) tfs_ctx_destroy(lane->ctx);
  delete eng;
  return TFS_OK;
}

uint32_t tfs_engine_lanes(const tfs_engine_t* eng) { return eng ? static_cast<uint32_t>(eng->lanes.size()) : 0u; }

tfs_ctx_t* tfs_engine_ctx(tfs_engine_t* eng, uint32_t lane) { return eng && lane < eng->lanes.size() ? eng->lanes[lane]->ctx : nullptr; }

tfs_status_t tfs_engine_encode_host(tfs_engine_t* eng, int32_t lane, const tfs_batch_t* batch, void* dst_host, uint64_t dst_capacity,
                                    uint32_t flags, uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out,
                                    tfs_ticket_t* ticket) {
  if (!eng || !batch || !ticket) return TFS_ERR_INVALID_ARG;
  auto job = std::make_shared<Job>();
  job->bytes = batch_bytes(batch);
  job->run = [=](tfs_ctx* ctx) { return tfs_encode_host(ctx, batch, dst_host, dst_capacity, flags, msg_off_out, msg_len_out, total_out); };
  *ticket = submit(eng, pick_lane(eng, lane), std::move(job));
  return TFS_OK;
}

tfs_status_t tfs_engine_decode_host_into(tfs_engine_t* eng, int32_t lane, const void* wire_host, uint64_t wire_bytes,
                                         const uint64_t* msg_off, const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind,
                                         uint32_t flags, void* pool_host, uint64_t pool_capacity, tfs_decoded_t** out,
                                         tfs_ticket_t* ticket) {
  if (!eng || !out || !ticket) return TFS_ERR_INVALID_ARG;
  *out = nullptr;
  const uint32_t li = pick_lane(eng, lane);
  auto job = std::make_shared<Job>();
  job->bytes = wire_bytes;
  job->run = [=](tfs_ctx* ctx) {
    const tfs_status_t st = tfs_decode_host_into(ctx, wire_host, wire_bytes, msg_off, msg_len, n_msgs, msg_kind, flags, pool_host, pool_capacity, out);
    if (*out) {
      std::lock_guard<std::mutex> lock(eng->m);
      eng->decoded_owner[*out] = li;
    }
    return st;
  };
  *ticket = submit(eng, li, std::move(job));
  return TFS_OK;
}

tfs_status_t tfs_engine_wait(tfs_engine_t* eng, tfs_ticket_t ticket, char* detail_out) {
  if (!eng) return TFS_ERR_INVALID_ARG;
  std::unique_lock<std::mutex> lock(eng->m);
  auto it = eng->tickets.find(ticket);
  if (it == eng->tickets.end()) return TFS_ERR_INVALID_ARG;  // unknown, or waited for already
  std::shared_ptr<Job> job = it->second;
  eng->cv.wait(lock, [&] { return job->done; });
  eng->tickets.erase(ticket);
  if (detail_out) {
    std::strncpy(detail_out, job->detail.c_str(), 255);
    detail_out[255] = 0;
  }
  return job->status;
}

tfs_status_t tfs_engine_decoded_free(tfs_engine_t* eng, tfs_decoded_t* dec) {
  if (!eng) return TFS_ERR_INVALID_ARG;
  if (!dec) return TFS_OK;
  uint32_t li = 0;
  {
    std::lock_guard<std::mutex> lock(eng->m);
    auto it = eng->decoded_owner.find(dec);
    if (it == eng->decoded_owner.end()) return TFS_ERR_INVALID_ARG;
    li = it->second;
    eng->decoded_owner.erase(it);
  }
  return tfs_decoded_free(eng->lanes[li]->ctx, dec);  // (host memory only: safe while the lane is busy)
}

tfs_status_t tfs_engine_decode_host_sharded(tfs_engine_t* eng, const void* wire_host, uint64_t wire_bytes, const uint64_t* msg_off,
                                            const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags,
                                            void* const* pool_host, const uint64_t* pool_capacity, tfs_decoded_t** out,
                                            uint32_t* first_msg_out) {
  if (!eng || !pool_host || !pool_capacity || !out || !first_msg_out || (n_msgs && (!msg_off || !msg_len))) return TFS_ERR_INVALID_ARG;
  const uint32_t n = static_cast<uint32_t>(eng->lanes.size());
  std::vector<uint32_t> first(n + 1);
  tfs_status_t st = tfs_shard_ranges(msg_len, n_msgs, n, first.data());
  if (st != TFS_OK) return st;
  // every run gets the sub-range of the wire it covers, its offsets rebased to that sub-range
  std::vector<std::vector<uint64_t>> offs(n);
  std::vector<tfs_ticket_t> tickets(n, 0);
  for (uint32_t r = 0; r < n; ++r) {
    first_msg_out[r] = first[r];
    out[r] = nullptr;
    const uint32_t a = first[r], b = first[r + 1];
    if (a == b) continue;
    uint64_t lo = ~0ull, hi = 0;
    for (uint32_t i = a; i < b; ++i) {
      lo = std::min<uint64_t>(lo, msg_off[i]);
      hi = std::max<uint64_t>(hi, msg_off[i] + msg_len[i]);
    }
    if (hi > wire_bytes) return TFS_ERR_INVALID_ARG;
    lo &= ~static_cast<uint64_t>(15);  // keep the byte phase of every message: pinned wires are read where they lie
    offs[r].resize(b - a);
    for (uint32_t i = a; i < b; ++i) offs[r][i - a] = msg_off[i] - lo;
    st = tfs_engine_decode_host_into(eng, static_cast<int32_t>(r), static_cast<const uint8_t*>(wire_host) + lo, hi - lo, offs[r].data(), msg_len + a,
                                     b - a, msg_kind, flags, pool_host[r], pool_capacity[r], &out[r], &tickets[r]);
    if (st != TFS_OK) break;
  }
  tfs_status_t first_bad = st;
  for (uint32_t r = 0; r < n; ++r) {
    if (!tickets[r]) continue;
    const tfs_status_t s2 = tfs_engine_wait(eng, tickets[r], nullptr);
    if (first_bad == TFS_OK && s2 != TFS_OK) first_bad = s2;
  }
  return first_bad;
}

}  // extern "C"

// =====================================================================================================
// coalescer
// =====================================================================================================
namespace {

struct Pending {  // a batch that is still open, or being encoded
  std::vector<tfs_msg_desc_t> msgs;
  std::vector<tfs_tensor_desc_t> tensors;
  std::vector<tfs_field_desc_t> fields;
  std::vector<tfs_str_t> filters;
  uint32_t flags = 0;
  std::chrono::steady_clock::time_point opened;
  bool closed = false, done = false;
  tfs_status_t status = TFS_OK;
  std::vector<uint64_t> off, len;
  tfs_slab_t slab{};
  uint32_t lane = 0;
  uint32_t refs = 0;  // callers that have not released their bytes yet
  uint64_t lease = 0;
};

}  // namespace

struct tfs_coalescer {
  tfs_engine* eng = nullptr;
  uint32_t window_us = 200, max_msgs = 64;
  std::mutex m;
  std::condition_variable cv;
  std::shared_ptr<Pending> open;
  std::map<uint64_t, std::shared_ptr<Pending>> leased;
  uint64_t next_lease = 1, batches = 0, messages = 0;
};

namespace {

// encode a closed batch (called WITHOUT the coalescer's lock by the caller that closed it)
void run_batch(tfs_coalescer* co, const std::shared_ptr<Pending>& p) {
  tfs_batch_t b{};
  b.msgs = p->msgs.data();
  b.tensors = p->tensors.data();
  b.fields = p->fields.data();
  b.filters = p->filters.data();
  b.n_msgs = static_cast<uint32_t>(p->msgs.size());
  b.n_tensors = static_cast<uint32_t>(p->tensors.size());
  b.n_fields = static_cast<uint32_t>(p->fields.size());
  b.n_filters = static_cast<uint32_t>(p->filters.size());
  p->off.assign(b.n_msgs, 0);
  p->len.assign(b.n_msgs, 0);
  tfs_status_t st = TFS_OK;
  const uint64_t cap = tfs_encode_bound(&b, p->flags);
  p->lane = pick_lane(co->eng, -1);
  tfs_ctx* pool_ctx = co->eng->lanes[p->lane]->ctx;
  if (cap == 0) st = TFS_ERR_INVALID_ARG;
  if (st == TFS_OK) st = tfs_stage_acquire(pool_ctx, cap, 0, &p->slab);
  if (st == TFS_OK) {
    uint64_t total = 0;
    tfs_ticket_t t = 0;
    st = tfs_engine_encode_host(co->eng, static_cast<int32_t>(p->lane), &b, p->slab.ptr, p->slab.capacity, p->flags, p->off.data(), p->len.data(), &total, &t);
    if (st == TFS_OK) st = tfs_engine_wait(co->eng, t, nullptr);
    if (st != TFS_OK) {
      tfs_stage_release(pool_ctx, &p->slab);
      p->slab = tfs_slab_t{};
    }
  }
  {
    std::lock_guard<std::mutex> lock(co->m);
    p->status = st;
    p->done = true;
    co->batches++;
    co->messages += b.n_msgs;
    if (st == TFS_OK) {
      p->lease = co->next_lease++;
      co->leased[p->lease] = p;
    }
  }
  co->cv.notify_all();
}

}  // namespace

extern "C" {

tfs_status_t tfs_coalescer_create(tfs_engine_t* eng, uint32_t window_us, uint32_t max_msgs, tfs_coalescer_t** out) {
  if (!eng || !out || max_msgs == 0) return TFS_ERR_INVALID_ARG;
  tfs_coalescer* co = new tfs_coalescer();
  co->eng = eng;
  co->window_us = window_us;
  co->max_msgs = max_msgs;
  *out = co;
  return TFS_OK;
}

tfs_status_t tfs_coalescer_destroy(tfs_coalescer_t* co) {
  if (!co) return TFS_OK;
  {
    std::lock_guard<std::mutex> lock(co->m);
    for (auto& kv : co->leased)  // bytes nobody released: give the slabs back
      if (kv.second->slab.ptr) tfs_stage_release(co->eng->lanes[kv.second->lane]->ctx, &kv.second->slab);
    co->leased.clear();
  }
  delete co;
  return TFS_OK;
}

tfs_status_t tfs_coalescer_encode(tfs_coalescer_t* co, const tfs_batch_t* one, uint32_t flags, const uint8_t** wire_out, uint64_t* len_out,
                                  uint64_t* lease_out) {
  if (!co || !one || !wire_out || !len_out || !lease_out || one->n_msgs != 1) return TFS_ERR_INVALID_ARG;
  const tfs_msg_desc_t& m = one->msgs[0];
  const bool input_kind = m.kind == TFS_MSG_INPUT || m.kind == TFS_MSG_CLASSIFY_REQUEST;
  // every index is checked BEFORE anything joins the shared batch
  if (static_cast<uint64_t>(m.first_tensor) + m.n_tensors > one->n_tensors) return TFS_ERR_INVALID_ARG;
  for (uint32_t i = 0; i < m.n_tensors; ++i) {
    const tfs_tensor_desc_t& td = one->tensors[m.first_tensor + i];
    if (static_cast<uint64_t>(td.first_field) + td.n_fields > one->n_fields) return TFS_ERR_INVALID_ARG;
  }
  if (!input_kind && static_cast<uint64_t>(m.first_filter) + m.n_filters > one->n_filters) return TFS_ERR_INVALID_ARG;
  std::shared_ptr<Pending> p;
  uint32_t my_index = 0;
  bool leader = false;
  {
    std::unique_lock<std::mutex> lock(co->m);
    if (co->open && (co->open->closed || co->open->flags != flags)) co->open.reset();  // (a batch has ONE set of flags)
    if (!co->open) {
      co->open = std::make_shared<Pending>();
      co->open->flags = flags;
      co->open->opened = std::chrono::steady_clock::now();
      leader = true;
    }
    p = co->open;
    // append this message's descriptors, indices rebased into the batch's arrays
    my_index = static_cast<uint32_t>(p->msgs.size());
    tfs_msg_desc_t md = m;
    md.first_tensor = static_cast<uint32_t>(p->tensors.size());
    md.first_filter = static_cast<uint32_t>(p->filters.size());
    for (uint32_t i = 0; i < m.n_tensors; ++i) {
      tfs_tensor_desc_t td = one->tensors[m.first_tensor + i];
      const uint32_t f0 = static_cast<uint32_t>(p->fields.size());
      for (uint32_t k = 0; k < td.n_fields; ++k) p->fields.push_back(one->fields[td.first_field + k]);
      td.first_field = f0;
      p->tensors.push_back(td);
    }
    if (!input_kind)
      for (uint32_t i = 0; i < m.n_filters; ++i) p->filters.push_back(one->filters[m.first_filter + i]);
    p->msgs.push_back(md);
    p->refs++;
    if (p->msgs.size() >= co->max_msgs) {  // full: whoever fills it closes it (the leader may still be asleep)
      p->closed = true;
      co->open.reset();
      co->cv.notify_all();
    }
    if (leader) {
      // the first caller of a batch keeps the window open, then closes the batch and encodes it
      const auto deadline = p->opened + std::chrono::microseconds(co->window_us);
      co->cv.wait_until(lock, deadline, [&] { return p->closed; });
      if (!p->closed) {
        p->closed = true;
        if (co->open == p) co->open.reset();
      }
    }
  }
  if (leader) {
    run_batch(co, p);
  } else {
    std::unique_lock<std::mutex> lock(co->m);
    co->cv.wait(lock, [&] { return p->done; });
  }
  if (p->status != TFS_OK) return p->status;
  const uint64_t prefix = (p->flags & TFS_F_GRPC_PREFIX) ? 5u : 0u;
  *wire_out = static_cast<const uint8_t*>(p->slab.ptr) + p->off[my_index];
  *len_out = p->len[my_index] + prefix;
  *lease_out = p->lease;
  return TFS_OK;
}

tfs_status_t tfs_coalescer_release(tfs_coalescer_t* co, uint64_t lease) {
  if (!co) return TFS_ERR_INVALID_ARG;
  std::shared_ptr<Pending> p;
  {
    std::lock_guard<std::mutex> lock(co->m);
    auto it = co->leased.find(lease);
    if (it == co->leased.end()) return TFS_ERR_INVALID_ARG;
    p = it->second;
    if (--p->refs != 0) return TFS_OK;
    co->leased.erase(it);
  }
  return tfs_stage_release(co->eng->lanes[p->lane]->ctx, &p->slab);  // the last reference: the slab goes back to the pool
}

tfs_status_t tfs_coalescer_stats(tfs_coalescer_t* co, uint64_t* batches, uint64_t* messages) {
  if (!co) return TFS_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(co->m);
  if (batches) *batches = co->batches;
  if (messages) *messages = co->messages;
  return TFS_OK;
}

}  // extern "C"
