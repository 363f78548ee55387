// tfs_host.cu — the host-memory entry points: what the Rust shim calls from inside tonic's body stream,
// where the message payloads (Vec<f32>, ...) and the BytesMut both live in host memory.  The ctx's device
// staging buffers bridge them to HBM; host<->device copies are part of these calls (the `e2e` number).
#include <cstring>

#include "tfs_internal.h"

struct tfs_decoded;

namespace tfs {
tfs_status_t encode_device(tfs_ctx* ctx, const tfs_batch_t* batch, void* dst_dev, uint64_t dst_capacity, uint32_t flags,
                           uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out, bool size_only);
tfs_status_t decode_scan_device(tfs_ctx* ctx, const void* wire_dev, uint64_t wire_bytes, const uint64_t* msg_off,
                                const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, tfs_decoded** out);
tfs_status_t decode_device(tfs_ctx* ctx, tfs_decoded* dec, void* pool_dev, uint64_t pool_bytes);
uint64_t decoded_pool_bytes(const tfs_decoded* dec);
void decoded_mark_host_staged(tfs_decoded* dec, uint64_t stage_generation);
bool decoded_stage_is_current(const tfs_decoded* dec, uint64_t stage_generation);
}  // namespace tfs

using namespace tfs;

extern "C" {

tfs_status_t tfs_encode_host(tfs_ctx_t* ctx, const tfs_batch_t* batch, void* dst_host, uint64_t dst_capacity, uint32_t flags,
                             uint64_t* msg_off_out, uint64_t* msg_len_out, uint64_t* total_out) {
  if (!ctx || !batch || (!dst_host && dst_capacity)) return TFS_ERR_INVALID_ARG;
  if ((batch->n_fields && !batch->fields)) return TFS_ERR_INVALID_ARG;
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->stage_generation++;  // the staging buffers are reused: a pending tfs_decode_host handle becomes stale
  // 1. stage every payload array into one device slab (256-byte aligned pieces), host -> device
  std::vector<tfs_field_desc_t> fields(batch->fields, batch->fields + batch->n_fields);
  std::vector<uint64_t> at(batch->n_fields, 0);
  uint64_t in_bytes = 0;
  for (uint32_t i = 0; i < batch->n_fields; ++i) {
    const uint64_t b = fields[i].n_elems * tfs_type_size(fields[i].src_type);
    if (b && (!fields[i].data || tfs_type_size(fields[i].src_type) == 0)) return set_err(ctx, TFS_ERR_INVALID_ARG, "payload field without data or type");
    at[i] = in_bytes;
    in_bytes = align_up(in_bytes + b, 256);
  }
  TFS_TRY(dev_reserve(ctx, &ctx->d_stage_in, in_bytes + 256));
  TFS_TRY(dev_reserve(ctx, &ctx->d_stage_out, dst_capacity + 256));
  uint8_t* din = static_cast<uint8_t*>(ctx->d_stage_in.p);
  for (uint32_t i = 0; i < batch->n_fields; ++i) {
    const uint64_t b = fields[i].n_elems * tfs_type_size(fields[i].src_type);
    if (b) TFS_CUDA(ctx, cudaMemcpyAsync(din + at[i], fields[i].data, b, cudaMemcpyHostToDevice, ctx->stream));
    fields[i].data = b ? din + at[i] : nullptr;
  }
  tfs_batch_t dev_batch = *batch;
  dev_batch.fields = fields.data();
  // 2. encode on the device, 3. wire bytes device -> host
  uint64_t total = 0;
  TFS_TRY(encode_device(ctx, &dev_batch, ctx->d_stage_out.p, dst_capacity, flags, msg_off_out, msg_len_out, &total, false));
  if (total) TFS_CUDA(ctx, cudaMemcpyAsync(dst_host, ctx->d_stage_out.p, total, cudaMemcpyDeviceToHost, ctx->stream));
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (total_out) *total_out = total;
  return TFS_OK;
}

tfs_status_t tfs_decode_host(tfs_ctx_t* ctx, const void* wire_host, uint64_t wire_bytes, const uint64_t* msg_off,
                             const uint64_t* msg_len, uint32_t n_msgs, int32_t msg_kind, uint32_t flags, tfs_decoded_t** out) {
  if (!ctx || !out || (n_msgs && (!msg_off || !msg_len)) || (wire_bytes && !wire_host)) return TFS_ERR_INVALID_ARG;
  *out = nullptr;
  TFS_CUDA(ctx, cudaSetDevice(ctx->device));
  TFS_TRY(dev_reserve(ctx, &ctx->d_stage_in, wire_bytes + 256));
  if (wire_bytes) TFS_CUDA(ctx, cudaMemcpyAsync(ctx->d_stage_in.p, wire_host, wire_bytes, cudaMemcpyHostToDevice, ctx->stream));
  const uint64_t gen = ++ctx->stage_generation;
  tfs_status_t st = decode_scan_device(ctx, ctx->d_stage_in.p, wire_bytes, msg_off, msg_len, n_msgs, msg_kind, flags, out);
  if (*out) decoded_mark_host_staged(*out, gen);
  return st;
}

tfs_status_t tfs_decode_host_finish(tfs_ctx_t* ctx, tfs_decoded_t* dec, void* pool_host, uint64_t pool_bytes) {
  if (!ctx || !dec) return TFS_ERR_INVALID_ARG;
  if (!decoded_stage_is_current(dec, ctx->stage_generation))
    return set_err(ctx, TFS_ERR_INVALID_ARG, "stale handle: another *_host call reused the staging buffers since tfs_decode_host");
  const uint64_t need = decoded_pool_bytes(dec);
  if (pool_bytes < need || (need && !pool_host)) return set_err(ctx, TFS_ERR_NOMEM, "value pool is smaller than tfs_decoded_view().pool_bytes");
  TFS_TRY(dev_reserve(ctx, &ctx->d_stage_out, need + 256));
  TFS_TRY(decode_device(ctx, dec, ctx->d_stage_out.p, need));
  if (need) TFS_CUDA(ctx, cudaMemcpyAsync(pool_host, ctx->d_stage_out.p, need, cudaMemcpyDeviceToHost, ctx->stream));
  TFS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return TFS_OK;
}

}  // extern "C"
