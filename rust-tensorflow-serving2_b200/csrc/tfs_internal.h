// tfs_internal.h — structures shared by the host orchestration and the sm_100a kernels.
// Nothing here is part of the ABI (include/tfs_gpu.h is).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <chrono>
#include <cstdio>
#include <memory>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/tfs_gpu.h"

namespace tfs {

// ---- geometry --------------------------------------------------------------------------------
constexpr int kVarintThreads = 256;                  // 8 warps per CTA
constexpr int kVarintRows = 8;                       // elements per lane (one per 32-element row)
constexpr int kVarintTile = kVarintThreads * kVarintRows;  // 2048 elements per tile
constexpr int kVarintMaxLen = 10;
constexpr int kCopyThreads = 256;
constexpr int kCopyTileBytes = 32768;                // output bytes per copy tile
// A copy segment of at most this many OUTPUT bytes owns no tiles at all: k_copy_small takes it, one warp per segment
// (a batch of 10^6 twelve-byte float_val bodies would otherwise launch two million CTAs).  The encoder's host tables,
// the decoder's structure walk and the copy kernels apply the same rule.
constexpr uint32_t kSmallSegBytes = 512;
constexpr int kDecThreads = 256;
constexpr int kDecTileBytes = 4096;                  // wire bytes per varint-decode tile (16 B per thread of the gather kernel)

// ---- element fetch modes of the varint encoder (source memory type -> u64 that is varint-coded) ----
enum VarintMode : int32_t {
  VM_I64 = 0,  // also u64: the bits are the value
  VM_I32,      // sign-extended (`as u64` of an i32)
  VM_U32,
  VM_I16,
  VM_U16,      // also f16 / bf16 bit patterns
  VM_I8,
  VM_U8,
  VM_F32_F16,  // f32 -> binary16 pattern (RNE)
  VM_F32_BF16, // f32 -> bfloat16 pattern (RNE)
  VM_COUNT
};

// ---- bulk copy ops ------------------------------------------------------------------------------
enum CopyOp : int32_t {
  CP_BYTES = 0,
  CP_BOOL,        // byte -> (byte != 0)
  CP_U8_F32,      // u8 -> f32
  CP_U8_F32_DIV,  // u8 -> f32 / param
  CP_U8_F32_MUL,  // u8 -> f32 * param
  CP_F32_F16,
  CP_F32_BF16,
  CP_F16_F32,
  CP_BF16_F32,
  CP_COUNT
};

struct StrRef {
  uint32_t off;  // into the string blob
  uint32_t len;
};

// One bulk copy / cast segment.  `n` counts OUTPUT elements of the op (bytes for CP_BYTES / CP_BOOL).
struct CopySeg {
  const uint8_t* src;
  uint8_t* dst;
  uint64_t n;
  uint64_t first_tile;  // exclusive prefix of tiles over the segment table (host- or device-built)
  int32_t op;
  float param;
};

struct EncField {
  const void* src;
  uint64_t n_elems;
  uint64_t body_len;   // wire bytes of the packed body (copy class: host-known; varint class: device scan)
  uint64_t dst_off;    // absolute byte offset of the body in dst (written by the header kernel)
  uint32_t first_tile; // varint class: first varint tile; copy class: index into the CopySeg table
  uint32_t n_tiles;
  int32_t kind;        // TFS_KIND_*
  int32_t mode;        // VarintMode or CopyOp
  float op_param;
  uint32_t is_varint;  // FieldClass
  uint64_t aux;        // ragged class: total bytes of the elements as the descriptor states them
};
enum FieldClass : uint32_t { FC_COPY = 0, FC_VARINT = 1, FC_RAGGED = 2 };

// One 2048-element tile of a varint-coded field (built on the device by k_tile_map).
struct VTile {
  const uint8_t* src;  // first element of the tile
  uint32_t n;          // elements in the tile (2048 except for a field's last tile)
  uint16_t mode;       // VarintMode
  uint16_t esize;      // bytes per source element
};

struct EncTensor {
  StrRef name;
  int32_t dtype;
  int32_t version_number;
  int32_t has_shape;
  int32_t unknown_rank;
  uint32_t n_dims;
  uint32_t first_field;
  uint32_t n_fields;  // sorted by kind (ascending tag order)
  uint32_t _pad;
  int64_t dims[TFS_MAX_DIMS];
};

struct EncMsg {
  int32_t kind;
  int32_t has_model_spec;
  StrRef model_name;
  StrRef signature_name;
  StrRef version_label;
  int32_t version_choice;
  uint32_t _pad;
  int64_t version;
  uint32_t first_tensor;
  uint32_t n_tensors;
  uint32_t first_filter;
  uint32_t n_filters;
};

// Device view of one compiled encode batch (all pointers are device pointers).
struct EncTables {
  EncMsg* msgs;
  EncTensor* tensors;
  EncField* fields;
  StrRef* filters;
  const uint8_t* strings;
  uint32_t* vfield_first_tile;  // [n_vfields + 1]
  uint32_t* vfield_index;       // [n_vfields] -> index into fields
  CopySeg* segs;                // [n_segs]
  uint32_t* rfield_first_tile;  // [n_rfields + 1] ragged fields: first sizing tile (2048 elements each)
  uint32_t* rfield_first_out;   // [n_rfields + 1] first copy tile (64 KiB of records each, from the host's upper bound)
  uint32_t* rfield_index;       // [n_rfields] -> index into fields
  // work arrays
  VTile* vtiles;                // [n_vtiles]
  uint32_t* tile_field;         // [n_vtiles] index into fields
  uint32_t* tile_bytes;         // [n_vtiles]
  uint64_t* tile_off;           // [n_vtiles] byte offset of the tile inside its field body
  uint32_t* ticket;             // [2] meeting counters of k_frame_small's CTAs (tfs_ctx::d_ctrl: zero between launches)
  uint64_t* tensor_len;         // [n_tensors]
  uint64_t* entry_len;          // [n_tensors]
  uint64_t* feat_rec;           // [n_tensors] Input kinds: bytes of the feature's map entry record (key + length + entry)
  int32_t* feat_ex;             // [n_tensors] Input kinds: index of the Example (compact copy for the per-message pass)
  uint64_t* entry_off;          // [n_tensors] Input kinds: where the feature's map entry starts in dst (per-message pass)
  uint64_t* spec_len;           // [n_msgs]
  uint64_t* msg_len;            // [n_msgs]
  uint64_t* msg_pay_off;        // [n_msgs] offset (inside the message) of its largest bulk payload
  uint64_t* msg_off;            // [n_msgs]
  uint64_t* rtile_bytes;        // [n_rtiles] wire bytes of the records of each ragged sizing tile
  uint64_t* rtile_off;          // [n_rtiles] exclusive prefix inside the field's record area
  uint64_t* result;             // [4]: total bytes, overflow flag, ragged offsets invalid, -
  uint32_t n_msgs, n_tensors, n_fields, n_filters;
  uint32_t n_vfields, n_vtiles, n_segs, n_rfields;
  uint32_t n_rtiles, n_routs;
};

// ---- decode ---------------------------------------------------------------------------------------
enum SpanClass : uint32_t {
  SC_VARINT = 0,  // packed or single varint(s) -> slot element type
  SC_COPY = 1,    // fixed-width packed body / tensor_content / single fixed value
  SC_STR = 2,     // one string_val element (repeated bytes: one span per element, never packed)
};

// One contiguous run of payload bytes found by the structure walk.
struct DecSpan {
  uint64_t wire_off;   // absolute offset in the wire buffer
  uint64_t len;        // bytes
  uint64_t elem_base;  // index of its first element inside the (tensor, slot) vector (finalize kernel)
  uint64_t n_elems;    // varint: terminators counted; copy: len / width
  uint32_t tensor;     // global tensor index
  uint32_t slot;
  uint32_t msg;
  uint32_t cls;        // SpanClass
  uint64_t first_tile; // exclusive prefix of tiles (per class)
  uint64_t elem_tile0; // varint: index into tile_count[] of its first tile (== first_tile); string: byte offset inside the tensor's string bytes
};

// One 4 KiB tile of a packed varint body (k_dtile_map) and where its values go (k_dtile_bind).
struct TileRef {
  uint64_t span_abs0;  // absolute address of the first byte of the span
  uint64_t span_end;   // absolute address one past its last byte
  uint64_t tile_abs;   // absolute (16-byte aligned) address of the tile's first byte
  uint32_t span;
  uint32_t msg;
};
struct TileDst {
  uint8_t* dst;    // pool address of the tile's first value
  int32_t type;    // TFS_T_* the values are written as, TFS_T_NONE = skip (the message failed)
  int32_t dtype;   // tensor dtype (19 / 14 select the f16 / bf16 widening)
};

struct DecTables {
  const uint8_t* wire;       // where the value kernels read the wire bytes (tile grids are cut on THESE addresses)
  const uint8_t* walk_wire;  // where the structure walk reads them: the same, or the pinned host copy while `wire` is
                             // still being filled by the copy engine (tfs_decode_host_into)
  uint64_t wire_bytes;
  const uint64_t* msg_off;  // device copies of the caller's arrays
  const uint64_t* msg_len;
  uint32_t n_msgs;
  int32_t msg_kind;
  uint32_t flags;
  uint32_t _pad;
  // pass-1 counts per message (then exclusive-scanned in place): tensors, filters, spans, string bytes
  uint32_t* cnt_tensors;
  uint32_t* cnt_filters;
  uint32_t* cnt_spans;
  uint32_t* cnt_strbytes;
  uint32_t* cnt_vtiles;
  uint32_t* cnt_ctiles;
  uint64_t* totals;  // [8]: tensors, filters, spans, strbytes, varint tiles, copy tiles, pool bytes, failed
  // pass-2 outputs
  tfs_msg_out_t* msgs;
  tfs_tensor_out_t* tensors;
  tfs_span_t* filters;
  uint8_t* strings;
  DecSpan* spans;
  uint32_t* span_vtile0;  // [n_spans] first varint tile of each span (non-decreasing over the table)
  uint32_t* span_ctile0;  // [n_spans] first copy tile of each span (non-decreasing over the table)
  uint64_t* tensor_pool;  // [n_tensors] pool bytes of each tensor
  uint32_t* tile_count;  // [n varint tiles] terminators per tile
  TileRef* dtiles;       // [n varint tiles]
  TileDst* dtile_dst;    // [n varint tiles]
  uint64_t* tile_base;   // [n varint tiles] exclusive prefix inside the span
  CopySeg* segs;         // copy-class spans as segments (dst filled once the pool is known)
  unsigned long long* err;  // [n_msgs] min over (offset << 8 | code) of device-detected errors, ~0 = none
};

}  // namespace tfs

// ---- context ---------------------------------------------------------------------------------------
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

struct tfs_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  std::string err;
  // where the last batch decode failed first: message index and byte offset inside that message (tfs_last_error_location)
  bool err_has_location = false;
  uint32_t err_msg_index = 0;
  uint64_t err_byte_offset = 0;
  uint64_t launches = 0;
  bool trace = false;  // TFS_TRACE=1 in the environment: host phase times of every batch call on stderr (HostTrace)
  DevBuf d_tables, d_work, d_stage_in, d_stage_out;
  DevBuf d_dec1, d_dec2, d_dec3;  // decode tables (pass 1 / pass 2 / varint tiles); owned by the newest tfs_decoded handle
  // small-batch decode plans its tables optimistically from what earlier batches on this context needed
  uint64_t dec_hint_tensors = 0, dec_hint_filters = 0, dec_hint_spans = 0, dec_hint_str = 0, dec_hint_blob = 0;
  bool copy_attr_set = false;
  bool dec_spec_tma = false;  // the last planned decode on this ctx was all bulk copies: launch the TMA ring ahead of the sync next time
  uint32_t* d_ctrl = nullptr;  // 256 bytes of control words, zeroed once (grid-wide meeting counters that clear themselves)
  // single-walk decode plan: per-message quota slots, from what the largest message of earlier batches needed
  uint32_t dec_q_t = 4, dec_q_s = 8, dec_q_f = 2, dec_q_str = 256;
  // Host entry points with pinned buffers: the copy engine brings the source to HBM in chunks on `copy_stream` while the
  // copy kernel already writes the finished chunks to the pinned destination.  `pipe` lists, for the NEXT launch_copy,
  // the event of every chunk and the device address its bytes end at.
  struct PipePiece {
    cudaEvent_t ready;
    uint64_t src_hi;
  };
  cudaStream_t copy_stream = nullptr;
  std::vector<cudaEvent_t> pipe_events;  // recycled
  std::vector<PipePiece> pipe;
  uint64_t dec_generation = 0;
  uint64_t stage_generation = 0;  // bumped by every *_host call that reuses the staging buffers
  void* h_up = nullptr;
  size_t h_up_cap = 0;
  void* h_down = nullptr;
  size_t h_down_cap = 0;
  // pageable host buffers: helper threads copy them through a pinned ring of four slots while the copy engine drains /
  // fills the others (tfs_host.cu: StageCrew, staged_h2d / staged_d2h)
  std::shared_ptr<void> stage_crew;
  void* h_ring = nullptr;
  cudaEvent_t ring_ev[4] = {nullptr, nullptr, nullptr, nullptr};
  bool ring_busy[4] = {false, false, false, false};
  uint32_t ring_next = 0;
  cudaEvent_t dec_tables_ev = nullptr;  // one-call decode: recorded when the result tables are final (before the value kernels)
  cudaEvent_t enc_layout_ev = nullptr;  // TFS_F_ASYNC: recorded when the framing is done; the results come back behind it
  std::shared_ptr<void> enc_scratch;  // the encoder's host tables (tfs_encode.cu: Compiled), kept between calls: grow-only
  void* h_rec = nullptr;  // the encoder compiles its message / tensor / field records straight into this pinned block
  size_t h_rec_cap = 0;
  // pinned staging pool (tfs_stage_acquire / release): one region pinned at tfs_ctx_create, carved into power-of-two
  // classes from 4 KiB up and recycled through per-class free lists
  void* h_slab = nullptr;
  size_t h_slab_cap = 0;
  std::mutex pool_mutex;
  size_t pool_bump = 0;                               // first byte of the region that was never handed out
  std::vector<std::vector<uint64_t>> pool_free;       // [class] offsets of free blocks
  std::vector<std::pair<void*, uint64_t>> pool_extra; // one-off allocations (the pool was exhausted): base, id
  uint64_t pool_in_use = 0, pool_fallbacks = 0, pool_next_id = 1;
  // optional per-kernel timing with CUDA events on the launching stream (tfs_prof_*)
  bool prof_on = false;
  std::vector<cudaEvent_t> prof_pool;                           // recycled events
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_open[12];  // (start, stop) pairs per kernel id, not yet read
  uint64_t prof_launches[12] = {0};
  double prof_ms[12] = {0};
};

namespace tfs {
enum ProfKernel { PK_COPY = 0, PK_VARINT_SIZE, PK_VARINT_SCATTER, PK_VARINT_COUNT, PK_VARINT_GATHER, PK_COPY_TMA, PK_WALK, PK_FRAME, PK_COPY_CAST, PK_COUNT_ };
// Brackets one kernel launch with events when profiling is on (a no-op otherwise).
struct ProfScope {
  tfs_ctx* ctx;
  int id;
  cudaEvent_t start = nullptr, stop = nullptr;
  ProfScope(tfs_ctx* c, int kernel_id);
  ~ProfScope();
};

// Host phase times of one call on stderr when TFS_TRACE is set (a no-op otherwise): `mark("name")` closes a phase.
struct HostTrace {
  const bool on;
  const char* what;
  std::chrono::steady_clock::time_point t0, last;
  std::string line;
  HostTrace(const tfs_ctx* c, const char* w) : on(c->trace), what(w) {
    if (on) t0 = last = std::chrono::steady_clock::now();
  }
  void mark(const char* phase) {
    if (!on) return;
    const auto now = std::chrono::steady_clock::now();
    char buf[64];
    std::snprintf(buf, sizeof buf, " %s %.3f", phase, std::chrono::duration<double, std::milli>(now - last).count());
    line += buf;
    last = now;
  }
  ~HostTrace() {
    if (on) std::fprintf(stderr, "[tfs] %s: total %.3f ms |%s\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(), line.c_str());
  }
};
}  // namespace tfs

namespace tfs {

tfs_status_t set_err(tfs_ctx* ctx, tfs_status_t st, const std::string& what);
tfs_status_t cuda_fail(tfs_ctx* ctx, cudaError_t e, const char* where);
tfs_status_t dev_reserve(tfs_ctx* ctx, DevBuf* b, size_t bytes);
tfs_status_t pinned_reserve(tfs_ctx* ctx, void** p, size_t* cap, size_t bytes);
// Pinned result blocks of the decoder.  A decode handle OWNS the block its tables were copied into (the view points into
// it: no second copy on the host); freed handles give their block to a small process-wide list (portable pinned memory,
// usable from every device), so a steady decode / read / free loop allocates nothing.
struct PinnedBlock {
  void* p = nullptr;
  size_t cap = 0;
};
tfs_status_t down_acquire(tfs_ctx* ctx, size_t bytes, PinnedBlock* out);
void down_release(PinnedBlock b);

#define TFS_CUDA(ctx, call)                                       \
  do {                                                            \
    cudaError_t e__ = (call);                                     \
    if (e__ != cudaSuccess) return tfs::cuda_fail(ctx, e__, #call); \
  } while (0)

#define TFS_TRY(expr)                      \
  do {                                     \
    tfs_status_t st__ = (expr);            \
    if (st__ != TFS_OK) return st__;       \
  } while (0)

// Bulk copy / cast kernel over a segment table that already lives on the device.
// With `gate` the kernels do nothing when *gate != 0.  `need_generic` = false promises that every non-empty
// segment is a plain byte copy whose source and destination are congruent mod 16 (TMA ring only).
// Host-side picture of a segment (only what a chunked launch needs to pick its tile range).
struct HostSeg {
  uint64_t src;             // device address of the first source byte
  uint64_t src_bytes;
  uint64_t first_tile;
  uint64_t n_tiles;
  uint64_t src_tile_min, src_tile_max;  // bounds of the source bytes one tile consumes (32 KiB for plain bytes, other for casts)
};
// `need_tma` = false promises that no segment is large enough for the TMA ring (>= 4 KiB of plain bytes).
// `host_segs` is needed only while ctx->pipe is set (sources still arriving in chunks): each chunk's launch then walks
// just the tiles that can end inside the chunk; without it the launch waits for the last chunk.
// `cast_kind`: 0 = no widening / narrowing segment can be large enough for the staged cast kernel, 1 = only widening ones
// (u8 -> f32, 16-bit -> f32), 2 = narrowing ones (f32 -> 16-bit) too (a larger input stage).
// `any_small`: some segment has at most kSmallSegBytes of output (it owns no tiles; k_copy_small takes it).
tfs_status_t launch_copy(tfs_ctx* ctx, const CopySeg* segs_dev, uint32_t n_segs, uint64_t n_tiles, const uint64_t* gate,
                         bool need_generic, const std::vector<HostSeg>* host_segs = nullptr, bool need_tma = true, int cast_kind = 0,
                         bool any_small = false);
// The TMA ring launched BEFORE the host knows the counts (the single-walk decode plan): the kernel reads the number of
// segments and tiles from device memory and does nothing unless *go != 0 (set by the plan kernel when the tables are final,
// the pool is large enough and every copy segment is bound).  `many_segs` only picks the warp-cooperative variant.
tfs_status_t launch_copy_tma_deferred(tfs_ctx* ctx, const CopySeg* segs_dev, const uint64_t* go, const uint64_t* n_segs_dev,
                                      const uint64_t* n_tiles_dev, bool many_segs);

inline uint64_t align_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

}  // namespace tfs
