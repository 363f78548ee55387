//! `mod ffi` of src/lib.rs: the C ABI of include/tfs_gpu.h.  UNCOMPILED here (no rustc); layouts are checked against
//! the C header by tests/test_abi.py through the ctypes mirror, which uses the same field order and padding.
#![allow(non_camel_case_types, dead_code)]
use std::os::raw::{c_char, c_void};

pub const TFS_ABI_VERSION: i32 = 3;
pub const TFS_MAX_DIMS: usize = 8;
pub const TFS_N_SLOTS: usize = 14;   // proto fields 4-13, 16, 17, 14, 15 (tfs_slot_of_kind)

pub const TFS_OK: i32 = 0;
pub const TFS_ERR_UNSUPPORTED: i32 = 12;
pub const TFS_ERR_BATCH: i32 = 23;
pub const TFS_ERR_TOO_LARGE: i32 = 24;

pub const TFS_KIND_TENSOR_CONTENT: i32 = 4;
pub const TFS_KIND_FLOAT_VAL: i32 = 5;
pub const TFS_KIND_DOUBLE_VAL: i32 = 6;
pub const TFS_KIND_INT_VAL: i32 = 7;
pub const TFS_KIND_SCOMPLEX_VAL: i32 = 9;
pub const TFS_KIND_INT64_VAL: i32 = 10;
pub const TFS_KIND_STRING_VAL: i32 = 8;
pub const TFS_KIND_BOOL_VAL: i32 = 11;
pub const TFS_KIND_DCOMPLEX_VAL: i32 = 12;
pub const TFS_KIND_HALF_VAL: i32 = 13;
pub const TFS_KIND_UINT32_VAL: i32 = 16;
pub const TFS_KIND_UINT64_VAL: i32 = 17;
pub const TFS_KIND_RESOURCE_HANDLE_VAL: i32 = 14;   // TFS_T_RAGGED: elements are ENCODED ResourceHandleProto bodies
pub const TFS_KIND_VARIANT_VAL: i32 = 15;           // TFS_T_RAGGED: elements are ENCODED VariantTensorDataProto bodies

pub const TFS_T_U8: i32 = 1;
pub const TFS_T_I32: i32 = 5;
pub const TFS_T_U32: i32 = 6;
pub const TFS_T_I64: i32 = 7;
pub const TFS_T_U64: i32 = 8;
pub const TFS_T_F32: i32 = 11;
pub const TFS_T_F64: i32 = 12;
pub const TFS_T_BOOL: i32 = 13;
pub const TFS_T_RAGGED: i32 = 14;

pub const TFS_MSG_PREDICT_REQUEST: i32 = 1;
pub const TFS_MSG_PREDICT_RESPONSE: i32 = 2;
pub const TFS_F_GRPC_PREFIX: u32 = 1;
pub const TFS_F_ALIGN_PAYLOAD: u32 = 2;

#[repr(C)] pub struct tfs_ctx_t { _private: [u8; 0] }
#[repr(C)] pub struct tfs_decoded_t { _private: [u8; 0] }

#[repr(C)] #[derive(Clone, Copy)]
pub struct tfs_slab_t { pub ptr: *mut c_void, pub capacity: u64, pub id: u64 }

#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct tfs_span_t { pub off: u64, pub len: u32, pub _pad: u32 }

#[repr(C)] #[derive(Clone, Copy)]
pub struct tfs_str_t { pub ptr: *const c_char, pub len: u32, pub _pad: u32 }

#[repr(C)] #[derive(Clone, Copy)]
pub struct tfs_field_desc_t {
    pub kind: i32, pub src_type: i32, pub dst_type: i32, pub op: i32, pub op_param: f32, pub _pad: u32,
    pub n_elems: u64, pub data: *const c_void,
    pub aux: u64, // string_val (TFS_T_RAGGED): total bytes of all elements
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct tfs_tensor_desc_t {
    pub name: tfs_str_t, pub dtype: i32, pub version_number: i32, pub has_shape: i32, pub unknown_rank: i32,
    pub n_dims: u32, pub first_field: u32, pub n_fields: u32, pub _pad: u32, pub dims: [i64; TFS_MAX_DIMS],
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct tfs_msg_desc_t {
    pub kind: i32, pub has_model_spec: i32, pub model_name: tfs_str_t, pub signature_name: tfs_str_t,
    pub version_label: tfs_str_t, pub version_choice: i32, pub _pad: u32, pub version: i64,
    pub first_tensor: u32, pub n_tensors: u32, pub first_filter: u32, pub n_filters: u32,
}

#[repr(C)]
pub struct tfs_batch_t {
    pub msgs: *const tfs_msg_desc_t, pub tensors: *const tfs_tensor_desc_t, pub fields: *const tfs_field_desc_t,
    pub filters: *const tfs_str_t, pub n_msgs: u32, pub n_tensors: u32, pub n_fields: u32, pub n_filters: u32,
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct tfs_tensor_out_t {
    pub key: tfs_span_t, pub msg_index: u32, pub dtype: i32, pub version_number: i32, pub has_shape: i32,
    pub unknown_rank: i32, pub n_dims: u32, pub dims: [i64; TFS_MAX_DIMS], pub n_elems: [u64; TFS_N_SLOTS],
    pub dst_off: [u64; TFS_N_SLOTS], pub dst_type: [i32; TFS_N_SLOTS],
}

#[repr(C)] #[derive(Clone, Copy)]
pub struct tfs_msg_out_t {
    pub status: i32, pub has_model_spec: i32, pub err_offset: u64, pub model_name: tfs_span_t,
    pub signature_name: tfs_span_t, pub version_label: tfs_span_t, pub version_choice: i32, pub _pad: u32,
    pub version: i64, pub first_tensor: u32, pub n_tensors: u32, pub first_filter: u32, pub n_filters: u32,
}

#[repr(C)]
pub struct tfs_decoded_view_t {
    pub msgs: *const tfs_msg_out_t, pub tensors: *const tfs_tensor_out_t, pub filters: *const tfs_span_t,
    pub strings: *const u8, pub strings_bytes: u64, pub pool_bytes: u64,
    pub n_msgs: u32, pub n_tensors: u32, pub n_filters: u32, pub n_failed: u32,
}

extern "C" {
    pub fn tfs_abi_version() -> i32;
    pub fn tfs_strerror(status: i32) -> *const c_char;
    pub fn tfs_ctx_create(device: i32, pinned_bytes: u64, out: *mut *mut tfs_ctx_t) -> i32;
    pub fn tfs_ctx_destroy(ctx: *mut tfs_ctx_t) -> i32;
    pub fn tfs_last_error_detail(ctx: *const tfs_ctx_t) -> *const c_char;
    pub fn tfs_last_error_location(ctx: *const tfs_ctx_t, msg_index: *mut u32, offset: *mut u64) -> i32;
    pub fn tfs_stage_acquire(ctx: *mut tfs_ctx_t, bytes: u64, align_skew: u32, out: *mut tfs_slab_t) -> i32;
    pub fn tfs_stage_release(ctx: *mut tfs_ctx_t, slab: *const tfs_slab_t) -> i32;
    pub fn tfs_stage_stats(ctx: *mut tfs_ctx_t, in_use: *mut u64, pool_bytes: *mut u64, fallbacks: *mut u64) -> i32;
    pub fn tfs_sync(ctx: *mut tfs_ctx_t) -> i32;
    pub fn tfs_host_alloc(ctx: *mut tfs_ctx_t, bytes: u64, out: *mut *mut c_void) -> i32;
    pub fn tfs_host_free(ctx: *mut tfs_ctx_t, p: *mut c_void) -> i32;
    pub fn tfs_encode_bound(batch: *const tfs_batch_t, flags: u32) -> u64;
    pub fn tfs_encode_size(ctx: *mut tfs_ctx_t, batch: *const tfs_batch_t, msg_len_out: *mut u64) -> i32;
    pub fn tfs_encode_host(ctx: *mut tfs_ctx_t, batch: *const tfs_batch_t, dst_host: *mut c_void, dst_capacity: u64, flags: u32,
                           msg_off_out: *mut u64, msg_len_out: *mut u64, total_out: *mut u64) -> i32;
    pub fn tfs_decode_host(ctx: *mut tfs_ctx_t, wire_host: *const c_void, wire_bytes: u64, msg_off: *const u64, msg_len: *const u64,
                           n_msgs: u32, msg_kind: i32, flags: u32, out: *mut *mut tfs_decoded_t) -> i32;
    pub fn tfs_decoded_view(dec: *const tfs_decoded_t, view: *mut tfs_decoded_view_t) -> i32;
    pub fn tfs_decode_host_finish(ctx: *mut tfs_ctx_t, dec: *mut tfs_decoded_t, pool_host: *mut c_void, pool_bytes: u64) -> i32;
    pub fn tfs_decoded_free(ctx: *mut tfs_ctx_t, dec: *mut tfs_decoded_t) -> i32;
    pub fn tfs_pack_int64_noalloc(src_dev: *const i64, n_elems: u64, dst_dev: *mut c_void, dst_capacity: u64,
                                  body_len_dev: *mut u64, cuda_stream: *mut c_void) -> i32;
}
