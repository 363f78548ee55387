//! `GpuCodec`: the tonic codec `predict` uses instead of the generated stub's `ProstCodec` (reference src/lib.rs:265).
//! UNCOMPILED here (no rustc).  Targets tonic 0.1.0 (`tonic::codec::Codec`, `tokio_util::codec::{Encoder, Decoder}`),
//! prost 0.6.0 generated structs (`PredictRequest`, `PredictResponse`, `TensorProto`) and bytes 0.5.3.
use std::{collections::HashMap, ffi::CStr, os::raw::c_void, ptr, sync::{Arc, Mutex}};

use bytes::{BufMut, BytesMut};
use tonic::{codec::Codec, Status};

use crate::ffi::*;
use crate::{ModelSpec, PredictRequest, PredictResponse, TensorProto, TensorShapeProto};

/// Owns the context; destroyed when the last clone of the handle goes (ctx buffers, pinned pool, stream).
pub struct CtxBox(pub *mut tfs_ctx_t);
unsafe impl Send for CtxBox {}
impl Drop for CtxBox { fn drop(&mut self) { unsafe { tfs_ctx_destroy(self.0) } } }

/// One context per GPU, shared by the clones of `TensorflowServing` (src/lib.rs:148-156); a ctx is not re-entrant.
#[derive(Clone)]
pub struct GpuHandle(pub Arc<Mutex<CtxBox>>);

/// Messages smaller than this stay on prost: below it the host-built descriptors cost more than the CPU needs to
/// write the bytes (4.5 M msgs/s on the GPU path, 9.8 M on 16 CPU threads).
pub const MIN_GPU_BYTES: usize = 64 << 10;

impl GpuHandle {
    pub fn new(device: i32) -> Result<Self, Status> {
        if unsafe { tfs_abi_version() } != TFS_ABI_VERSION { return Err(Status::internal("libtfs_gpu: ABI mismatch")); }
        let mut ctx = ptr::null_mut();
        check(unsafe { tfs_ctx_create(device, 64 << 20, &mut ctx) })?; // no device => no codec: there is no CPU fallback
        Ok(GpuHandle(Arc::new(Mutex::new(CtxBox(ctx)))))
    }
}

/// A pinned slab of the context's staging pool, given back on drop.  `skew` shifts the first usable byte so that a
/// message received at `ptr()` has its bulk payload on a 16-byte boundary (tfs_stage_acquire, include/tfs_gpu.h).
pub struct Slab { gpu: GpuHandle, slab: tfs_slab_t }
impl Slab {
    pub fn acquire(gpu: &GpuHandle, bytes: usize, skew: u32) -> Result<Slab, Status> {
        let mut slab: tfs_slab_t = unsafe { std::mem::zeroed() };
        let ctx = gpu.0.lock().unwrap();
        check(unsafe { tfs_stage_acquire(ctx.0, bytes as u64, skew, &mut slab) })?;
        Ok(Slab { gpu: gpu.clone(), slab })
    }
    pub fn ptr(&self) -> *mut u8 { self.slab.ptr as *mut u8 }
    pub fn capacity(&self) -> usize { self.slab.capacity as usize }
}
impl Drop for Slab { fn drop(&mut self) { let ctx = self.gpu.0.lock().unwrap(); unsafe { tfs_stage_release(ctx.0, &self.slab) }; } }

fn check(status: i32) -> Result<(), Status> {
    if status == TFS_OK { return Ok(()); }
    // same gRPC code as tonic's ProstDecoder (Status::internal(err.to_string())); the TEXT is the library's class name
    // ("invalid varint", "buffer underflow" ...) - prost's own text adds "failed to decode Protobuf message: " and
    // the field path, which the library does not reproduce (tfs_last_error_location gives message index + offset)
    let text = unsafe { CStr::from_ptr(tfs_strerror(status)) }.to_string_lossy().into_owned();
    Err(Status::internal(text))
}

#[derive(Clone)]
pub struct GpuCodec { pub gpu: GpuHandle }
impl Default for GpuCodec { fn default() -> Self { GpuCodec { gpu: crate::default_gpu() } } }

impl Codec for GpuCodec {
    type Encode = PredictRequest;
    type Decode = PredictResponse;
    type Encoder = GpuEncoder;
    type Decoder = GpuDecoder;
    fn encoder(&mut self) -> GpuEncoder { GpuEncoder { gpu: self.gpu.clone() } }
    fn decoder(&mut self) -> GpuDecoder { GpuDecoder { gpu: self.gpu.clone() } }
}

pub struct GpuEncoder { gpu: GpuHandle }
pub struct GpuDecoder { gpu: GpuHandle }

/// Plain-data picture of one PredictRequest: the descriptors point into the message's Vecs; only the ragged fields
/// (string_val, resource_handle_val, variant_val) are flattened into TFS_T_RAGGED blocks that `ragged` keeps alive.
struct Described { msg: tfs_msg_desc_t, tensors: Vec<tfs_tensor_desc_t>, fields: Vec<tfs_field_desc_t>, ragged: Vec<Vec<u64>> }

fn s(x: &str) -> tfs_str_t { tfs_str_t { ptr: x.as_ptr() as *const _, len: x.len() as u32, _pad: 0 } }

/// The body of a nested message, as prost writes it (no key, no length).
fn body_of<M: prost::Message>(m: &M) -> Vec<u8> { let mut b = Vec::with_capacity(m.encoded_len()); m.encode(&mut b).unwrap(); b }

/// TFS_T_RAGGED source block: (n + 1) u64 offsets, then the element bytes; u64 storage keeps it 8-byte aligned.
fn ragged_block(elems: &[&[u8]]) -> (Vec<u64>, usize) {
    let n = elems.len();
    let total: usize = elems.iter().map(|e| e.len()).sum();
    let mut block = vec![0u64; n + 1 + (total + 7) / 8];
    let mut at = 0usize;
    for (i, e) in elems.iter().enumerate() {
        block[i] = at as u64;
        unsafe { std::ptr::copy_nonoverlapping(e.as_ptr(), (block.as_mut_ptr().add(n + 1) as *mut u8).add(at), e.len()) };
        at += e.len();
    }
    block[n] = at as u64;
    (block, total)
}

/// Payload bytes of a request (what decides between the GPU path and prost).
fn payload_bytes(req: &PredictRequest) -> usize {
    req.inputs.values().map(|t| t.tensor_content.len() + 4 * t.float_val.len() + 8 * t.double_val.len() + 4 * t.int_val.len()
        + 8 * t.int64_val.len() + 4 * t.half_val.len() + 4 * t.uint32_val.len() + 8 * t.uint64_val.len() + t.bool_val.len()
        + 4 * t.scomplex_val.len() + 8 * t.dcomplex_val.len() + t.string_val.iter().map(|e| e.len()).sum::<usize>()).sum()
}

fn describe(req: &PredictRequest) -> Described {
    let mut tensors = Vec::new();
    let mut fields = Vec::new();
    let mut ragged: Vec<Vec<u64>> = Vec::new();
    for (name, t) in req.inputs.iter() {                    // HashMap order = the order prost itself would emit
        let first_field = fields.len() as u32;
        let mut push = |kind: i32, ty: i32, n: usize, p: *const c_void| if n > 0 {
            fields.push(tfs_field_desc_t { kind, src_type: ty, dst_type: 0, op: 0, op_param: 0.0, _pad: 0, n_elems: n as u64, data: p, aux: 0 });
        };
        push(TFS_KIND_TENSOR_CONTENT, TFS_T_U8, t.tensor_content.len(), t.tensor_content.as_ptr() as *const _);
        push(TFS_KIND_FLOAT_VAL, TFS_T_F32, t.float_val.len(), t.float_val.as_ptr() as *const _);
        push(TFS_KIND_DOUBLE_VAL, TFS_T_F64, t.double_val.len(), t.double_val.as_ptr() as *const _);
        push(TFS_KIND_INT_VAL, TFS_T_I32, t.int_val.len(), t.int_val.as_ptr() as *const _);
        push(TFS_KIND_SCOMPLEX_VAL, TFS_T_F32, t.scomplex_val.len(), t.scomplex_val.as_ptr() as *const _);
        push(TFS_KIND_INT64_VAL, TFS_T_I64, t.int64_val.len(), t.int64_val.as_ptr() as *const _);
        push(TFS_KIND_BOOL_VAL, TFS_T_BOOL, t.bool_val.len(), t.bool_val.as_ptr() as *const _);
        push(TFS_KIND_DCOMPLEX_VAL, TFS_T_F64, t.dcomplex_val.len(), t.dcomplex_val.as_ptr() as *const _);
        push(TFS_KIND_HALF_VAL, TFS_T_I32, t.half_val.len(), t.half_val.as_ptr() as *const _);
        push(TFS_KIND_UINT32_VAL, TFS_T_U32, t.uint32_val.len(), t.uint32_val.as_ptr() as *const _);
        push(TFS_KIND_UINT64_VAL, TFS_T_U64, t.uint64_val.len(), t.uint64_val.as_ptr() as *const _);
        // repeated bytes / repeated message fields (after the closure's last use: it borrows `fields`; the library
        // sorts a tensor's fields by field number itself).  For the message-typed fields an element is the message's
        // encoded BODY: prost writes it, the library frames it.
        let string_elems: Vec<&[u8]> = t.string_val.iter().map(|e| e.as_slice()).collect();
        let handle_bodies: Vec<Vec<u8>> = t.resource_handle_val.iter().map(body_of).collect();
        let variant_bodies: Vec<Vec<u8>> = t.variant_val.iter().map(body_of).collect();
        for (kind, elems) in [(TFS_KIND_STRING_VAL, string_elems),
                              (TFS_KIND_RESOURCE_HANDLE_VAL, handle_bodies.iter().map(|e| e.as_slice()).collect()),
                              (TFS_KIND_VARIANT_VAL, variant_bodies.iter().map(|e| e.as_slice()).collect())] {
            if elems.is_empty() { continue; }
            let (block, total) = ragged_block(&elems);
            fields.push(tfs_field_desc_t { kind, src_type: TFS_T_RAGGED, dst_type: 0, op: 0, op_param: 0.0, _pad: 0,
                                           n_elems: elems.len() as u64, data: block.as_ptr() as *const _, aux: total as u64 });
            ragged.push(block);                                           // moving the Vec does not move its heap buffer
        }
        let mut dims = [0i64; TFS_MAX_DIMS];
        let shape = t.tensor_shape.as_ref();
        let n_dims = shape.map(|sh| sh.dim.len()).unwrap_or(0);
        for (i, d) in shape.iter().flat_map(|sh| sh.dim.iter()).enumerate().take(TFS_MAX_DIMS) { dims[i] = d.size; }
        tensors.push(tfs_tensor_desc_t {
            name: s(name), dtype: t.dtype, version_number: t.version_number, has_shape: shape.is_some() as i32,
            unknown_rank: shape.map(|sh| sh.unknown_rank as i32).unwrap_or(0), n_dims: n_dims as u32,
            first_field, n_fields: fields.len() as u32 - first_field, _pad: 0, dims,
        });
    }
    let ms = req.model_spec.as_ref();
    let (choice, version) = match ms.and_then(|m| m.version_choice.as_ref()) {
        Some(crate::model_spec::VersionChoice::Version(v)) => (1, *v),
        Some(crate::model_spec::VersionChoice::VersionLabel(_)) => (2, 0),
        None => (0, 0),
    };
    let empty = "";
    let msg = tfs_msg_desc_t {
        kind: TFS_MSG_PREDICT_REQUEST, has_model_spec: ms.is_some() as i32,
        model_name: s(ms.map(|m| m.name.as_str()).unwrap_or(empty)),
        signature_name: s(ms.map(|m| m.signature_name.as_str()).unwrap_or(empty)),
        version_label: s(match ms.and_then(|m| m.version_choice.as_ref()) {
            Some(crate::model_spec::VersionChoice::VersionLabel(l)) => l.as_str(), _ => empty }),
        version_choice: choice, _pad: 0, version,
        first_tensor: 0, n_tensors: tensors.len() as u32, first_filter: 0, n_filters: 0,
    };
    Described { msg, tensors, fields, ragged }
}

impl tokio_util::codec::Encoder for GpuEncoder {
    type Item = PredictRequest;
    type Error = Status;
    /// Appends exactly the protobuf bytes of `item` (tonic adds the 5-byte frame prefix itself).
    fn encode(&mut self, item: PredictRequest, dst: &mut BytesMut) -> Result<(), Status> {
        // what the GPU path leaves to prost: rank > 8, named dims; and small messages, where building the descriptors
        // costs more than writing the bytes (MIN_GPU_BYTES)
        if payload_bytes(&item) < MIN_GPU_BYTES
            || item.inputs.values().any(|t| t.tensor_shape.as_ref().map(|sh| sh.dim.len() > TFS_MAX_DIMS || sh.dim.iter().any(|d| !d.name.is_empty())).unwrap_or(false)) {
            return prost::Message::encode(&item, dst).map_err(|e| Status::internal(e.to_string()));
        }
        let d = describe(&item);
        let filters: Vec<tfs_str_t> = item.output_filter.iter().map(|f| s(f)).collect();
        let mut msg = d.msg;
        msg.n_filters = filters.len() as u32;
        let batch = tfs_batch_t { msgs: &msg, tensors: d.tensors.as_ptr(), fields: d.fields.as_ptr(), filters: filters.as_ptr(),
                                  n_msgs: 1, n_tensors: d.tensors.len() as u32, n_fields: d.fields.len() as u32, n_filters: filters.len() as u32 };
        let bound = unsafe { tfs_encode_bound(&batch, 0) } as usize;
        dst.reserve(bound);
        let (mut off, mut len, mut total) = (0u64, 0u64, 0u64);
        let ctx = self.gpu.0.lock().unwrap();
        check(unsafe { tfs_encode_host(ctx.0, &batch, dst.bytes_mut().as_mut_ptr() as *mut c_void, bound as u64, 0, &mut off, &mut len, &mut total) })?;
        unsafe { dst.advance_mut(len as usize) };
        Ok(())
    }
}

impl tokio_util::codec::Decoder for GpuDecoder {
    type Item = PredictResponse;
    type Error = Status;
    /// `src` holds exactly one complete message body (tonic's Streaming has already removed the frame prefix).
    fn decode(&mut self, src: &mut BytesMut) -> Result<Option<PredictResponse>, Status> {
        if src.len() < MIN_GPU_BYTES {                               // small: prost is faster than the descriptors
            let r = <PredictResponse as prost::Message>::decode(&src[..]).map_err(|e| Status::internal(e.to_string()));
            src.clear();
            return r.map(Some);
        }
        let (off, len) = (0u64, src.len() as u64);
        let mut dec = ptr::null_mut();
        let guard = self.gpu.0.lock().unwrap();                      // held to the end: a ctx is not re-entrant
        let ctx = guard.0;
        let st = unsafe { tfs_decode_host(ctx, src.as_ptr() as *const c_void, len, &off, &len, 1, TFS_MSG_PREDICT_RESPONSE, 0, &mut dec) };
        if st != TFS_OK && st != TFS_ERR_BATCH { return Err(check(st).unwrap_err()); }
        let mut view: tfs_decoded_view_t = unsafe { std::mem::zeroed() };
        check(unsafe { tfs_decoded_view(dec, &mut view) })?;
        let m = unsafe { *view.msgs };
        if m.status == TFS_ERR_UNSUPPORTED {                         // leave this message to prost
            unsafe { tfs_decoded_free(ctx, dec) };
            let r = <PredictResponse as prost::Message>::decode(&src[..]).map_err(|e| Status::internal(e.to_string()));
            src.clear();
            return r.map(Some);
        }
        if m.status != TFS_OK { unsafe { tfs_decoded_free(ctx, dec) }; return Err(check(m.status).unwrap_err()); }
        // the value pool: a pinned slab (the device writes it directly; 4 KiB aligned, so every 64-byte aligned region
        // of the pool is aligned for its element type - a Vec<u8> would only promise 1).  Acquired through the raw
        // call because this thread already holds the ctx lock; released below.
        let mut slab: tfs_slab_t = unsafe { std::mem::zeroed() };
        let st = unsafe { tfs_stage_acquire(ctx, view.pool_bytes.max(1), 0, &mut slab) };
        if st != TFS_OK { unsafe { tfs_decoded_free(ctx, dec) }; return Err(check(st).unwrap_err()); }
        let pool = slab.ptr as *const u8;
        let st = unsafe { tfs_decode_host_finish(ctx, dec, slab.ptr, view.pool_bytes) };
        if st != TFS_OK { unsafe { tfs_stage_release(ctx, &slab); tfs_decoded_free(ctx, dec) }; return Err(check(st).unwrap_err()); }
        let strings = unsafe { std::slice::from_raw_parts(view.strings, view.strings_bytes as usize) };
        let text = |sp: tfs_span_t| String::from_utf8_lossy(&strings[sp.off as usize..sp.off as usize + sp.len as usize]).into_owned();
        let mut outputs = HashMap::new();
        for i in 0..m.n_tensors as usize {
            let t = unsafe { *view.tensors.add(m.first_tensor as usize + i) };
            let mut tp = TensorProto::default();
            tp.dtype = t.dtype;
            tp.version_number = t.version_number;
            if t.has_shape != 0 {
                let mut sh = TensorShapeProto::default();
                sh.unknown_rank = t.unknown_rank != 0;
                for k in 0..t.n_dims as usize { sh.dim.push(crate::tensor_shape_proto::Dim { size: t.dims[k], name: String::new() }); }
                tp.tensor_shape = Some(sh);
            }
            // one exact-capacity copy per field out of the pool (slot order of tfs_slot_of_kind)
            macro_rules! take { ($slot:expr, $field:ident, $ty:ty) => {{
                let n = t.n_elems[$slot] as usize;
                if n > 0 { let p = unsafe { pool.add(t.dst_off[$slot] as usize) } as *const $ty;
                           tp.$field = unsafe { std::slice::from_raw_parts(p, n) }.to_vec(); } }} }
            take!(0, tensor_content, u8); take!(1, float_val, f32); take!(2, double_val, f64); take!(3, int_val, i32);
            // ragged slots: the flattened block ((n + 1) u64 offsets, then bytes) back into one Vec per element
            let ragged = |slot: usize| -> Vec<Vec<u8>> {
                let ns = t.n_elems[slot] as usize;
                if ns == 0 { return Vec::new(); }
                let block = unsafe { pool.add(t.dst_off[slot] as usize) };
                let off = |i: usize| unsafe { std::ptr::read(block.add(8 * i) as *const u64) } as usize;
                let bytes = unsafe { block.add(8 * (ns + 1)) };
                (0..ns).map(|i| unsafe { std::slice::from_raw_parts(bytes.add(off(i)), off(i + 1) - off(i)) }.to_vec()).collect()
            };
            tp.string_val = ragged(4);
            // message-typed fields: the library has validated every body the way the generated merge would, so these
            // decodes cannot fail; they only build the structs
            for body in ragged(12) { tp.resource_handle_val.push(prost::Message::decode(&body[..]).map_err(|e| Status::internal(e.to_string()))?); }
            for body in ragged(13) { tp.variant_val.push(prost::Message::decode(&body[..]).map_err(|e| Status::internal(e.to_string()))?); }
            take!(5, scomplex_val, f32); take!(6, int64_val, i64); take!(7, bool_val, bool); take!(8, dcomplex_val, f64);
            take!(9, half_val, i32); take!(10, uint32_val, u32); take!(11, uint64_val, u64);
            outputs.insert(text(t.key), tp);                         // a repeated key: the last entry wins, as in prost
        }
        let model_spec = if m.has_model_spec != 0 {
            Some(ModelSpec { name: text(m.model_name), signature_name: text(m.signature_name),
                             version_choice: match m.version_choice { 1 => Some(crate::model_spec::VersionChoice::Version(m.version)),
                                                                      2 => Some(crate::model_spec::VersionChoice::VersionLabel(text(m.version_label))),
                                                                      _ => None } })
        } else { None };
        unsafe { tfs_stage_release(ctx, &slab); tfs_decoded_free(ctx, dec) };
        src.clear();
        Ok(Some(PredictResponse { outputs, model_spec }))
    }
}
