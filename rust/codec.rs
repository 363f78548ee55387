//! `GpuCodec`: the tonic codec `predict` uses instead of the generated stub's `ProstCodec` (reference src/lib.rs:265).
//! UNCOMPILED here (no rustc).  Targets tonic 0.1.0 (`tonic::codec::Codec`, `tokio_util::codec::{Encoder, Decoder}`),
//! prost 0.6.0 generated structs (`PredictRequest`, `PredictResponse`, `TensorProto`) and bytes 0.5.3.
use std::{collections::HashMap, ffi::CStr, os::raw::c_void, ptr, sync::{Arc, Mutex}};

use bytes::{BufMut, BytesMut};
use tonic::{codec::Codec, Status};

use crate::ffi::*;
use crate::{ModelSpec, PredictRequest, PredictResponse, TensorProto, TensorShapeProto};

/// One context per GPU, shared by the clones of `TensorflowServing` (src/lib.rs:148-156); a ctx is not re-entrant.
#[derive(Clone)]
pub struct GpuHandle(pub Arc<Mutex<*mut tfs_ctx_t>>);
unsafe impl Send for GpuHandle {}
unsafe impl Sync for GpuHandle {}

impl GpuHandle {
    pub fn new(device: i32) -> Result<Self, Status> {
        let mut ctx = ptr::null_mut();
        check(unsafe { tfs_ctx_create(device, 64 << 20, &mut ctx) })?; // no device => no codec: there is no CPU fallback
        Ok(GpuHandle(Arc::new(Mutex::new(ctx))))
    }
}

fn check(status: i32) -> Result<(), Status> {
    if status == TFS_OK { return Ok(()); }
    // same text class tonic's ProstDecoder derives from prost::DecodeError (Status::internal(err.to_string()))
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

/// Plain-data picture of one PredictRequest: the descriptors point into the message's Vecs; only string_val
/// (`Vec<Vec<u8>>`) is flattened into a TFS_T_RAGGED block ((n + 1) u64 offsets, then the bytes) that `ragged` keeps alive.
struct Described { msg: tfs_msg_desc_t, tensors: Vec<tfs_tensor_desc_t>, fields: Vec<tfs_field_desc_t>, ragged: Vec<Vec<u64>> }

fn s(x: &str) -> tfs_str_t { tfs_str_t { ptr: x.as_ptr() as *const _, len: x.len() as u32, _pad: 0 } }

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
        if !t.string_val.is_empty() {
            let n = t.string_val.len();
            let total: usize = t.string_val.iter().map(|e| e.len()).sum();
            let mut block = vec![0u64; n + 1 + (total + 7) / 8];         // u64 storage: the block is 8-byte aligned
            let mut at = 0usize;
            for (i, e) in t.string_val.iter().enumerate() {
                block[i] = at as u64;
                unsafe { std::ptr::copy_nonoverlapping(e.as_ptr(), (block.as_mut_ptr().add(n + 1) as *mut u8).add(at), e.len()) };
                at += e.len();
            }
            block[n] = at as u64;
            fields.push(tfs_field_desc_t { kind: TFS_KIND_STRING_VAL, src_type: TFS_T_RAGGED, dst_type: 0, op: 0, op_param: 0.0, _pad: 0,
                                           n_elems: n as u64, data: block.as_ptr() as *const _, aux: total as u64 });
            ragged.push(block);                                           // moving the Vec does not move its heap buffer
        }
        push(TFS_KIND_SCOMPLEX_VAL, TFS_T_F32, t.scomplex_val.len(), t.scomplex_val.as_ptr() as *const _);
        push(TFS_KIND_INT64_VAL, TFS_T_I64, t.int64_val.len(), t.int64_val.as_ptr() as *const _);
        push(TFS_KIND_BOOL_VAL, TFS_T_BOOL, t.bool_val.len(), t.bool_val.as_ptr() as *const _);
        push(TFS_KIND_DCOMPLEX_VAL, TFS_T_F64, t.dcomplex_val.len(), t.dcomplex_val.as_ptr() as *const _);
        push(TFS_KIND_HALF_VAL, TFS_T_I32, t.half_val.len(), t.half_val.as_ptr() as *const _);
        push(TFS_KIND_UINT32_VAL, TFS_T_U32, t.uint32_val.len(), t.uint32_val.as_ptr() as *const _);
        push(TFS_KIND_UINT64_VAL, TFS_T_U64, t.uint64_val.len(), t.uint64_val.as_ptr() as *const _);
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
        // messages that use what the GPU path leaves to prost (variant_val, resource handles, rank > 8, named dims)
        // go through prost unchanged
        if item.inputs.values().any(|t| !t.variant_val.is_empty() || !t.resource_handle_val.is_empty()
            || t.tensor_shape.as_ref().map(|sh| sh.dim.len() > TFS_MAX_DIMS || sh.dim.iter().any(|d| !d.name.is_empty())).unwrap_or(false)) {
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
        check(unsafe { tfs_encode_host(*ctx, &batch, dst.bytes_mut().as_mut_ptr() as *mut c_void, bound as u64, 0, &mut off, &mut len, &mut total) })?;
        unsafe { dst.advance_mut(len as usize) };
        Ok(())
    }
}

impl tokio_util::codec::Decoder for GpuDecoder {
    type Item = PredictResponse;
    type Error = Status;
    /// `src` holds exactly one complete message body (tonic's Streaming has already removed the frame prefix).
    fn decode(&mut self, src: &mut BytesMut) -> Result<Option<PredictResponse>, Status> {
        let ctx = self.gpu.0.lock().unwrap();
        let (off, len) = (0u64, src.len() as u64);
        let mut dec = ptr::null_mut();
        let st = unsafe { tfs_decode_host(*ctx, src.as_ptr() as *const c_void, len, &off, &len, 1, TFS_MSG_PREDICT_RESPONSE, 0, &mut dec) };
        if st != TFS_OK && st != TFS_ERR_BATCH { return Err(check(st).unwrap_err()); }
        let mut view: tfs_decoded_view_t = unsafe { std::mem::zeroed() };
        check(unsafe { tfs_decoded_view(dec, &mut view) })?;
        let m = unsafe { *view.msgs };
        if m.status == TFS_ERR_UNSUPPORTED {                         // leave this message to prost
            unsafe { tfs_decoded_free(*ctx, dec) };
            let r = <PredictResponse as prost::Message>::decode(&src[..]).map_err(|e| Status::internal(e.to_string()));
            src.clear();
            return r.map(Some);
        }
        if m.status != TFS_OK { unsafe { tfs_decoded_free(*ctx, dec) }; return Err(check(m.status).unwrap_err()); }
        let mut pool = vec![0u8; view.pool_bytes as usize];
        check(unsafe { tfs_decode_host_finish(*ctx, dec, pool.as_mut_ptr() as *mut c_void, pool.len() as u64) })?;
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
                if n > 0 { let p = unsafe { pool.as_ptr().add(t.dst_off[$slot] as usize) } as *const $ty;
                           tp.$field = unsafe { std::slice::from_raw_parts(p, n) }.to_vec(); } }} }
            take!(0, tensor_content, u8); take!(1, float_val, f32); take!(2, double_val, f64); take!(3, int_val, i32);
            let ns = t.n_elems[4] as usize;                          // string_val: the flattened block back into Vec<Vec<u8>>
            if ns > 0 {
                let block = unsafe { pool.as_ptr().add(t.dst_off[4] as usize) };
                let off = |i: usize| unsafe { std::ptr::read_unaligned(block.add(8 * i) as *const u64) } as usize;
                let bytes = unsafe { block.add(8 * (ns + 1)) };
                tp.string_val = (0..ns).map(|i| unsafe { std::slice::from_raw_parts(bytes.add(off(i)), off(i + 1) - off(i)) }.to_vec()).collect();
            }
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
        unsafe { tfs_decoded_free(*ctx, dec) };
        src.clear();
        Ok(Some(PredictResponse { outputs, model_spec }))
    }
}
