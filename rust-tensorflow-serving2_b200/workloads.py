"""Synthetic inputs of the BASELINE.json configs (definitions: SURVEY.md §8d).  Host numpy only.

C1 half_plus_two request/response + the examples/prediction.rs shape; C2 ResNet-50 input 64x224x224x3
DT_FLOAT as float_val and as tensor_content; C3 BERT-base input_ids / attention_mask DT_INT64 256x512;
C4 DT_HALF / DT_BFLOAT16 half_val, 2^20 elements, from f32; C5 mixed-dtype PredictResponses.
"""
from __future__ import annotations

from typing import List

import numpy as np

from . import abi
from .messages import ClassificationRequest, Feature, Input, ModelSpec, Payload, PredictRequest, PredictResponse, Ragged, TensorProto, TensorShapeProto, ragged_block


def c1_request() -> PredictRequest:
    return PredictRequest(ModelSpec("half_plus_two", signature_name="serving_default"),
                          [("x", TensorProto(dtype=abi.DT_FLOAT, tensor_shape=TensorShapeProto([3]),
                                             float_val=np.array([1.0, 2.0, 5.0], np.float32)))])


def c1_response() -> PredictResponse:
    return PredictResponse([("y", TensorProto(dtype=abi.DT_FLOAT, tensor_shape=TensorShapeProto([3]),
                                              float_val=np.array([2.5, 3.0, 4.5], np.float32)))],
                           ModelSpec("half_plus_two", version=123, signature_name="serving_default"))


def c1_image_request(width=224, height=224, fused=True) -> PredictRequest:
    """The message predict_with_preprocessing builds (reference src/lib.rs:229-263): key "input", dims
    [1, width, height, 3], dtype 1, float_val = u8 pixels / 255 (examples/prediction.rs:43)."""
    px = np.random.default_rng(1001).integers(0, 256, width * height * 3, dtype=np.uint8)
    if fused:
        fv = Payload(px, abi.T_U8, abi.T_F32, abi.OP_DIV, 255.0)
    else:
        fv = px.astype(np.float32) / np.float32(255.0)
    return PredictRequest(ModelSpec("resnet", signature_name="serving_default"),
                          [("input", TensorProto(dtype=abi.DT_FLOAT, tensor_shape=TensorShapeProto([1, width, height, 3]), float_val=fv))])


def c2_payload(batch=64, seed=2001) -> np.ndarray:
    return np.random.default_rng(seed).standard_normal(batch * 224 * 224 * 3, dtype=np.float32)


def c2_request(x: np.ndarray, variant="float_val", batch=64) -> PredictRequest:
    shape = TensorShapeProto([batch, 224, 224, 3])
    if variant == "float_val":
        t = TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, float_val=x)
    elif variant == "tensor_content":
        t = TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, tensor_content=x)
    elif variant == "tensor_content_bf16":
        t = TensorProto(dtype=abi.DT_BFLOAT16, tensor_shape=shape, tensor_content=Payload(x, abi.T_F32, abi.T_BF16))
    else:
        raise ValueError(variant)
    return PredictRequest(ModelSpec("resnet", signature_name="serving_default"), [("input", t)])


def c3_arrays(rows=256, cols=512, negatives=0.0):
    ids = np.random.default_rng(3001).integers(0, 30522, (rows, cols), dtype=np.int64)
    lens = np.random.default_rng(3002).integers(1, cols + 1, rows)
    mask = (np.arange(cols)[None, :] < lens[:, None]).astype(np.int64)
    if negatives > 0:
        neg = np.random.default_rng(3003).random(ids.shape) < negatives
        ids = np.where(neg, np.int64(-1), ids)
    return np.ascontiguousarray(ids.reshape(-1)), np.ascontiguousarray(mask.reshape(-1))


def c3_request(ids: np.ndarray, mask: np.ndarray, rows=256, cols=512) -> PredictRequest:
    shape = TensorShapeProto([rows, cols])
    return PredictRequest(ModelSpec("bert", signature_name="serving_default"),
                          [("input_ids", TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=ids)),
                           ("attention_mask", TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=mask))])


def c4_payload(n=1 << 20, seed=4001) -> np.ndarray:
    return np.random.default_rng(seed).standard_normal(n, dtype=np.float32)


def c4_tensor(x: np.ndarray, fmt="f16") -> TensorProto:
    ty, dt = (abi.T_F16, abi.DT_HALF) if fmt == "f16" else (abi.T_BF16, abi.DT_BFLOAT16)
    return TensorProto(dtype=dt, tensor_shape=TensorShapeProto([x.size]), half_val=Payload(x, abi.T_F32, ty))


def c5_responses(n_msgs=8192, seed=5001, max_log2=18) -> List[PredictResponse]:
    """Mixed-dtype PredictResponses: k ~ U{1,2,3} outputs "out0".., dtype uniform over 7 layouts, rank
    U{1..4}, element count log-uniform in [1, 2^max_log2]; normals for FP, full-range ints with 10 % small."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n_msgs):
        k = int(rng.integers(1, 4))
        entries = []
        for j in range(k):
            n = int(2 ** rng.uniform(0, max_log2))
            rank = int(rng.integers(1, 5))
            dims = [1] * (rank - 1) + [n]
            shape = TensorShapeProto(dims)
            which = int(rng.integers(0, 7))
            small = rng.random(n) < 0.1
            if which == 0:
                t = TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, float_val=rng.standard_normal(n, dtype=np.float32))
            elif which == 1:
                t = TensorProto(dtype=abi.DT_DOUBLE, tensor_shape=shape, double_val=rng.standard_normal(n))
            elif which == 2:
                v = rng.integers(-(1 << 31), 1 << 31, n, dtype=np.int64)
                v = np.where(small, rng.integers(-127, 128, n), v).astype(np.int32)
                t = TensorProto(dtype=abi.DT_INT32, tensor_shape=shape, int_val=v)
            elif which == 3:
                v = rng.integers(-(1 << 63), (1 << 63) - 1, n, dtype=np.int64)
                v = np.where(small, rng.integers(-127, 128, n), v).astype(np.int64)
                t = TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=v)
            elif which == 4:
                h = rng.standard_normal(n, dtype=np.float32).astype(np.float16).view(np.uint16)
                t = TensorProto(dtype=abi.DT_HALF, tensor_shape=shape, half_val=h)
            elif which == 5:
                t = TensorProto(dtype=abi.DT_BOOL, tensor_shape=shape, bool_val=rng.random(n) < 0.5)
            else:
                t = TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, tensor_content=rng.standard_normal(n, dtype=np.float32))
            entries.append((f"out{j}", t))
        out.append(PredictResponse(entries, ModelSpec("mixed", version=int(rng.integers(1, 100)), signature_name="serving_default")))
    return out


def to_device(messages, kind, device):
    """Copy every payload array of the messages to `device` (torch), keeping the message structure."""
    import copy

    import torch

    def mv_raw(v):
        """-> (device tensor, TFS_T_* to declare when torch cannot express the numpy dtype, else T_NONE)"""
        a = np.ascontiguousarray(v)
        if a.dtype == np.bool_:
            return torch.from_numpy(a.view(np.uint8)).to(device).view(torch.bool), abi.T_NONE
        if a.dtype == np.uint16:  # torch's unsigned support is partial: carry the bits in the signed type
            return torch.from_numpy(a.view(np.int16)).to(device), abi.T_U16
        if a.dtype == np.uint32:
            return torch.from_numpy(a.view(np.int32)).to(device), abi.T_U32
        if a.dtype == np.uint64:
            return torch.from_numpy(a.view(np.int64)).to(device), abi.T_U64
        return torch.from_numpy(a).to(device), abi.T_NONE

    def mv(v):
        if v is None:
            return None
        if isinstance(v, tuple):  # several arrays concatenated into one field
            return tuple(mv(x) for x in v)
        if isinstance(v, Ragged):
            return Ragged(torch.as_tensor(np.ascontiguousarray(v.block)).to(device), v.n_elems, v.total_bytes)
        if isinstance(v, list):  # string_val given as a list of bytes-like elements
            blk = ragged_block(v)
            return Ragged(torch.from_numpy(blk).to(device), len(v), blk.size - 8 * (len(v) + 1))
        if isinstance(v, Payload):
            t, implied = mv_raw(v.data)
            return Payload(t, v.src_type if v.src_type != abi.T_NONE else implied, v.dst_type, v.op, v.op_param)
        t, implied = mv_raw(v)
        return t if implied == abi.T_NONE else Payload(t, implied)

    def mv_tensor(t: TensorProto) -> TensorProto:
        t2 = copy.copy(t)
        for k in abi.SLOT_KINDS:
            name = abi.KIND_FIELD_NAMES[k]
            setattr(t2, name, mv(getattr(t, name)))
        return t2

    def mv_input(inp: Input) -> Input:
        return Input([[(k, Feature(mv(f.bytes_list), mv(f.float_list), mv(f.int64_list))) for k, f in ex] for ex in inp.examples])

    out = []
    for m in messages:
        if kind == abi.MSG_TENSOR_PROTO:
            out.append(mv_tensor(m))
        elif kind == abi.MSG_PREDICT_REQUEST:
            out.append(PredictRequest(m.model_spec, [(k, mv_tensor(t)) for k, t in m.inputs], list(m.output_filter)))
        elif kind == abi.MSG_INPUT:
            out.append(mv_input(m))
        elif kind == abi.MSG_CLASSIFY_REQUEST:
            out.append(ClassificationRequest(m.model_spec, mv_input(m.input)))
        else:
            out.append(PredictResponse([(k, mv_tensor(t)) for k, t in m.outputs], m.model_spec))
    return out
