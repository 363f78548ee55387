"""B200-native TensorProto / PredictRequest / PredictResponse wire codec (the one hot path of
simonrw/rust-tensorflow-serving2, reference src/lib.rs:229-266), behind the C ABI of include/tfs_gpu.h.

Python here is only the harness above the C ABI (the role the Rust shim plays in the reference crate):
message objects, descriptor building and a ctypes binding.  All compute is in csrc/ (CUDA, sm_100a).
There is no CPU fallback: using the codec without the built library or without a GPU raises.
"""
from . import abi, messages, workloads, _lib, codec, sharding  # noqa: F401  (all submodules, so aliases of the package see them)
from ._lib import TfsError  # noqa: F401
from .codec import Coalescer, Engine, GpuCodec, GpuContext, GpuDecoder, GpuEncoder, Slab  # noqa: F401
from .messages import (  # noqa: F401
    AnyProto, ClassificationRequest, Feature, GetModelMetadataRequest, GetModelMetadataResponse, Input, ModelSpec, Payload, PredictRequest, PredictResponse, Ragged, TensorProto, TensorShapeProto, build_batch, ragged_block,
    messages_from_view, replicate_batch, BuiltBatch, ResourceHandleProto, VariantTensorDataProto, resource_handle_body, variant_body,
    parse_resource_handle, parse_variant,
)

__all__ = ["abi", "workloads", "codec", "ModelSpec", "Payload", "PredictRequest", "PredictResponse", "TensorProto",
           "TensorShapeProto", "build_batch", "messages_from_view", "GpuCodec", "GpuContext", "GpuEncoder", "GpuDecoder",
           "TfsError"]
