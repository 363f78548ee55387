"""The headline e2e step (C2, pinned buffers, TFS_F_ALIGN_PAYLOAD) with encode_host and decode_host_into timed apart."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tfs_b200 as tfs  # noqa: E402
from tfs_b200 import abi  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 32
ELEMS = 64 * 224 * 224 * 3
codec = tfs.GpuCodec(tfs.GpuContext(0))
enc, dec = codec.encoder(), codec.decoder()
kind, flags = abi.MSG_PREDICT_REQUEST, abi.F_ALIGN_PAYLOAD
h_src = torch.randn(R, ELEMS).pin_memory()
shape = tfs.TensorShapeProto([64, 224, 224, 3])
msgs = [tfs.PredictRequest(tfs.ModelSpec("resnet50", signature_name="serving_default"),
                           [("input", tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, float_val=h_src[i].numpy()))]) for i in range(R)]
built = tfs.build_batch(msgs, kind)
h_wire = torch.empty(enc.bound(built, flags) + 4096, dtype=torch.uint8).pin_memory().numpy()
h_pool = torch.empty(R * (ELEMS * 4 + 4096), dtype=torch.uint8).pin_memory().numpy()
for rep in range(5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    eb = enc.encode_host(built, kind, flags, dst=h_wire)
    t1 = time.perf_counter()
    d = dec.decode_host_into(h_wire[: eb.total], eb.offsets_np if hasattr(eb, "offsets_np") else eb.offsets, eb.lengths_np if hasattr(eb, "lengths_np") else eb.lengths, kind, h_pool)
    t2 = time.perf_counter()
    d.close()
    print(f"encode_host {1e3 * (t1 - t0):.2f} ms ({eb.total / (t1 - t0) / 1e9:.1f} GB/s), decode_host_into {1e3 * (t2 - t1):.2f} ms ({eb.total / (t2 - t1) / 1e9:.1f} GB/s)")
