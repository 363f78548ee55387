"""C2 requests through tfs_encode_host / tfs_decode_host_into with PAGEABLE buffers (what prost's Vecs and tonic's BytesMut
are): the pinned ring + helper threads path of csrc/tfs_host.cu.  TFS_TRACE=1 prints the phases; TFS_STAGE_THREADS=n."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfs_b200 as tfs  # noqa: E402
from tfs_b200 import abi  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 8
ELEMS = 64 * 224 * 224 * 3
codec = tfs.GpuCodec(tfs.GpuContext(0))
enc, dec = codec.encoder(), codec.decoder()
kind = abi.MSG_PREDICT_REQUEST
src = np.random.default_rng(1).standard_normal((R, ELEMS)).astype(np.float32)
shape = tfs.TensorShapeProto([64, 224, 224, 3])
msgs = [tfs.PredictRequest(tfs.ModelSpec("resnet50", signature_name="serving_default"),
                           [("input", tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, float_val=src[i]))]) for i in range(R)]
built = tfs.build_batch(msgs, kind)
wire = np.empty(enc.bound(built, 0) + 4096, np.uint8)
pool = np.empty(R * (ELEMS * 4 + 4096), np.uint8)
for rep in range(4):
    t0 = time.perf_counter()
    eb = enc.encode_host(built, kind, 0, dst=wire)
    t1 = time.perf_counter()
    d = dec.decode_host_into(wire[: eb.total], eb.offsets_np, eb.lengths_np, kind, pool)
    t2 = time.perf_counter()
    d.close()
    print(f"pageable: encode_host {1e3 * (t1 - t0):.1f} ms ({eb.total / (t1 - t0) / 1e9:.1f} GB/s), "
          f"decode_host_into {1e3 * (t2 - t1):.1f} ms ({eb.total / (t2 - t1) / 1e9:.1f} GB/s)", flush=True)
