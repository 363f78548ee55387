"""DT_STRING tensors with many short elements (64 requests x 512 tokens): decode timings / ncu launch list."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import tfs_b200 as tfs  # noqa: E402
from tfs_b200 import abi, workloads  # noqa: E402

codec = tfs.GpuCodec(tfs.GpuContext(0))
enc, dec = codec.encoder(), codec.decoder()
dev = torch.device("cuda:0")
rng = np.random.default_rng(3)
msgs = [tfs.PredictRequest(tfs.ModelSpec("strings", signature_name="serving_default"),
                           [("examples", tfs.TensorProto(dtype=7, tensor_shape=tfs.TensorShapeProto([512]),
                                                         string_val=[rng.integers(0, 256, int(rng.integers(0, 41)), dtype=np.uint8).tobytes() for _ in range(512)]))])
        for _ in range(64)]
built = tfs.build_batch(workloads.to_device(msgs, abi.MSG_PREDICT_REQUEST, dev), abi.MSG_PREDICT_REQUEST)
eb = enc.encode(built, abi.MSG_PREDICT_REQUEST)
d = dec.scan(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST)
pool = torch.empty(d.pool_bytes + 64, dtype=torch.uint8, device=dev)
d.close()
for rep in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    dec.decode_into(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST, pool).close()
    torch.cuda.synchronize()
    print(f"tokens decode_into {1e3 * (time.perf_counter() - t0):.3f} ms")
