"""C1 (63-byte half_plus_two request) x N (default 16384; argv[1]) in one batch; TFS_TRACE=1 prints the host phases: encode + decode, plain timings or an ncu launch list."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfs_b200 as tfs  # noqa: E402
from tfs_b200 import abi  # noqa: E402

codec = tfs.GpuCodec(tfs.GpuContext(0))
enc, dec = codec.encoder(), codec.decoder()
dev = torch.device("cuda:0")
x3 = torch.tensor([1.0, 2.0, 5.0], device=dev)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
one = tfs.PredictRequest(tfs.ModelSpec("half_plus_two", signature_name="serving_default"),
                         [("x", tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=tfs.TensorShapeProto([3]), float_val=x3))])
built = tfs.replicate_batch(tfs.build_batch([one], abi.MSG_PREDICT_REQUEST), N)
wire = torch.empty(enc.bound(built, 0), dtype=torch.uint8, device=dev)
eb = enc.encode(built, abi.MSG_PREDICT_REQUEST, 0, dst=wire)
d = dec.scan(eb.wire, eb.offsets_np, eb.lengths_np, abi.MSG_PREDICT_REQUEST)
pool = torch.empty(d.pool_bytes + 64, dtype=torch.uint8, device=dev)
d.close()
for rep in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    eb = enc.encode(built, abi.MSG_PREDICT_REQUEST, 0, dst=wire)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    dd = dec.decode_into(eb.wire, eb.offsets_np, eb.lengths_np, abi.MSG_PREDICT_REQUEST, pool)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    dd.close()
    print(f"C1 x {N}: encode {1e3 * (t1 - t0):.3f} ms, decode_into {1e3 * (t2 - t1):.3f} ms")
