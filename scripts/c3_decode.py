"""C3 (256 BERT requests, int64_val) encode + decode a few times: plain for timings, under ncu for the launch list."""
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
R = 256
g = torch.Generator(device=dev)
g.manual_seed(3001)
ids = torch.randint(0, 30522, (R, 256 * 512), device=dev, dtype=torch.int64, generator=g)
lens = torch.randint(1, 513, (R, 256, 1), device=dev, generator=g)
mask = (torch.arange(512, device=dev)[None, None, :] < lens).to(torch.int64).reshape(R, -1).contiguous()
shape = tfs.TensorShapeProto([256, 512])
msgs = [tfs.PredictRequest(tfs.ModelSpec("bert", signature_name="serving_default"),
                           [("input_ids", tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=ids[i])),
                            ("attention_mask", tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=mask[i]))]) for i in range(R)]
built = tfs.build_batch(msgs, abi.MSG_PREDICT_REQUEST)
wire = torch.empty(enc.bound(built, 0), dtype=torch.uint8, device=dev)
eb = enc.encode(built, abi.MSG_PREDICT_REQUEST, 0, dst=wire)
d = dec.scan(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST)
pool = torch.empty(d.pool_bytes + 64, dtype=torch.uint8, device=dev)
d.close()
for rep in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    eb = enc.encode(built, abi.MSG_PREDICT_REQUEST, 0, dst=wire)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    dec.decode_into(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST, pool).close()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"C3 encode {1e3 * (t1 - t0):.3f} ms, decode_into {1e3 * (t2 - t1):.3f} ms (wire {eb.total / 1e6:.1f} MB)")
