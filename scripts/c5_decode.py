"""C5 (1024 mixed-dtype PredictResponses) decode and the ClassificationRequest encode, a few repetitions each: run plain for
timings, or under `ncu --metrics gpu__time_duration.sum` for the per-kernel launch list."""
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
msgs = workloads.to_device(workloads.c5_responses(1024, seed=5001), abi.MSG_PREDICT_RESPONSE, dev)
eb = enc.encode(msgs, abi.MSG_PREDICT_RESPONSE)
d = dec.scan(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_RESPONSE)
pool = torch.empty(d.pool_bytes + 64, dtype=torch.uint8, device=dev)
d.close()
for rep in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    dec.decode_into(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_RESPONSE, pool).close()
    torch.cuda.synchronize()
    print(f"C5 decode_into {1e3 * (time.perf_counter() - t0):.3f} ms  (wire {eb.total / 1e6:.1f} MB)")

codec.ctx.prof_enable(True)
for k in range(8):
    codec.ctx.prof_read(k, reset=True)
for rep in range(5):
    dec.decode_into(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_RESPONSE, pool).close()
codec.ctx.sync()
names = ["k_copy", "k_varint_size", "k_varint_scatter", "k_varint_count", "k_varint_gather", "k_copy_tma", "walk/plan", "k_frame_small"]
print("C5 decode, event-timed kernels (avg ms per launch):", {names[k]: round(ms / n, 4) for k in range(8) for n, ms in [codec.ctx.prof_read(k, reset=True)] if n})
codec.ctx.prof_enable(False)
rng = np.random.default_rng(1)
reqs = []
for r in range(64):
    exs = [[("ids", tfs.Feature(int64_list=rng.integers(0, 30522, 128, dtype=np.int64))), ("weights", tfs.Feature(float_list=rng.standard_normal(128).astype(np.float32))),
            ("tag", tfs.Feature(bytes_list=[b"label-%d" % (r % 7)]))] for e in range(64)]
    reqs.append(tfs.ClassificationRequest(tfs.ModelSpec("wide_deep", signature_name="classify"), tfs.Input(exs)))
reqs = workloads.to_device(reqs, abi.MSG_CLASSIFY_REQUEST, dev)
built = tfs.build_batch(reqs, abi.MSG_CLASSIFY_REQUEST)
wire = torch.empty(enc.bound(built, 0), dtype=torch.uint8, device=dev)
for rep in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    eb2 = enc.encode(built, abi.MSG_CLASSIFY_REQUEST, 0, dst=wire)
    torch.cuda.synchronize()
    print(f"ClassificationRequest encode {1e3 * (time.perf_counter() - t0):.3f} ms  (wire {eb2.total / 1e6:.2f} MB)")
