"""BASELINE config 5 from ONE process: 8 192 mixed-dtype PredictResponses in pinned host memory, decoded by the engine
(one lane = context + stream + worker thread per visible GPU), the batch cut into contiguous runs of about equal bytes.
Prints the time of the sharded call for 1 lane and for all lanes."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import tfs_b200 as tfs  # noqa: E402
from tfs_b200 import abi, workloads  # noqa: E402
import oracle_bind as ob  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
msgs = workloads.c5_responses(N, seed=5001, max_log2=18)
wire, offs, lens = ob.encode(tfs.build_batch(msgs, abi.MSG_PREDICT_RESPONSE))
pw = torch.empty(wire.size, dtype=torch.uint8).pin_memory()
pw.numpy()[:] = wire
wire_p = pw.numpy()
print(f"{N} responses, wire {wire.size / 1e6:.1f} MB, {torch.cuda.device_count()} GPUs visible")
for devs in ([0], list(range(torch.cuda.device_count()))):
    eng = tfs.Engine(devs, lanes_per_device=1, pinned_bytes_per_lane=0)
    keep = [torch.empty(int(wire.size * 1.6 / len(devs)) + (64 << 20), dtype=torch.uint8).pin_memory() for _ in devs]
    pools = [k.numpy() for k in keep]
    best = None
    for rep in range(4):
        t0 = time.perf_counter()
        st, res, first = eng.decode_host_sharded(wire_p, offs, lens, abi.MSG_PREDICT_RESPONSE, pools)
        dt = time.perf_counter() - t0
        n_msgs = 0
        for r in res:
            if r is not None:
                n_msgs += r.view.n_msgs
                r.close()
        assert st == 0 and n_msgs == N
        best = dt if best is None or (rep and dt < best) else best
    print(f"{len(devs)} lane(s): sharded decode (pinned host wire -> pinned host values) {1e3 * best:.2f} ms = {wire.size / best / 1e9:.1f} GB/s of wire")
    eng.close()
    del keep, pools
