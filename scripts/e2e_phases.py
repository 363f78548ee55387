"""Where the host-buffer (e2e) time goes for a varint-coded batch: C3 (BERT ids, int64_val) with pinned buffers.

Prints, per phase, the library's host entry points beside a hand-rolled "one big copy each way + device entry point"
variant, which is what the PCIe link allows with today's kernels.
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import tfs_b200 as tfs  # noqa: E402
from tfs_b200 import abi  # noqa: E402
from test_gpu_parity import pinned_like  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 32
REPS = 10
codec = tfs.GpuCodec(tfs.GpuContext(0))
enc, dec = codec.encoder(), codec.decoder()
dev = torch.device("cuda:0")
rng = np.random.default_rng(3001)
ids = rng.integers(0, 30522, (R, 256 * 512), dtype=np.int64)
lens = rng.integers(1, 513, (R, 256, 1))
mask = (np.arange(512)[None, None, :] < lens).astype(np.int64).reshape(R, -1)
shape = tfs.TensorShapeProto([256, 512])
KIND = abi.MSG_PREDICT_REQUEST


def c3(a, b):
    return tfs.PredictRequest(tfs.ModelSpec("bert", signature_name="serving_default"),
                              [("input_ids", tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=a)),
                               ("attention_mask", tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=b))])


def pin(n):
    t = torch.empty(n, dtype=torch.uint8).pin_memory()
    return t, t.numpy()


def timed(fn, reps=REPS):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / reps


host_msgs = [c3(ids[i], mask[i]) for i in range(R)]
pmsgs, keep = pinned_like(host_msgs, KIND)
built = tfs.build_batch(pmsgs, KIND)
cap = enc.bound(built, 0) + 4096
kw, wbuf = pin(cap)
eb = enc.encode_host(built, KIND, 0, dst=wbuf)
total = eb.total
d = dec.decode_host(wbuf[:total], eb.offsets, eb.lengths, KIND, 0)
pb = d.pool_bytes + 64
d.close()
kp, pbuf = pin(pb)
values_mb = 2 * R * 256 * 512 * 8 / 1e6
print(f"C3 x {R}: values {values_mb:.1f} MB, wire {total / 1e6:.1f} MB, pool {pb / 1e6:.1f} MB")

t_enc = timed(lambda: enc.encode_host(built, KIND, 0, dst=wbuf))
t_dec = timed(lambda: dec.decode_host_into(wbuf[:total], eb.offsets, eb.lengths, KIND, pbuf, 0).close())
print(f"library   : encode_host {t_enc:.3f} ms, decode_host_into {t_dec:.3f} ms -> {2 * total / (t_enc + t_dec) / 1e6:.2f} GB/s of wire")

# hand-rolled: one H2D of all values, device encode, one D2H of the wire; one H2D of the wire, device decode, one D2H
allv_t, allv = pin(2 * R * 256 * 512 * 8)
av = allv.view(np.int64).reshape(2, R, -1)
av[0], av[1] = ids, mask
dvals = torch.empty(2 * R * 256 * 512, dtype=torch.int64, device=dev)
dmsgs = [c3(dvals[i * 256 * 512:(i + 1) * 256 * 512], dvals[(R + i) * 256 * 512:(R + i + 1) * 256 * 512]) for i in range(R)]
dbuilt = tfs.build_batch(dmsgs, KIND)
dwire = torch.empty(cap, dtype=torch.uint8, device=dev)
dpool = torch.empty(pb, dtype=torch.uint8, device=dev)
wire_t = kw


def hand_enc():
    dvals.view(torch.uint8).copy_(allv_t, non_blocking=True)
    e = enc.encode(dbuilt, KIND, 0, dst=dwire)
    wire_t[:total].copy_(dwire[:total], non_blocking=True)
    torch.cuda.synchronize()
    return e


def hand_dec():
    dwire[:total].copy_(wire_t[:total], non_blocking=True)
    dec.decode_into(dwire[:total], eb.offsets, eb.lengths, KIND, dpool).close()
    kp.copy_(dpool, non_blocking=True)
    torch.cuda.synchronize()


# GpuContext follows torch's current stream (the legacy default stream here), so torch's copies order with the library
t_henc = timed(hand_enc)
t_hdec = timed(hand_dec)
print(f"hand-rolled: encode {t_henc:.3f} ms, decode {t_hdec:.3f} ms -> {2 * total / (t_henc + t_hdec) / 1e6:.2f} GB/s of wire")
t_h2d = timed(lambda: dvals.view(torch.uint8).copy_(allv_t, non_blocking=True))
t_d2h = timed(lambda: kp.copy_(dpool, non_blocking=True))
print(f"copies alone: values H2D {t_h2d:.3f} ms ({values_mb / t_h2d:.1f} GB/s), pool D2H {t_d2h:.3f} ms ({pb / 1e6 / t_d2h:.1f} GB/s)")
