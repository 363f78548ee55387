"""Device-resident timing of the varint paths (C3 BERT int64_val, C4 half_val from f32) — development aid.
Prints per-kernel event times (tfs_prof_*) and algorithmic GB/s."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import torch
import tfs_b200 as tfs
from tfs_b200 import abi, workloads

codec = tfs.GpuCodec(tfs.GpuContext(0))
ctx, enc, dec = codec.ctx, codec.encoder(), codec.decoder()
dev = torch.device("cuda:0")
NAMES = dict(abi.PK_NAMES)

def run(name, msgs, kind, in_bytes, reps=10, flags=0, dflags=0):
    built = tfs.build_batch(msgs, kind)
    wire = torch.empty(enc.bound(built, flags), dtype=torch.uint8, device=dev)
    eb = enc.encode(built, kind, flags, dst=wire)
    W = sum(eb.lengths)
    d = dec.scan(eb.wire, eb.offsets, eb.lengths, kind, dflags)
    pool = torch.empty(d.pool_bytes + 64, dtype=torch.uint8, device=dev)
    d.close()
    for _ in range(3):
        eb = enc.encode(built, kind, flags, dst=wire); dd = dec.decode_into(eb.wire, eb.offsets, eb.lengths, kind, pool, dflags); dd.close()
    ctx.prof_enable(True)
    for k in range(abi.PK_COUNT): ctx.prof_read(k, reset=True)
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    te = td = 0.0
    for _ in range(reps):
        e[0].record(); eb = enc.encode(built, kind, flags, dst=wire); e[1].record()
        dd = dec.decode_into(eb.wire, eb.offsets, eb.lengths, kind, pool, dflags); e[2].record(); torch.cuda.synchronize()
        te += e[0].elapsed_time(e[1]); td += e[1].elapsed_time(e[2]); dd.close()
    ctx.prof_enable(False)
    out_bytes = dd.pool_bytes
    print(f"== {name}: {len(msgs)} msgs, input {in_bytes/1e6:.1f} MB, wire {W/1e6:.1f} MB, decoded {out_bytes/1e6:.1f} MB")
    print(f"   encode call {te/reps*1e3:8.1f} us -> {W/(te/reps*1e-3)/1e9:7.1f} GB/s wire, {(in_bytes+W)/(te/reps*1e-3)/1e9:7.1f} GB/s algorithmic")
    print(f"   decode call {td/reps*1e3:8.1f} us -> {W/(td/reps*1e-3)/1e9:7.1f} GB/s wire, {(out_bytes+W)/(td/reps*1e-3)/1e9:7.1f} GB/s algorithmic")
    for k in range(abi.PK_COUNT):
        n, ms = ctx.prof_read(k, reset=True)
        if n:
            alg = {1: in_bytes, 2: in_bytes + W, 3: W, 4: W + out_bytes}.get(k)
            extra = f"  {alg/(ms/n*1e-3)/1e9:7.1f} GB/s algorithmic" if alg else ""
            print(f"   {NAMES[k]:15s} {ms/n*1e3:8.1f} us/launch x{n//reps}/step{extra}")

R = int(os.environ.get("R3", "256"))
g = torch.Generator(device=dev); g.manual_seed(3001)
ids = torch.randint(0, 30522, (R, 256 * 512), device=dev, dtype=torch.int64, generator=g)
lens = torch.randint(1, 513, (R, 256, 1), device=dev, generator=g)
mask = (torch.arange(512, device=dev)[None, None, :] < lens).to(torch.int64).reshape(R, -1).contiguous()
shape = tfs.TensorShapeProto([256, 512])
msgs = [tfs.PredictRequest(tfs.ModelSpec("bert", signature_name="serving_default"),
        [("input_ids", tfs.TensorProto(dtype=9, tensor_shape=shape, int64_val=ids[i])),
         ("attention_mask", tfs.TensorProto(dtype=9, tensor_shape=shape, int64_val=mask[i]))]) for i in range(R)]
run(f"C3 BERT int64_val x{R}", msgs, abi.MSG_PREDICT_REQUEST, 2 * R * 256 * 512 * 8)
del ids, mask, msgs
R4 = int(os.environ.get("R4", "64"))
x = torch.randn(R4, 1 << 20, device=dev, generator=g)
for fmt, ty, dt in (("f16", abi.T_F16, 19), ("bf16", abi.T_BF16, 14)):
    msgs = [tfs.TensorProto(dtype=dt, tensor_shape=tfs.TensorShapeProto([1 << 20]), half_val=tfs.Payload(x[i], abi.T_F32, ty)) for i in range(R4)]
    run(f"C4 half_val {fmt} from f32 x{R4}, decode to f32", msgs, abi.MSG_TENSOR_PROTO, R4 * (1 << 22), dflags=abi.F_HALF_TO_F32)
