"""Device-resident timing of the staged cast kernel (k_copy_cast): C2 payloads narrowed f32 -> bf16 into tensor_content, widened
back on decode, and u8 -> f32 / 255 into float_val.  Run plain for timings, or under ncu for a capture of k_copy_cast."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import tfs_b200 as tfs
from tfs_b200 import abi

R = int(os.environ.get("R", "16"))
N = 64 * 224 * 224 * 3
dev = torch.device("cuda:0")
codec = tfs.GpuCodec(tfs.GpuContext(0)); ctx, enc, dec = codec.ctx, codec.encoder(), codec.decoder()
g = torch.Generator(device=dev); g.manual_seed(7)
src = torch.randn(R, N, device=dev, generator=g)
px = torch.randint(0, 256, (R, N), device=dev, dtype=torch.uint8, generator=g)
shape = tfs.TensorShapeProto([64, 224, 224, 3])
spec = tfs.ModelSpec("resnet", signature_name="serving_default")
cases = {
    "f32->bf16 content": ([tfs.PredictRequest(spec, [("input", tfs.TensorProto(dtype=14, tensor_shape=shape, tensor_content=tfs.Payload(src[i], abi.T_F32, abi.T_BF16)))]) for i in range(R)], R * N * 6, abi.F_CONTENT_HALF_TO_F32),
    "u8->f32 /255": ([tfs.PredictRequest(spec, [("input", tfs.TensorProto(dtype=1, tensor_shape=shape, float_val=tfs.Payload(px[i], abi.T_U8, abi.T_F32, abi.OP_DIV, 255.0)))]) for i in range(R)], R * N * 5, 0),
}
for name, (msgs, alg, dflags) in cases.items():
    for eflags, label in ((abi.F_ALIGN_PAYLOAD, "aligned"), (0, "back to back")):
        built = tfs.build_batch(msgs, abi.MSG_PREDICT_REQUEST)
        wire = torch.empty(enc.bound(built, eflags), dtype=torch.uint8, device=dev)
        eb = enc.encode(built, abi.MSG_PREDICT_REQUEST, eflags, dst=wire)
        d = dec.scan(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST, dflags); pool = torch.empty(d.pool_bytes + 64, dtype=torch.uint8, device=dev); d.close()
        for _ in range(2):
            enc.encode(built, abi.MSG_PREDICT_REQUEST, eflags, dst=wire); dec.decode_into(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST, pool, dflags).close()
        ctx.prof_enable(True); ctx.prof_read(abi.PK_COPY_CAST, reset=True)
        for _ in range(5):
            enc.encode(built, abi.MSG_PREDICT_REQUEST, eflags, dst=wire)
        n, ms = ctx.prof_read(abi.PK_COPY_CAST, reset=True)
        print(f"{name:18s} {label:13s} encode  k_copy_cast {ms / n * 1e3:7.1f} us/launch  {alg / (ms / n * 1e-3) / 1e9:7.1f} GB/s algorithmic")
        if dflags:
            for _ in range(5):
                dec.decode_into(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST, pool, dflags).close()
            n, ms = ctx.prof_read(abi.PK_COPY_CAST, reset=True)
            print(f"{name:18s} {label:13s} decode  k_copy_cast {ms / n * 1e3:7.1f} us/launch  {alg / (ms / n * 1e-3) / 1e9:7.1f} GB/s algorithmic")
        ctx.prof_enable(False)
