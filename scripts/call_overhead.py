"""Host wall-clock of the synchronous API calls at several batch sizes (development aid)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import tfs_b200 as tfs
from tfs_b200 import abi
C2 = 64 * 224 * 224 * 3
codec = tfs.GpuCodec(tfs.GpuContext(0))
enc, dec = codec.encoder(), codec.decoder()
dev = torch.device("cuda:0")
for R in (1, 4, 16):
    src = torch.randn(R, C2, device=dev)
    msgs = [tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"),
            [("input", tfs.TensorProto(dtype=1, tensor_shape=tfs.TensorShapeProto([64, 224, 224, 3]), float_val=src[i]))]) for i in range(R)]
    built = tfs.build_batch(msgs, abi.MSG_PREDICT_REQUEST)
    wire = torch.empty(enc.bound(built, abi.F_ALIGN_PAYLOAD), dtype=torch.uint8, device=dev)
    pool = torch.empty(R * (C2 * 4 + 4096), dtype=torch.uint8, device=dev)
    te = td = ts = tv = 0.0
    N = 30
    for it in range(N + 5):
        t0 = time.perf_counter(); eb = enc.encode(built, abi.MSG_PREDICT_REQUEST, abi.F_ALIGN_PAYLOAD, dst=wire); t1 = time.perf_counter()
        d = dec.decode_into(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST, pool); t2 = time.perf_counter()
        d.close()
        s = dec.scan(eb.wire, eb.offsets, eb.lengths, abi.MSG_PREDICT_REQUEST); t3 = time.perf_counter()
        dec.values(s, pool); t4 = time.perf_counter()
        s.close()
        if it >= 5:
            te += t1 - t0; td += t2 - t1; ts += t3 - t2; tv += t4 - t3
    ideal = 2 * R * C2 * 4 / 6.4e12 * 1e6
    print(f"R={R:3d}: encode {te/N*1e6:7.1f} us  decode_into {td/N*1e6:7.1f} us  (scan {ts/N*1e6:7.1f} + values {tv/N*1e6:7.1f})   copy-kernel ideal {ideal:6.1f} us")
