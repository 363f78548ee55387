#!/usr/bin/env python
"""bench.py — TensorProto / PredictRequest wire encode+decode throughput (BASELINE.json's metric).

A step = one pass of the hot path over one batch of synthetic requests: for each of the two C2 variants
(ResNet-50 input 64x224x224x3 DT_FLOAT carried as `float_val`, and as `tensor_content`) encode R
PredictRequests to wire bytes and decode those wire bytes back to values.  `value` counts WIRE bytes
(encode output + decode input) per second with everything resident in HBM; `e2e` is the same work through
the host-memory entry points (payloads and wire in pinned host memory, H2D/D2H inside the timed region).

  python bench.py --gpus N --steps K --warmup W            our arm (one process per GPU; torchrun for N > 1)
  python bench.py --impl reference ...                      the reference's CPU path (oracle port of prost) on host cores

Requests are independent, so N GPUs each take their own shard of the batch: no collective on the data
path; torch.distributed (nccl) is only used for the barrier and the max-over-ranks time.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

METRIC = "tensorproto_encode_decode_wire_GBps"
UNIT = "GB/s"
C2_ELEMS = 64 * 224 * 224 * 3
C2_WIRE = 38535239  # prost encoded_len of one C2 PredictRequest
VARIANTS = ("float_val", "tensor_content")


def workload_name(r):
    return (f"C2 ResNet-50 input 64x224x224x3 DT_FLOAT as float_val and as tensor_content; per step and per GPU: "
            f"{r} PredictRequests per variant encoded, then decoded")


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc, self.thread = device, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except Exception:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# =====================================================================================================
# our arm
# =====================================================================================================
def pin_to_gpu_numa_node(index):
    """Bind this process to the CPUs NVML names as local to GPU `index` (one process per GPU: its pinned staging buffers
    are then first-touched on that socket).  Returns a short description for the bench line; failing is harmless."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w in range(words) for b in range(64) if (mask[w] >> b) & 1}
        everything = os.sched_getaffinity(0)
        cpus &= everything
        if cpus:
            os.sched_setaffinity(0, cpus)
            pin_to_gpu_numa_node.restore = lambda: os.sched_setaffinity(0, everything)
            return f"{len(cpus)} of {len(everything)} cpus (local to gpu {index}) during the GPU measurements"
    except Exception as e:  # noqa: BLE001
        return f"not bound ({type(e).__name__})"
    return "not bound"


def run_ours(args):
    import torch
    import torch.distributed as dist

    import tfs_b200 as tfs
    from tfs_b200 import abi, workloads

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the codec has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    numa = pin_to_gpu_numa_node(local)  # before any pinned allocation: the staging buffers should be local to the GPU's socket
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    codec = tfs.GpuCodec(tfs.GpuContext(local))
    ctx, enc, dec = codec.ctx, codec.encoder(), codec.decoder()
    R = args.requests
    kind = abi.MSG_PREDICT_REQUEST
    flags = abi.F_ALIGN_PAYLOAD

    # ---- device-resident inputs: R distinct payloads per variant (R * 38.5 MB > L2 for R >= 4)
    g = torch.Generator(device=dev)
    g.manual_seed(2001 + rank)
    src = torch.randn(R, C2_ELEMS, device=dev, generator=g)

    def requests(payload_rows, variant):
        out = []
        for i in range(R):
            shape = tfs.TensorShapeProto([64, 224, 224, 3])
            t = (tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, float_val=payload_rows[i]) if variant == "float_val"
                 else tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, tensor_content=payload_rows[i]))
            out.append(tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"), [("input", t)]))
        return out

    built = {v: tfs.build_batch(requests(src, v), kind) for v in VARIANTS}  # the "prost structs", built once
    cap = enc.bound(built["float_val"], flags)
    wire = {v: torch.empty(cap, dtype=torch.uint8, device=dev) for v in VARIANTS}
    pool = {v: torch.empty(R * (C2_ELEMS * 4 + 4096), dtype=torch.uint8, device=dev) for v in VARIANTS}

    def step():
        n = 0
        for v in VARIANTS:
            eb = enc.encode(built[v], kind, flags, dst=wire[v])
            d = dec.decode_into(eb.wire, eb.offsets, eb.lengths, kind, pool[v])
            n += 2 * sum(eb.lengths)
            d.close()
        return n

    # correctness of what is being timed: decoded values == source, wire length == prost's encoded_len
    eb = enc.encode(built["float_val"], kind, flags, dst=wire["float_val"])
    assert eb.lengths == [C2_WIRE] * R, eb.lengths[:2]
    d = dec.decode(eb.wire, eb.offsets, eb.lengths, kind, 0, pool=pool["float_val"])
    t0 = d.view.tensors[d.view.msgs[R - 1].first_tensor]
    got = pool["float_val"][t0.dst_off[1]: t0.dst_off[1] + 4 * C2_ELEMS].view(torch.float32)
    assert torch.equal(got, src[R - 1]), "decode(encode(x)) != x"
    d.close()

    for _ in range(args.warmup):
        step()
    ctx.prof_enable(True)
    for k in range(8):
        ctx.prof_read(k, reset=True)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    launches0 = ctx.launch_count()
    e0.record()
    wire_bytes = 0
    for _ in range(args.steps):
        wire_bytes += step()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = ctx.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    n_copy, ms_copy = ctx.prof_read(abi.PK_COPY_TMA)   # the TMA ring moves every payload byte of this workload
    n_gen, ms_gen = ctx.prof_read(abi.PK_COPY)          # the generic kernel only finds nothing to do here
    n_walk, ms_walk = ctx.prof_read(abi.PK_WALK)        # decode plan (both structure walks), one launch per decode
    n_frame, ms_frame = ctx.prof_read(abi.PK_FRAME)     # encode framing, one launch per encode
    ctx.prof_enable(False)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())

    # ---- end to end: pinned host payloads -> host wire -> host values, smaller batch (PCIe-bound)
    Re = min(args.e2e_requests, R)
    h_src = torch.randn(Re, C2_ELEMS).pin_memory()
    h_wire = torch.empty(enc.bound(built["float_val"], flags) // R * Re + 4096, dtype=torch.uint8).pin_memory()
    h_pool = torch.empty(Re * (C2_ELEMS * 4 + 4096), dtype=torch.uint8).pin_memory()
    rows = [h_src[i].numpy() for i in range(Re)]

    def host_requests(variant):
        out = []
        for i in range(Re):
            shape = tfs.TensorShapeProto([64, 224, 224, 3])
            t = (tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, float_val=rows[i]) if variant == "float_val"
                 else tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, tensor_content=rows[i]))
            out.append(tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"), [("input", t)]))
        return out

    hbuilt = {v: tfs.build_batch(host_requests(v), kind) for v in VARIANTS}
    h_wire_np, h_pool_np = h_wire.numpy(), h_pool.numpy()

    def e2e_step():
        n = h2d = d2h = 0
        for v in VARIANTS:
            eb = enc.encode_host(hbuilt[v], kind, flags, dst=h_wire_np)
            h2d += Re * C2_ELEMS * 4
            d2h += eb.total
            d = dec.decode_host_into(h_wire_np[: eb.total], eb.offsets, eb.lengths, kind, h_pool_np)
            h2d += eb.total
            d2h += d.pool_bytes
            n += 2 * sum(eb.lengths)
            d.close()
        return n, h2d, d2h

    e2e_step()
    e2e_steps = max(1, min(args.steps, 5))
    barrier()
    torch.cuda.synchronize()
    t_start = time.perf_counter()
    e2e_bytes = h2d = d2h = 0
    for _ in range(e2e_steps):
        a, b, c = e2e_step()
        e2e_bytes += a
        h2d, d2h = b, c
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t_start
    if world > 1:
        t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    last = pool_check = h_pool[: 16].clone()  # touch the result (device -> host read is part of every step)
    del last, pool_check

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = load_peaks()
    value = world * wire_bytes / (ms * 1e-3) / 1e9
    # K5 k_copy: one launch per encode and per decode; algorithmic bytes of one launch = read + write of
    # R payloads (fixed / content encode or decode = N*w_src + N*w_dst)
    alg_per_launch = 2.0 * R * C2_ELEMS * 4
    achieved = alg_per_launch / (ms_copy / max(1, n_copy) * 1e-3) / 1e9 if n_copy else None
    # DRAM bytes of one k_copy_tma launch from the committed `ncu --set full` capture (only if it is of this batch size)
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")))["k_copy_tma"]
        if tj["requests_per_variant"] == R:
            traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]
    except (OSError, KeyError, ValueError):
        pass
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(R), "requests_per_variant_per_gpu": R, "wire_bytes_per_request": C2_WIRE,
                   "l2": f"inputs larger than L2: each pass streams {R * C2_WIRE / 1e6:.0f} MB in and out (L2 is 126 MB)",
                   "flags": "TFS_F_ALIGN_PAYLOAD", "sharding": f"dp{world}: independent requests per GPU, no collective"},
        "clocks": clocks,
        "e2e": {"value": round(world * e2e_bytes / e2e_s / 1e9, 3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "requests_per_variant_per_gpu": Re, "steps": e2e_steps, "host_affinity": numa, "path": "tfs_encode_host + tfs_decode_host_into on pinned host buffers: copy engine brings the source to HBM in 8 MB chunks, the copy kernel follows chunk by chunk and stores straight into the pinned destination"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_copy_tma (K5 bulk copy through the TMA ring; one launch per encode and per decode)",
                     "other_copy_kernel_avg_ms": round(ms_gen / max(1, n_gen), 4),
                     "achieved": round(achieved, 1) if achieved else None,
                     "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4) if achieved else None, "peak_source": peak_src,
                     "frac_of_nominal_8000_GBps": round(achieved / 8000.0, 4) if achieved else None,
                     "traffic": traffic, "traffic_source": traffic_src, "launches_timed": n_copy, "avg_launch_ms": round(ms_copy / max(1, n_copy), 4),
                     "algorithmic_bytes_per_launch": alg_per_launch,
                     "step_frac_of_peak": round(2 * value / world / peak, 4),
                     "step_breakdown_ms": {"k_copy_tma": round(ms_copy / args.steps, 4), "k_plan_small": round(ms_walk / args.steps, 4),
                                           "k_frame_small": round(ms_frame / args.steps, 4),
                                           "rest (host syncs, table copies, k_dec_bind)": round((ms - ms_copy - ms_walk - ms_frame) / args.steps, 4)}},
    }
    getattr(pin_to_gpu_numa_node, "restore", lambda: None)()  # the CPU baseline gets every host core again
    line["cpu_baseline"] = cpu_baseline(R, budget_s=args.cpu_budget)
    if world == 1 and not args.no_extras:
        del src, wire, pool
        torch.cuda.empty_cache()
        line["other_configs"] = other_configs(codec, dev, peak)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# =====================================================================================================
# the other BASELINE configs (N = 1 only, a few repetitions each): parity-tested workloads, reported next to the
# headline so that the varint kernels have measured numbers too
# =====================================================================================================
def other_configs(codec, dev, peak):
    import torch
    import tfs_b200 as tfs
    from tfs_b200 import abi

    ctx, enc, dec = codec.ctx, codec.encoder(), codec.decoder()
    names = {abi.PK_COPY: "k_copy", abi.PK_VARINT_SIZE: "k_varint_size", abi.PK_VARINT_SCATTER: "k_varint_scatter",
             abi.PK_VARINT_COUNT: "k_varint_count", abi.PK_VARINT_GATHER: "k_varint_gather", abi.PK_COPY_TMA: "k_copy_tma"}

    def measure(msgs, kind, in_bytes, reps=5, dflags=0, decode_only=False):
        built = tfs.build_batch(msgs, kind)
        wire = torch.empty(enc.bound(built, 0), dtype=torch.uint8, device=dev)
        eb = enc.encode(built, kind, 0, dst=wire)
        W = sum(eb.lengths)
        d = dec.scan(eb.wire, eb.offsets, eb.lengths, kind, dflags)
        out_bytes = d.pool_bytes
        pool = torch.empty(out_bytes + 64, dtype=torch.uint8, device=dev)
        d.close()
        for _ in range(2):
            if not decode_only:
                eb = enc.encode(built, kind, 0, dst=wire)
            dec.decode_into(eb.wire, eb.offsets, eb.lengths, kind, pool, dflags).close()
        ctx.prof_enable(True)
        for k in names:
            ctx.prof_read(k, reset=True)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        te = td = 0.0
        for _ in range(reps):
            ev[0].record()
            if not decode_only:
                eb = enc.encode(built, kind, 0, dst=wire)
            ev[1].record()
            dec.decode_into(eb.wire, eb.offsets, eb.lengths, kind, pool, dflags).close()
            ev[2].record()
            torch.cuda.synchronize()
            te += ev[0].elapsed_time(ev[1]) / reps
            td += ev[1].elapsed_time(ev[2]) / reps
        ctx.prof_enable(False)
        alg = {abi.PK_VARINT_SIZE: in_bytes, abi.PK_VARINT_SCATTER: in_bytes + W, abi.PK_VARINT_COUNT: W, abi.PK_VARINT_GATHER: W + out_bytes}
        kernels = {}
        for k, nm in names.items():
            n, ms = ctx.prof_read(k, reset=True)
            if n and k in alg and not decode_only:  # (C5's wire is mostly fixed-width payload: no per-kernel figure)
                gbs = alg[k] / (ms / n * 1e-3) / 1e9
                kernels[nm] = {"avg_launch_ms": round(ms / n, 4), "algorithmic_GBps": round(gbs, 1), "frac_of_measured_peak": round(gbs / peak, 3)}
        res = {"messages": len(msgs), "wire_MB": round(W / 1e6, 2), "values_MB": round(out_bytes / 1e6, 2),
               "decode_ms": round(td, 4), "decode_wire_GBps": round(W / (td * 1e-3) / 1e9, 1),
               "decode_algorithmic_GBps": round((W + out_bytes) / (td * 1e-3) / 1e9, 1), "kernels": kernels}
        if not decode_only:
            res.update({"encode_ms": round(te, 4), "encode_wire_GBps": round(W / (te * 1e-3) / 1e9, 1),
                        "encode_algorithmic_GBps": round((in_bytes + W) / (te * 1e-3) / 1e9, 1),
                        "encode_decode_wire_GBps": round(2 * W / ((te + td) * 1e-3) / 1e9, 1)})
        return res

    out = {}
    g = torch.Generator(device=dev)
    g.manual_seed(3001)
    R = 256
    ids = torch.randint(0, 30522, (R, 256 * 512), device=dev, dtype=torch.int64, generator=g)
    lens = torch.randint(1, 513, (R, 256, 1), device=dev, generator=g)
    mask = (torch.arange(512, device=dev)[None, None, :] < lens).to(torch.int64).reshape(R, -1).contiguous()
    shape = tfs.TensorShapeProto([256, 512])
    msgs = [tfs.PredictRequest(tfs.ModelSpec("bert", signature_name="serving_default"),
                               [("input_ids", tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=ids[i])),
                                ("attention_mask", tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=mask[i]))]) for i in range(R)]
    out["C3 BERT input_ids+attention_mask int64_val 256x512, 256 requests"] = measure(msgs, abi.MSG_PREDICT_REQUEST, 2 * R * 256 * 512 * 8)
    del ids, mask, msgs
    R4 = 64
    x = torch.randn(R4, 1 << 20, device=dev, generator=g)
    msgs = [tfs.TensorProto(dtype=abi.DT_HALF, tensor_shape=tfs.TensorShapeProto([1 << 20]), half_val=tfs.Payload(x[i], abi.T_F32, abi.T_F16))
            for i in range(R4)]
    out["C4 DT_HALF half_val 2^20 elements from f32 (decode widens to f32), 64 tensors"] = measure(
        msgs, abi.MSG_TENSOR_PROTO, R4 * (1 << 22), dflags=abi.F_HALF_TO_F32)
    del x, msgs
    # C5: the per-GPU share (1024 of 8192) of mixed-dtype PredictResponses, decoded
    import numpy as np
    rng = np.random.default_rng(5001)
    msgs = []
    for _ in range(1024):
        entries = []
        for j in range(int(rng.integers(1, 4))):
            n = int(2 ** rng.uniform(0, 18))
            sh = tfs.TensorShapeProto([1] * int(rng.integers(0, 4)) + [n])
            which = int(rng.integers(0, 7))
            if which == 0:
                t = tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=sh, float_val=torch.randn(n, device=dev, generator=g))
            elif which == 1:
                t = tfs.TensorProto(dtype=abi.DT_DOUBLE, tensor_shape=sh, double_val=torch.randn(n, device=dev, generator=g, dtype=torch.float64))
            elif which == 2:
                v = torch.randint(-(1 << 31), 1 << 31, (n,), device=dev, generator=g, dtype=torch.int64).to(torch.int32)
                t = tfs.TensorProto(dtype=abi.DT_INT32, tensor_shape=sh, int_val=v)
            elif which == 3:
                v = torch.randint(-(1 << 62), 1 << 62, (n,), device=dev, generator=g, dtype=torch.int64)
                t = tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=sh, int64_val=v)
            elif which == 4:
                h = torch.randn(n, device=dev, generator=g).to(torch.float16)
                t = tfs.TensorProto(dtype=abi.DT_HALF, tensor_shape=sh, half_val=h)
            elif which == 5:
                t = tfs.TensorProto(dtype=abi.DT_BOOL, tensor_shape=sh, bool_val=torch.rand(n, device=dev, generator=g) < 0.5)
            else:
                t = tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=sh, tensor_content=torch.randn(n, device=dev, generator=g))
            entries.append((f"out{j}", t))
        msgs.append(tfs.PredictResponse(entries, tfs.ModelSpec("mixed", version=int(rng.integers(1, 100)), signature_name="serving_default")))
    out["C5 mixed-dtype PredictResponses, 1024 messages (one GPU's share of 8192), decode"] = measure(
        msgs, abi.MSG_PREDICT_RESPONSE, 0, decode_only=True)
    del msgs

    # C1: the half_plus_two request (63 wire bytes), a large batch of copies: latency / descriptor bound by construction
    x3 = torch.tensor([1.0, 2.0, 5.0], device=dev)
    c1 = [tfs.PredictRequest(tfs.ModelSpec("half_plus_two", signature_name="serving_default"),
                             [("x", tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=tfs.TensorShapeProto([3]), float_val=x3))]) for _ in range(16384)]
    r1 = measure(c1, abi.MSG_PREDICT_REQUEST, 12 * 16384, reps=3)
    r1["encode_msgs_per_s"] = round(16384 / (r1["encode_ms"] * 1e-3))
    r1["decode_msgs_per_s"] = round(16384 / (r1["decode_ms"] * 1e-3))
    out["C1 half_plus_two PredictRequest (63 B), 16 384 copies in one batch"] = r1
    del c1

    # DT_STRING tensors (string_val): encoded images, and short tokens
    def string_requests(n_req, n_str, lo, hi):
        msgs, in_bytes = [], 0
        for _ in range(n_req):
            lens = torch.randint(lo, hi + 1, (n_str,), device=dev, generator=g, dtype=torch.int64)
            offs = torch.zeros(n_str + 1, device=dev, dtype=torch.int64)
            offs[1:] = torch.cumsum(lens, 0)
            total = int(offs[-1].item())
            block = torch.empty(8 * (n_str + 1) + total, device=dev, dtype=torch.uint8)
            block[: 8 * (n_str + 1)] = offs.view(torch.uint8)
            block[8 * (n_str + 1):] = torch.randint(0, 256, (total,), device=dev, generator=g, dtype=torch.uint8)
            t = tfs.TensorProto(dtype=abi.DT_STRING, tensor_shape=tfs.TensorShapeProto([n_str]), string_val=tfs.Ragged(block, n_str, total))
            msgs.append(tfs.PredictRequest(tfs.ModelSpec("strings", signature_name="serving_default"), [("examples", t)]))
            in_bytes += block.numel()
        return msgs, in_bytes
    msgs, nb = string_requests(64, 32, 100_000, 300_000)
    out["DT_STRING string_val, 64 requests x 32 encoded images of 100-300 kB"] = measure(msgs, abi.MSG_PREDICT_REQUEST, nb, reps=3)
    del msgs
    # Classify / Regress inputs (section 8(f) row 1), encode only: 64 requests x 64 Examples x {ids int64[128], weights f32[128], tag bytes}
    def classify_requests(n_req, n_ex, n_val):
        reqs, in_bytes = [], 0
        for r in range(n_req):
            ids = torch.randint(0, 30522, (n_ex, n_val), device=dev, generator=g, dtype=torch.int64)
            w = torch.randn(n_ex, n_val, device=dev, generator=g)
            tag = tfs.ragged_block([b"label-%d" % (r % 7)])
            tag_dev = torch.from_numpy(tag).to(dev)
            exs = [[("ids", tfs.Feature(int64_list=ids[e])), ("weights", tfs.Feature(float_list=w[e])),
                    ("tag", tfs.Feature(bytes_list=tfs.Ragged(tag_dev, 1, tag.size - 16)))] for e in range(n_ex)]
            reqs.append(tfs.ClassificationRequest(tfs.ModelSpec("wide_deep", signature_name="classify"), tfs.Input(exs)))
            in_bytes += n_ex * n_val * 12
        return reqs, in_bytes
    reqs, nb = classify_requests(64, 64, 128)
    built = tfs.build_batch(reqs, abi.MSG_CLASSIFY_REQUEST)
    wire = torch.empty(enc.bound(built, 0), dtype=torch.uint8, device=dev)
    for _ in range(2):
        eb = enc.encode(built, abi.MSG_CLASSIFY_REQUEST, 0, dst=wire)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(3):
        eb = enc.encode(built, abi.MSG_CLASSIFY_REQUEST, 0, dst=wire)
    ev1.record()
    torch.cuda.synchronize()
    te = ev0.elapsed_time(ev1) / 3
    W = sum(eb.lengths)
    out["ClassificationRequest, 64 requests x 64 Examples x (int64[128] + f32[128] + 1 tag), encode"] = {
        "messages": 64, "features": 64 * 64 * 3, "wire_MB": round(W / 1e6, 2), "encode_ms": round(te, 4),
        "encode_wire_GBps": round(W / (te * 1e-3) / 1e9, 1), "encode_algorithmic_GBps": round((nb + W) / (te * 1e-3) / 1e9, 1),
        "note": "12 288 features: framed by one GPU thread each (k_feature_layout / k_feature_write); most of the call is the host walking 12 288 descriptors and their name strings"}
    del reqs, built, wire
    msgs, nb = string_requests(64, 512, 0, 40)
    out["DT_STRING string_val, 64 requests x 512 tokens of 0-40 bytes (the decode walk is serial per message: ~1.6 us per element)"] = measure(msgs, abi.MSG_PREDICT_REQUEST, nb, reps=3)
    return out


# =====================================================================================================
# CPU side: the oracle port of prost's algorithm, timed on the host cores
# =====================================================================================================
def cpu_pass(threads, n_msgs, reps_budget_s, min_reps=1):
    """Encode + decode n_msgs C2 requests per variant with `threads` host threads; returns (GB/s wire, reps, seconds)."""
    import tfs_b200 as tfs
    from tfs_b200 import abi
    import oracle_bind as ob

    x = np.random.default_rng(2001).standard_normal(C2_ELEMS, dtype=np.float32)
    total_b, total_s, reps = 0, 0.0, 0
    preps = []
    for v in VARIANTS:
        shape = tfs.TensorShapeProto([64, 224, 224, 3])
        msgs = []
        for _ in range(n_msgs):
            t = (tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, float_val=x) if v == "float_val"
                 else tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, tensor_content=x))
            msgs.append(tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"), [("input", t)]))
        preps.append(ob.Prepared(tfs.build_batch(msgs, abi.MSG_PREDICT_REQUEST)))
    t_begin = time.perf_counter()
    while reps < min_reps or (time.perf_counter() - t_begin) < reps_budget_s:
        for p in preps:
            s_e, wb = p.run_encode(threads)
            s_d, wb2, _ = p.run_decode(threads)
            assert wb == wb2 == n_msgs * C2_WIRE
            total_b += wb + wb2
            total_s += s_e + s_d
        reps += 1
    for p in preps:
        p.close()
    return total_b / total_s / 1e9, reps, total_s


def cpu_baseline(R, budget_s=12.0):
    cores = os.cpu_count() or 1
    threads = max(1, min(cores, 16))
    n_msgs = threads
    gbs, reps, secs = cpu_pass(threads, n_msgs, budget_s)
    return {"value": round(gbs, 3), "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{n_msgs} C2 requests per variant (one per thread), encode+decode, {reps} passes, {secs:.1f} s of CPU work; "
                      f"oracle/cpu_ref.cpp = C++ port of prost 0.6.0's pass structure (prost itself: unavailable on this box, no rustc/cargo)",
            "host_cores_available": cores}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    threads = max(1, min(cores, 64))
    n_msgs = threads
    # warm-up + K timed steps, each a bounded sample (one request per thread and variant)
    per_step_budget = 0.0
    t0 = time.perf_counter()
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_pass(threads, n_msgs, 0.0)
    steps = max(1, min(args.steps, 5))
    tot_b = tot_s = 0.0
    for _ in range(steps):
        gbs, reps, secs = cpu_pass(threads, n_msgs, per_step_budget)
        tot_b += gbs * secs
        tot_s += secs
    value = tot_b / tot_s
    sample = (f"{n_msgs} C2 requests per variant per step (one per host thread), {steps} steps, oracle/cpu_ref.cpp "
              f"(C++ port of prost 0.6.0; the Rust reference cannot be built here: no rustc/cargo, crates not vendored)")
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": min(args.warmup, 1), "ms_per_step": round(tot_s / steps * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(args.requests), "requests_per_variant_per_gpu": args.requests,
                       "wire_bytes_per_request": C2_WIRE, "sample_requests_per_variant_per_step": n_msgs},
            "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                             "host_cores_available": cores},
            "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": round(time.perf_counter() - t0, 1)}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--requests", type=int, default=32, help="C2 requests per variant, per GPU and per step")
    ap.add_argument("--e2e-requests", type=int, default=8)
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU baseline sampling")
    ap.add_argument("--no-extras", action="store_true", help="skip the C3 / C4 / C5 side measurements (N = 1 only)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
