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

Next to the headline the line carries, for the other BASELINE configs (C1, C3, C4, C5) and for the C2 variants that
do not ride the aligned fast path (back-to-back wire, fused u8 -> f32, fused f32 -> bf16), call-level roofline
fractions, the dominant kernel's fraction, the CPU port on the same inputs (1 thread and all cores) and an end-to-end
figure for pinned and for pageable host buffers: nested under `per_config` and, flattened to scalars, inside `roofline`.
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
    return f"C2 ResNet-50 input 64x224x224x3 f32 as float_val + tensor_content, {r} requests per variant: encode then decode"


def config_of(args, world):
    """The `config` object: the same keys in both arms (the driver compares them)."""
    return {"workload": workload_name(args.requests), "requests_per_variant_per_gpu": args.requests, "wire_bytes_per_request": C2_WIRE,
            "l2": "inputs larger than L2 (126 MB): each pass streams %d MB" % (args.requests * C2_WIRE // 1000000),
            "flags": "TFS_F_ALIGN_PAYLOAD", "sharding": f"dp{world}: independent requests per GPU, no collective"}


def host_threads():
    """One rule for every CPU number of this file: all the host cores the process may use."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


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
def run_ours(args):
    import torch
    import torch.distributed as dist

    import tfs_b200 as tfs
    from tfs_b200 import abi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the codec has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # the codec runs on torch's CURRENT stream, a stream of our own (the legacy default stream serialises against every other
    # blocking stream of the process); the events below are recorded on the same stream
    side = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(side)
    codec = tfs.GpuCodec(tfs.GpuContext(local))
    ctx, enc, dec = codec.ctx, codec.encoder(), codec.decoder()
    R = args.requests
    kind = abi.MSG_PREDICT_REQUEST
    flags = abi.F_ALIGN_PAYLOAD

    # ---- device-resident inputs: R distinct payloads per variant (R * 38.5 MB > L2 for R >= 4)
    g = torch.Generator(device=dev)
    g.manual_seed(2001 + rank)
    src = torch.randn(R, C2_ELEMS, device=dev, generator=g)

    def c2_requests(rows, variant, n):
        out = []
        for i in range(n):
            shape = tfs.TensorShapeProto([64, 224, 224, 3])
            t = (tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, float_val=rows[i]) if variant == "float_val"
                 else tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape, tensor_content=rows[i]))
            out.append(tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"), [("input", t)]))
        return out

    built = {v: tfs.build_batch(c2_requests(src, v, R), kind) for v in VARIANTS}  # the "prost structs", built once
    cap = enc.bound(built["float_val"], flags)
    wire = {v: torch.empty(cap, dtype=torch.uint8, device=dev) for v in VARIANTS}
    pool = {v: torch.empty(R * (C2_ELEMS * 4 + 4096), dtype=torch.uint8, device=dev) for v in VARIANTS}

    def step():
        n = 0
        for v in VARIANTS:
            # F_ASYNC: the encode call returns when the message offsets are known; the decode of the same bytes is queued
            # behind the payload kernel on the same stream (everything is waited for before the timed region ends)
            eb = enc.encode(built[v], kind, flags | abi.F_ASYNC, dst=wire[v])
            d = dec.decode_into(eb.wire, eb.offsets_np, eb.lengths_np, kind, pool[v])
            n += 2 * int(eb.lengths_np.sum())
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
    for k in range(abi.PK_COUNT):
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
    ctx.sync()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = ctx.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    n_copy, ms_copy = ctx.prof_read(abi.PK_COPY_TMA)   # the TMA ring moves every payload byte of this workload
    n_gen, ms_gen = ctx.prof_read(abi.PK_COPY)          # the generic kernel only finds nothing to do here
    n_walk, ms_walk = ctx.prof_read(abi.PK_WALK)        # decode plan (structure walk), one launch per decode
    n_frame, ms_frame = ctx.prof_read(abi.PK_FRAME)     # encode framing, one launch per encode
    ctx.prof_enable(False)
    ms = max_over_ranks(ms)

    # ---- a sustained side figure: the same step for >= 1 s, with its own clock samples
    sustained = None
    if rank == 0 and world == 1 and args.sustain_s > 0:
        s2 = ClockSampler(local)
        s2.start()
        torch.cuda.synchronize()
        e0.record()
        t_begin, nb, ns = time.perf_counter(), 0, 0
        while time.perf_counter() - t_begin < args.sustain_s:
            nb += step()
            ns += 1
        ctx.sync()
        e1.record()
        torch.cuda.synchronize()
        sms = e0.elapsed_time(e1)
        ck = s2.stop()
        sustained = {"GBps": round(nb / (sms * 1e-3) / 1e9, 1), "seconds": round(sms * 1e-3, 2), "steps": ns, "sm_mhz": ck.get("sm_mhz"),
                     "reasons": ",".join(ck.get("reasons") or [])}

    # ---- end to end: pinned host payloads -> host wire -> host values (PCIe-bound), same batch size as `value`
    Re = min(args.e2e_requests, R)
    h_src = torch.randn(Re, C2_ELEMS).pin_memory()
    h_wire = torch.empty(enc.bound(built["float_val"], flags) // R * Re + 4096, dtype=torch.uint8).pin_memory()
    h_pool = torch.empty(Re * (C2_ELEMS * 4 + 4096), dtype=torch.uint8).pin_memory()
    rows = [h_src[i].numpy() for i in range(Re)]
    hbuilt = {v: tfs.build_batch(c2_requests(rows, v, Re), kind) for v in VARIANTS}
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
    e2e_s = max_over_ranks(time.perf_counter() - t_start)
    assert np.array_equal(h_pool_np[: 16], h_pool_np[: 16])  # the result is read on the host (device -> host is part of every step)

    # ---- what the box can move: pinned host <-> HBM in both directions at once (copy engines, 256 MB buffers), every rank
    #      at the same time, so that the aggregate at N GPUs is the ceiling the e2e figure at N GPUs lives under
    def pcie_ceiling():
        nb = 256 << 20
        hp_in, hp_out = torch.empty(nb, dtype=torch.uint8).pin_memory(), torch.empty(nb, dtype=torch.uint8).pin_memory()
        d_in, d_out = torch.empty(nb, dtype=torch.uint8, device=dev), torch.empty(nb, dtype=torch.uint8, device=dev)
        s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        best = 0.0
        for rep in range(4):
            barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(4):
                with torch.cuda.stream(s1):
                    d_in.copy_(hp_in, non_blocking=True)
                with torch.cuda.stream(s2):
                    hp_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()
            dt = max_over_ranks(time.perf_counter() - t0)
            if rep:
                best = max(best, 4 * 2 * nb / dt / 1e9)
        return best * world  # every rank moved the same bytes in the (max over ranks) time

    pcie_bidir = pcie_ceiling()
    e2e_moved_GBps = world * e2e_steps * (h2d + d2h) / e2e_s / 1e9

    # ---- the same calls with PAGEABLE buffers (what prost's Vecs and tonic's BytesMut are): the library's pinned ring and
    #      staging threads are inside the timed region; 8 requests per variant (side figure)
    e2e_pageable = None
    if not args.no_extras:
        Rp = min(8, Re)
        p_rows = [np.array(h_src[i].numpy(), copy=True) for i in range(Rp)]
        p_built = {v: tfs.build_batch(c2_requests(p_rows, v, Rp), kind) for v in VARIANTS}
        p_wire = np.empty(enc.bound(p_built["float_val"], flags) + 4096, np.uint8)
        p_pool = np.empty(Rp * (C2_ELEMS * 4 + 4096), np.uint8)

        def pageable_step():
            n = 0
            for v in VARIANTS:
                eb = enc.encode_host(p_built[v], kind, flags, dst=p_wire)
                dec.decode_host_into(p_wire[: eb.total], eb.offsets_np, eb.lengths_np, kind, p_pool).close()
                n += 2 * int(eb.lengths_np.sum())
            return n

        pageable_step()
        barrier()
        t_start = time.perf_counter()
        nb = 0
        for _ in range(3):
            nb += pageable_step()
        e2e_pageable = world * nb / max_over_ranks(time.perf_counter() - t_start) / 1e9
        del p_rows, p_built, p_wire, p_pool

    # ---- the same work through the engine: two lanes (context + stream + worker thread each) on this GPU, the two variants in
    #      flight at once, so that one job's upload runs under the other's download (side figure; `e2e` above is one call at a time)
    e2e_engine = None
    if not args.no_extras:
        eng = tfs.Engine([local], lanes_per_device=2, pinned_bytes_per_lane=0)
        h_wire2 = torch.empty(h_wire.numel(), dtype=torch.uint8).pin_memory()
        h_pool2 = torch.empty(h_pool.numel(), dtype=torch.uint8).pin_memory()
        wires, pools = {"float_val": h_wire_np, "tensor_content": h_wire2.numpy()}, {"float_val": h_pool_np, "tensor_content": h_pool2.numpy()}

        def engine_step():
            te = {v: eng.encode_host(hbuilt[v], wires[v], flags, lane=i) for i, v in enumerate(VARIANTS)}
            ebs = {v: eng.wait(t) for v, t in te.items()}
            td = {v: eng.decode_host_into(wires[v][: ebs[v].total], ebs[v].offsets, ebs[v].lengths, kind, pools[v], lane=i) for i, v in enumerate(VARIANTS)}
            n = 0
            for v, t in td.items():
                eng.wait(t).close()
                n += 2 * sum(ebs[v].lengths)
            return n

        engine_step()
        barrier()
        t_start = time.perf_counter()
        nb = 0
        for _ in range(e2e_steps):
            nb += engine_step()
        e2e_engine = world * nb / max_over_ranks(time.perf_counter() - t_start) / 1e9
        eng.close()
        del h_wire2, h_pool2
    del h_src, h_wire, h_pool, hbuilt

    # ---- N > 1: BASELINE config 5, "8192 independent mixed-dtype PredictResponses decoded, sharded across the GPUs": every
    #      rank decodes its own 8192 / N responses (wire produced and, on rank 0, values checked by the oracle), no collective
    c5 = None
    if world > 1 and not args.no_extras:
        c5 = c5_sharded(codec, dev, rank, world, barrier, max_over_ranks)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = load_peaks()
    value = world * wire_bytes / (ms * 1e-3) / 1e9
    # K5 k_copy_tma: one launch per encode and per decode; algorithmic bytes of one launch = read + write of
    # R payloads (fixed / content encode or decode = N*w_src + N*w_dst)
    alg_per_launch = 2.0 * R * C2_ELEMS * 4
    achieved = alg_per_launch / (ms_copy / max(1, n_copy) * 1e-3) / 1e9 if n_copy else None
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_copy_tma"]
        if tj["requests_per_variant"] == R:
            traffic = tj["dram_bytes_per_launch"]
    except (OSError, KeyError, ValueError):
        pass
    roof = {"bound": "hbm", "kernel": "k_copy_tma (K5 bulk copy through the TMA ring; one launch per encode and per decode)",
            "achieved": round(achieved, 1) if achieved else None, "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 4) if achieved else None, "peak_source": peak_src,
            "frac_of_nominal_8000_GBps": round(achieved / 8000.0, 4) if achieved else None,
            "encode_call": "TFS_F_ASYNC: returns when offsets are known; stream waited for inside the timed region",
            "traffic": traffic, "launches_timed": n_copy, "avg_launch_ms": round(ms_copy / max(1, n_copy), 4),
            "algorithmic_bytes_per_launch": alg_per_launch, "step_frac_of_peak": round(2 * value / world / peak, 4),
            "step_ms_k_copy_tma": round(ms_copy / args.steps, 4), "step_ms_walk": round(ms_walk / args.steps, 4),
            "step_ms_frame": round(ms_frame / args.steps, 4), "step_ms_other_copy": round(ms_gen / args.steps, 4),
            "step_ms_rest": round((ms - ms_copy - ms_walk - ms_frame - ms_gen) / args.steps, 4)}
    if c5:
        roof.update(c5)
    if sustained:
        roof.update({"sustained_GBps": sustained["GBps"], "sustained_seconds": sustained["seconds"], "sustained_sm_mhz": sustained["sm_mhz"],
                     "sustained_reasons": sustained["reasons"]})
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "config": config_of(args, world), "clocks": clocks,
        "e2e": {"value": round(world * e2e_bytes / e2e_s / 1e9, 3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "requests_per_variant_per_gpu": Re, "steps": e2e_steps, "engine_two_lanes_GBps": round(e2e_engine, 3) if e2e_engine else None,
                "pageable_buffers_GBps": round(e2e_pageable, 3) if e2e_pageable else None,
                "pcie_bidir_measured_GBps": round(pcie_bidir, 1), "pcie_moved_GBps": round(e2e_moved_GBps, 1),
                "frac_of_pcie_bidir": round(e2e_moved_GBps / pcie_bidir, 3) if pcie_bidir else None,
                "path": "tfs_encode_host + tfs_decode_host_into, pinned host buffers"},
        "gpu_launches": int(launches), "roofline": roof,
    }
    del src, wire, pool, built
    torch.cuda.empty_cache()
    line["cpu_baseline"] = cpu_baseline_c2(budget_s=args.cpu_budget)
    if world == 1 and not args.no_extras:
        pc = per_config(codec, dev, peak, args)
        line["per_config"] = pc
        for name, res in pc.items():  # the driver's parser keeps scalars inside `roofline`: a flat copy
            for k, v in res.items():
                if isinstance(v, (int, float)) or v is None:
                    roof[f"pc_{name}_{k}"] = v
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def c5_sharded(codec, dev, rank, world, barrier, max_over_ranks, total_msgs=8192, reps=5):
    """-> flat dict for the bench line (rank 0), None elsewhere.  Aggregate = bytes of all ranks / slowest rank's time."""
    import torch
    import torch.distributed as dist
    import tfs_b200 as tfs
    from tfs_b200 import abi, workloads
    import oracle_bind as ob
    import helpers

    n_mine = total_msgs // world
    msgs = workloads.c5_responses(n_mine, seed=5001 + rank, max_log2=18)  # one shard per rank, its own seed
    wire, offs, lens = ob.encode(tfs.build_batch(msgs, abi.MSG_PREDICT_RESPONSE))
    dwire = torch.from_numpy(wire).to(dev)
    dec, ctx = codec.decoder(), codec.ctx
    d = dec.scan(dwire, offs, lens, abi.MSG_PREDICT_RESPONSE)
    out_bytes = d.pool_bytes
    pool = torch.empty(out_bytes + 64, dtype=torch.uint8, device=dev)
    d.close()
    d = dec.decode_into(dwire, offs, lens, abi.MSG_PREDICT_RESPONSE, pool)
    ctx.sync()
    if rank == 0:  # values against the oracle, once
        st_o, want = ob.decode(wire, offs, lens, abi.MSG_PREDICT_RESPONSE)
        helpers.assert_same_decoded_fast(want, d.messages(), "response", "bench C5 shard of rank 0")
        del want
    d.close()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        dec.decode_into(dwire, offs, lens, abi.MSG_PREDICT_RESPONSE, pool).close()
    ctx.sync()
    e1.record()
    torch.cuda.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1) / reps)
    t = torch.tensor([float(wire.size), float(out_bytes)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    if rank != 0:
        return None
    W, V = float(t[0].item()), float(t[1].item())
    return {"c5_sharded_messages": n_mine * world, "c5_sharded_ms": round(ms, 4), "c5_sharded_wire_GBps": round(W / (ms * 1e-3) / 1e9, 1),
            "c5_sharded_algorithmic_GBps": round((W + V) / (ms * 1e-3) / 1e9, 1), "c5_sharded_checked_against_oracle": 1}


# =====================================================================================================
# the other configs (N = 1 only): per direction the call-level algorithmic GB/s and fraction of the measured copy
# peak, the dominant kernel's fraction, the CPU port on the same inputs, end to end for pinned and pageable buffers
# =====================================================================================================
def _pin(a):
    import torch
    t = torch.from_numpy(a).pin_memory()
    return t, t.numpy()


def per_config(codec, dev, peak, args):
    import torch
    import tfs_b200 as tfs
    from tfs_b200 import abi, workloads
    import oracle_bind as ob

    ctx, enc, dec = codec.ctx, codec.encoder(), codec.decoder()
    names = dict(abi.PK_NAMES)
    threads = host_threads()

    def device_side(msgs_dev, kind, in_bytes, eflags=0, dflags=0, reps=5, decode_wire=None, encode=True, decode=True):
        """-> dict with call-level times and the largest kernel of each direction"""
        res = {}
        if encode:
            built = tfs.build_batch(msgs_dev, kind) if not isinstance(msgs_dev, tfs.BuiltBatch) else msgs_dev
            wire = torch.empty(enc.bound(built, eflags), dtype=torch.uint8, device=dev)
            eb = enc.encode(built, kind, eflags, dst=wire)
            W = int(eb.lengths_np.sum())
            wire_t, offs, lens = eb.wire, eb.offsets_np, eb.lengths_np  # arrays: no per-message Python work in the timed calls
        else:
            wire_t, offs, lens = decode_wire
            W = int(np.sum(lens))
        out_bytes = 0
        if decode:
            d = dec.scan(wire_t, offs, lens, kind, dflags)
            out_bytes = d.pool_bytes
            pool = torch.empty(out_bytes + 64, dtype=torch.uint8, device=dev)
            d.close()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

        def timed(fn, alg):
            for _ in range(2):
                fn()
            ctx.prof_enable(True)
            for k in names:
                ctx.prof_read(k, reset=True)
            ctx.sync()
            ev[0].record()
            for _ in range(reps):
                fn()
            ctx.sync()
            ev[1].record()
            torch.cuda.synchronize()
            ms = ev[0].elapsed_time(ev[1]) / reps
            ctx.prof_enable(False)
            top, top_ms = None, 0.0
            for k, nm in names.items():
                n, kms = ctx.prof_read(k, reset=True)
                if n and kms / reps > top_ms:
                    top, top_ms = nm, kms / reps
            return ms, top, top_ms

        res["wire_MB"] = round(W / 1e6, 2)
        if encode:
            alg = in_bytes + W
            # back-to-back calls, each returning when its offsets are known (F_ASYNC); the stream is waited for before the
            # closing event, so every kernel of every call is inside the timed region
            ms, top, top_ms = timed(lambda: enc.encode(built, kind, eflags | abi.F_ASYNC, dst=wire), alg)
            res.update({"enc_ms": round(ms, 4), "enc_GBps": round(alg / (ms * 1e-3) / 1e9, 1), "enc_frac": round(alg / (ms * 1e-3) / 1e9 / peak, 3),
                        "enc_kernel": top, "enc_kernel_ms": round(top_ms, 4),
                        "enc_kernel_frac": round(alg / (top_ms * 1e-3) / 1e9 / peak, 3) if top_ms else None})
        if decode:
            alg = W + out_bytes
            ms, top, top_ms = timed(lambda: dec.decode_into(wire_t, offs, lens, kind, pool, dflags).close(), alg)
            res.update({"dec_ms": round(ms, 4), "dec_GBps": round(alg / (ms * 1e-3) / 1e9, 1), "dec_frac": round(alg / (ms * 1e-3) / 1e9 / peak, 3),
                        "dec_kernel": top, "dec_kernel_ms": round(top_ms, 4),
                        "dec_kernel_frac": round(alg / (top_ms * 1e-3) / 1e9 / peak, 3) if top_ms else None})
        res["values_MB"] = round(out_bytes / 1e6, 2)
        return res

    def cpu_side(msgs_host, kind, budget_s=1.5, construct=False):
        """The oracle port on (a sample of) the same inputs: wire GB/s of encode+decode, 1 thread and all cores.
        construct=True also times the message construction (the reference's u8 -> f32 cast loop, src/lib.rs:237-242)."""
        built = tfs.build_batch(msgs_host, kind)
        t0 = time.perf_counter()
        p = ob.Prepared(built)
        t_build = time.perf_counter() - t0
        out = {}
        for label, th in (("cpu1_GBps", 1), ("cpuN_GBps", threads)):
            tot_b = tot_s = 0.0
            t_begin = time.perf_counter()
            while True:
                s_e, wb = p.run_encode(th)
                s_d, wb2, _ = p.run_decode(th)
                tot_b += wb + wb2
                tot_s += s_e + s_d + (t_build / th if construct else 0.0)
                if time.perf_counter() - t_begin > budget_s:
                    break
            out[label] = round(tot_b / tot_s / 1e9, 3)
        p.close()
        out["cpu_threads"] = threads
        return out

    def e2e_side(msgs_host, kind, eflags=0, dflags=0, reps=5):
        """tfs_encode_host + tfs_decode_host_into on the host copies of the inputs: pinned wire / pool, then pageable"""
        out = {}
        from test_gpu_parity import pinned_like  # the same messages with every payload array in pinned host memory
        for label, pinned in (("e2e_pinned_GBps", True), ("e2e_pageable_GBps", False)):
            if pinned:
                pmsgs, keep_payloads = pinned_like(msgs_host, kind)
                built = tfs.build_batch(pmsgs, kind)
            else:
                built = tfs.build_batch(msgs_host, kind)
            cap = enc.bound(built, eflags) + 4096
            if pinned:
                keep_w, wbuf = _pin(np.empty(cap, np.uint8))
            else:
                wbuf = np.empty(cap, np.uint8)
            eb = enc.encode_host(built, kind, eflags, dst=wbuf)
            d = dec.decode_host(wbuf[: eb.total], eb.offsets, eb.lengths, kind, dflags)
            pb = d.pool_bytes + 64
            d.close()
            if pinned:
                keep_p, pbuf = _pin(np.empty(pb, np.uint8))
            else:
                pbuf = np.empty(pb, np.uint8)
            # one untimed round trip: the first decode_host_into grows the ctx's buffers and touches the new pool's pages
            dec.decode_host_into(wbuf[: eb.total], eb.offsets, eb.lengths, kind, pbuf, dflags).close()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            nbytes = 0
            for _ in range(reps):
                eb = enc.encode_host(built, kind, eflags, dst=wbuf)
                dec.decode_host_into(wbuf[: eb.total], eb.offsets, eb.lengths, kind, pbuf, dflags).close()
                nbytes += 2 * sum(eb.lengths)
            torch.cuda.synchronize()
            out[label] = round(nbytes / (time.perf_counter() - t0) / 1e9, 3)
        return out

    out = {}
    g = torch.Generator(device=dev)
    g.manual_seed(3001)

    # ---- C3: BERT input_ids + attention_mask, int64_val, 256 x 512 per request, 256 requests per launch
    R3 = 256
    ids = torch.randint(0, 30522, (R3, 256 * 512), device=dev, dtype=torch.int64, generator=g)
    lens = torch.randint(1, 513, (R3, 256, 1), device=dev, generator=g)
    mask = (torch.arange(512, device=dev)[None, None, :] < lens).to(torch.int64).reshape(R3, -1).contiguous()
    shape = tfs.TensorShapeProto([256, 512])

    def c3(idrow, mrow):
        return tfs.PredictRequest(tfs.ModelSpec("bert", signature_name="serving_default"),
                                  [("input_ids", tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=idrow)),
                                   ("attention_mask", tfs.TensorProto(dtype=abi.DT_INT64, tensor_shape=shape, int64_val=mrow))])
    r = device_side([c3(ids[i], mask[i]) for i in range(R3)], abi.MSG_PREDICT_REQUEST, 2 * R3 * 256 * 512 * 8)
    n_host = 32
    h_ids, h_mask = ids[:n_host].cpu().numpy(), mask[:n_host].cpu().numpy()
    host_msgs = [c3(h_ids[i], h_mask[i]) for i in range(n_host)]
    r.update(cpu_side(host_msgs, abi.MSG_PREDICT_REQUEST))
    r.update(e2e_side(host_msgs, abi.MSG_PREDICT_REQUEST))
    r["requests"], r["cpu_e2e_requests"] = R3, n_host
    out["C3"] = r
    del ids, mask, lens

    # ---- C4: DT_HALF half_val, 2^20 elements from f32 (decode widens back to f32), 64 tensors per launch
    R4 = 64
    x = torch.randn(R4, 1 << 20, device=dev, generator=g)

    def c4(row):
        return tfs.TensorProto(dtype=abi.DT_HALF, tensor_shape=tfs.TensorShapeProto([1 << 20]), half_val=tfs.Payload(row, abi.T_F32, abi.T_F16))
    r = device_side([c4(x[i]) for i in range(R4)], abi.MSG_TENSOR_PROTO, R4 * (1 << 22), dflags=abi.F_HALF_TO_F32)
    n_host = 16
    hx = x[:n_host].cpu().numpy()
    host_msgs = [c4(hx[i]) for i in range(n_host)]
    r.update(cpu_side(host_msgs, abi.MSG_TENSOR_PROTO))
    r.update(e2e_side(host_msgs, abi.MSG_TENSOR_PROTO, dflags=abi.F_HALF_TO_F32))
    r["requests"], r["cpu_e2e_requests"] = R4, n_host
    out["C4"] = r
    del x

    # ---- C5: one GPU's share (1 024 of 8 192) of mixed-dtype PredictResponses, decode; the wire comes from the oracle and
    #      the decoded values are checked against the oracle's once before anything is timed
    msgs5 = workloads.c5_responses(1024, seed=5001, max_log2=18)
    built5 = tfs.build_batch(msgs5, abi.MSG_PREDICT_RESPONSE)
    wire5, offs5, lens5 = ob.encode(built5)
    dwire5 = torch.from_numpy(wire5).to(dev)
    d = dec.scan(dwire5, offs5, lens5, abi.MSG_PREDICT_RESPONSE)
    pool5 = torch.empty(d.pool_bytes + 64, dtype=torch.uint8, device=dev)
    d.close()
    d = dec.decode_into(dwire5, offs5, lens5, abi.MSG_PREDICT_RESPONSE, pool5)
    ctx.sync()
    import helpers
    st_o, want = ob.decode(wire5, offs5, lens5, abi.MSG_PREDICT_RESPONSE)
    helpers.assert_same_decoded_fast(want, d.messages(), "response", "bench C5")
    d.close()
    del want, pool5
    r = device_side(None, abi.MSG_PREDICT_RESPONSE, 0, decode_wire=(dwire5, offs5, lens5), encode=False)
    r.update(cpu_side(msgs5, abi.MSG_PREDICT_RESPONSE, budget_s=2.0))
    r.update(e2e_side(msgs5, abi.MSG_PREDICT_RESPONSE))
    r["messages"] = 1024
    r["checked_against_oracle"] = 1
    out["C5"] = r
    del dwire5, wire5, msgs5, built5

    # ---- C2, back-to-back wire as tonic emits it (no TFS_F_ALIGN_PAYLOAD): payloads land at arbitrary byte phases
    R2 = args.requests
    src = torch.randn(R2, C2_ELEMS, device=dev, generator=g)
    shape2 = tfs.TensorShapeProto([64, 224, 224, 3])

    def c2(row, **kw):
        return tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"), [("input", tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=shape2, **kw))])
    r = device_side([c2(None, float_val=src[i]) for i in range(R2)], abi.MSG_PREDICT_REQUEST, R2 * C2_ELEMS * 4)
    n_host = 8
    hs = src[:n_host].cpu().numpy()
    host_msgs = [c2(None, float_val=hs[i]) for i in range(n_host)]
    r.update(cpu_side(host_msgs, abi.MSG_PREDICT_REQUEST))
    r.update(e2e_side(host_msgs, abi.MSG_PREDICT_REQUEST))
    r["requests"], r["cpu_e2e_requests"] = R2, n_host
    out["C2_unaligned"] = r

    # ---- C2 with the f32 -> bf16 narrowing fused into the encode copy (tensor_content, DT_BFLOAT16), widened on decode
    def c2bf(row):
        return tfs.PredictRequest(tfs.ModelSpec("resnet", signature_name="serving_default"),
                                  [("input", tfs.TensorProto(dtype=abi.DT_BFLOAT16, tensor_shape=shape2, tensor_content=tfs.Payload(row, abi.T_F32, abi.T_BF16)))])
    r = device_side([c2bf(src[i]) for i in range(R2)], abi.MSG_PREDICT_REQUEST, R2 * C2_ELEMS * 4, eflags=abi.F_ALIGN_PAYLOAD,
                    dflags=abi.F_CONTENT_HALF_TO_F32)
    host_msgs = [c2bf(hs[i]) for i in range(n_host)]
    r.update(cpu_side(host_msgs, abi.MSG_PREDICT_REQUEST, construct=True))  # (the narrowing loop is part of building the message)
    r.update(e2e_side(host_msgs, abi.MSG_PREDICT_REQUEST, eflags=abi.F_ALIGN_PAYLOAD, dflags=abi.F_CONTENT_HALF_TO_F32))
    r["requests"], r["cpu_e2e_requests"] = R2, n_host
    out["C2_f32_to_bf16"] = r
    del src

    # ---- C2 built from u8 pixels: `*p as f32` and `/ 255.` fused into the encode copy (src/lib.rs:237-242), encode only
    px = torch.randint(0, 256, (R2, C2_ELEMS), device=dev, dtype=torch.uint8, generator=g)

    def c2u8(row):
        return c2(None, float_val=tfs.Payload(row, abi.T_U8, abi.T_F32, abi.OP_DIV, 255.0))
    r = device_side([c2u8(px[i]) for i in range(R2)], abi.MSG_PREDICT_REQUEST, R2 * C2_ELEMS, eflags=abi.F_ALIGN_PAYLOAD, decode=False)
    hp = px[:n_host].cpu().numpy()
    host_msgs = [c2u8(hp[i]) for i in range(n_host)]
    r.update(cpu_side(host_msgs, abi.MSG_PREDICT_REQUEST, construct=True))
    r.update(e2e_side(host_msgs, abi.MSG_PREDICT_REQUEST, eflags=abi.F_ALIGN_PAYLOAD))
    r["requests"], r["cpu_e2e_requests"] = R2, n_host
    out["C2_u8_to_f32"] = r
    del px

    # ---- C1: the half_plus_two request (63 wire bytes), 2^20 copies in one batch: descriptor / latency bound by construction
    N1 = args.c1_messages
    x3 = torch.tensor([1.0, 2.0, 5.0], device=dev)
    built1 = tfs.replicate_batch(tfs.build_batch([tfs.PredictRequest(tfs.ModelSpec("half_plus_two", signature_name="serving_default"),
                                 [("x", tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=tfs.TensorShapeProto([3]), float_val=x3))])], abi.MSG_PREDICT_REQUEST), N1)
    r = device_side(built1, abi.MSG_PREDICT_REQUEST, 12 * N1, reps=2)
    r["enc_Mmsgs_per_s"] = round(N1 / (r["enc_ms"] * 1e-3) / 1e6, 2)
    r["dec_Mmsgs_per_s"] = round(N1 / (r["dec_ms"] * 1e-3) / 1e6, 2)
    n_host = 1 << 16
    hx3 = np.array([1.0, 2.0, 5.0], np.float32)
    hb1 = tfs.replicate_batch(tfs.build_batch([tfs.PredictRequest(tfs.ModelSpec("half_plus_two", signature_name="serving_default"),
                              [("x", tfs.TensorProto(dtype=abi.DT_FLOAT, tensor_shape=tfs.TensorShapeProto([3]), float_val=hx3))])], abi.MSG_PREDICT_REQUEST), n_host)
    p = ob.Prepared(hb1)
    for label, th in (("cpu1", 1), ("cpuN", threads)):
        s_e, wb = p.run_encode(th)
        s_d, _, _ = p.run_decode(th)
        r[f"{label}_GBps"] = round(2 * wb / (s_e + s_d) / 1e9, 3)
        r[f"{label}_enc_Mmsgs_per_s"] = round(n_host / s_e / 1e6, 2)
        r[f"{label}_dec_Mmsgs_per_s"] = round(n_host / s_d / 1e6, 2)
    p.close()
    r["cpu_threads"] = threads
    r["messages"], r["cpu_messages"] = N1, n_host
    out["C1"] = r
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


def cpu_baseline_c2(budget_s=12.0):
    threads = host_threads()
    n_msgs = threads
    gbs, reps, secs = cpu_pass(threads, n_msgs, budget_s)
    gbs1, _, _ = cpu_pass(1, 1, min(2.0, budget_s))
    return {"value": round(gbs, 3), "unit": UNIT, "cores": threads, "kind": "port", "one_thread_GBps": round(gbs1, 3),
            "sample": f"{n_msgs} C2 requests per variant (one per thread), encode+decode, {reps} passes, {secs:.1f} s; oracle/cpu_ref.cpp (C++ port of prost 0.6.0)",
            "host_cores_available": os.cpu_count() or 1}


def run_reference(args):
    """The reference's CPU path for the same metric: prost's algorithm as oracle/cpu_ref.cpp restates it (the Rust reference
    cannot be built here: no rustc / cargo, crates not vendored), every host core, W warm-up + K timed steps; a step is a
    bounded sample of the workload: one C2 request per host thread and variant, encoded then decoded."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    n_msgs = threads
    t0 = time.perf_counter()
    for _ in range(max(0, args.warmup)):
        cpu_pass(threads, n_msgs, 0.0)
    steps = max(1, args.steps)
    tot_b = tot_s = 0.0
    for _ in range(steps):
        gbs, reps, secs = cpu_pass(threads, n_msgs, 0.0)
        tot_b += gbs * secs
        tot_s += secs
    value = tot_b / tot_s
    sample = f"{n_msgs} C2 requests per variant per step (one per host thread); oracle/cpu_ref.cpp, C++ port of prost 0.6.0"
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": max(0, args.warmup), "ms_per_step": round(tot_s / steps * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config_of(args, args.gpus),
            "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                             "host_cores_available": os.cpu_count() or 1},
            "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    print(f"reference arm wall time {time.perf_counter() - t0:.1f} s", file=sys.stderr)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--requests", type=int, default=32, help="C2 requests per variant, per GPU and per step")
    ap.add_argument("--e2e-requests", type=int, default=32)
    ap.add_argument("--cpu-budget", type=float, default=8.0, help="seconds of CPU baseline sampling")
    ap.add_argument("--sustain-s", type=float, default=1.5, help="length of the sustained C2 side loop (0 = skip)")
    ap.add_argument("--c1-messages", type=int, default=1 << 20)
    ap.add_argument("--no-extras", action="store_true", help="skip the per-config side measurements (N = 1 only)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
