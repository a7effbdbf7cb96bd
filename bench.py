#!/usr/bin/env python
"""Headline benchmark: pipelined ResNet-50 inference through the DEFER API on N B200s.

    python bench.py --gpus 1 --steps 30 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference --gpus N --steps K --warmup W      # restated CPU pipeline

One "step" = one micro-batch (default 128 images, BASELINE config 4's micro-batch) through the
whole pipeline.  Rank 0 prints ONE JSON line:
  value     images/s with the input batch already resident in the first stage's HBM slots, timed
            with CUDA events on every stage's stream, max over stages/ranks (first window of K steps;
            two more windows and a seconds-long window are reported under `checks`);
  e2e       images/s through the public API (input_q -> DEFER -> output_q) with pinned host
            inputs: host->device copy of every step's inputs and device->host read of every
            step's (B, 1000) result inside the timed region (wall clock);
  roofline  the dominant kernel class (conv_gemm_kernel: every Conv2D/Dense of the bottleneck stage)
            timed INSIDE a CUDA graph of its own (programmatic dependent launch overlap kept), against the
            measured bf16 tensor peak (burst figure: the timed window is tens of milliseconds);
  checks    every delivered micro-batch of both legs against the un-partitioned model on one GPU (bit for bit:
            the cuts are fused-group boundaries);
  baseline_configs   BASELINE.json's other configurations measured in the same run: config 1's B200 side
            (batch 1, also timed the reference's own way) at N=1, config 2 at N=2, config 3 at N=4;
  cpu_baseline  the restated reference CPU pipeline (oracle/cpu_pipeline.py) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import queue
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

FLOP_PER_IMAGE = 2 * 3_857_973_248  # Keras ResNet-50 v1, SURVEY 8(d)
CONFIG_PART_AT = {
    1: [],
    2: ["conv3_block4_out"],                                                 # BASELINE config 2
    4: ["conv2_block3_out", "conv3_block4_out", "conv4_block6_out"],         # BASELINE config 3
}
CONFIG_MICRO_BATCH = {1: 1, 2: 32, 4: 64}                                    # BASELINE configs 1, 2, 3
NCU_CAPTURE = os.path.join("profiles", "r2_conv_full_metrics.csv")


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"bf16_sustained": p.get("bf16_tflops_sustained", 1388.0), "bf16_burst": p.get("bf16_tflops", 1653.4),
                "hbm": p.get("hbm_gbs", 6548.5), "source": "measured"}
    return {"bf16_sustained": 1400.0, "bf16_burst": 1590.0, "hbm": 6650.0, "source": "fallback"}


def ncu_traffic_per_launch(expect_launches: int):
    """Mean DRAM bytes (read + write) per conv launch from the committed `ncu --set full` capture of one N=1 step --
    only when that capture holds exactly the launches the current build makes (else the numbers describe another
    kernel list and null is reported)."""
    import csv

    path = os.path.join(ROOT, NCU_CAPTURE)
    if not os.path.exists(path):
        return None, "no capture committed for this build"
    rows = list(csv.reader(open(path)))
    col = {h: i for i, h in enumerate(rows[0])}
    unit = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
    tot, n = 0.0, 0
    for r in rows[2:]:
        try:
            tot += sum(float(r[col[k]]) * unit.get(rows[1][col[k]], 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            n += 1
        except (ValueError, KeyError, IndexError):
            continue
    if n != expect_launches:
        return None, f"{NCU_CAPTURE} holds {n} launches, this build makes {expect_launches} per step"
    return round(tot / n), f"mean dram__bytes_read.sum + dram__bytes_write.sum per conv launch, {NCU_CAPTURE} (cold-cache ncu replays)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self):
        self.path = tempfile.mktemp(prefix="adapt_clocks_", suffix=".csv")
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [c.strip() for c in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        try:
            os.unlink(self.path)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # samples under load = those drawing well above idle power (the timed legs, the seconds-long window)
        load = [s for s, p in zip(sm, pw) if p >= 0.6 * max(pw)] or sorted(sm)[len(sm) // 2:]
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "samples_under_load": len(load), "power_w_max": max(pw)}


def part_at_for(model, n: int, cuts: str = "balanced"):
    """Cut points of the N-stage bench line.  "balanced" (default): the time-balanced partition the dispatcher
    computes from its roofline cost model; "config": the split points BASELINE.json's configs 2 and 3 name for 2 and
    4 stages (block-group boundaries, unbalanced -- they are also measured, in `baseline_configs`)."""
    from adapt_b200.dispatcher import balanced_part_at

    if n == 1:
        return []
    if os.environ.get("ADAPT_BENCH_PART_AT"):      # developer override: comma-separated layer names
        return os.environ["ADAPT_BENCH_PART_AT"].split(",")
    if cuts == "config" and n in CONFIG_PART_AT:
        return CONFIG_PART_AT[n]
    return balanced_part_at(model, n)


def workload_name(n: int, batch: int, part_at) -> str:
    return (f"ResNet-50 (Keras v1, random init) {n}-stage pipeline, micro-batch {batch}, bf16 activations / fp32 accumulate, "
            f"part_at={part_at if part_at else '[] (single chunk)'}")


def bench_config(n: int, batch: int, part_at) -> dict:
    """`config` of the JSON line -- the same dict in both arms (what is run, not how it is timed)."""
    return {"workload": workload_name(n, batch, part_at), "part_at": list(part_at), "micro_batch": batch}


# ---------------------------------------------------------------------------------------------
def run_reference(args) -> int:
    """Reference arm: the restated reference CPU pipeline on the box's host cores (rank 0 only), on the arm's own
    configuration: same cuts, same micro-batch."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    import adapt_b200
    from oracle.cpu_pipeline import time_cpu_pipeline

    model = adapt_b200.ResNet50(weights=None)
    weights = {l.name: l.get_weights() for l in model.layers if l.get_weights()}
    part_at = part_at_for(model, args.gpus, args.cuts)
    B = args.ref_batch or args.batch
    r = time_cpu_pipeline(weights, part_at, B, args.steps, args.warmup)
    # the reference's own way of timing (test/test.py:25-37): clock started before set-up, 10 x batch-1 items
    ref_style = reference_style_cpu(weights, part_at)
    line = {
        "impl": "reference", "metric": "images_per_sec", "value": r["value"], "unit": "img/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["seconds"] / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(args.gpus, B, part_at),
        "cpu_baseline": {"value": r["value"], "unit": "img/s", "cores": r["cores"], "kind": "port",
                         "sample": f"{args.steps} micro-batches of {B} images through {r['stages']} localhost worker "
                                   f"process(es), {r['threads_per_stage']} torch threads each; restated reference CPU "
                                   f"pipeline, TensorFlow unavailable, lossless zfp+lz4 codec omitted"},
        "e2e": {"value": r["value"], "unit": "img/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_style": ref_style,
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def reference_style_cpu(weights, part_at) -> dict:
    """10 batch-1 inputs, wall clock from BEFORE the pipeline is set up to the last result (reference test/test.py:25-37)."""
    from oracle import keras_resnet50_oracle as KO
    from oracle.cpu_pipeline import CpuPipeline

    x = KO.synthetic_images(1, seed=0, kind="caffe")
    t0 = time.perf_counter()
    pipe = CpuPipeline(weights, part_at)
    try:
        pipe.run([x] * 10)
        dt = time.perf_counter() - t0
    finally:
        pipe.close()
    return {"value": round(10 / dt, 3), "unit": "img/s", "items": 10, "batch": 1, "seconds": round(dt, 3),
            "clock": "started before partitioning / worker set-up, stopped at the 10th result (reference test/test.py:25-37)"}


# ---------------------------------------------------------------------------------------------
def drive_steady(defer, in_q, out_q, items, warm, timeout=60.0):
    """Feed `items` back to back; returns (results, seconds between the arrival of result `warm` - 1 and the last
    one), i.e. the steady-state time of the last len(items) - warm steps with the pipeline kept full -- every one of
    those steps has its host->device copy and its device->host result read inside the region -- and the wall time
    of the whole drive (fill and drain included) for reference."""
    timeout = float(os.environ.get("ADAPT_BENCH_DRIVE_TIMEOUT_S", timeout))
    t = threading.Thread(target=lambda: [in_q.put(x) for x in items], daemon=True)
    t0 = time.perf_counter()
    t.start()
    outs, stamps = [], []
    while len(outs) < len(items):
        try:
            outs.append(out_q.get(timeout=0.5))
            stamps.append(time.perf_counter())
        except queue.Empty:
            if defer.error:
                raise RuntimeError(defer.error)
            if time.perf_counter() - t0 > timeout:
                raise TimeoutError(f"bench: {len(outs)}/{len(items)} results; {defer.debug_state()}")
    t.join()
    return outs, stamps[-1] - stamps[warm - 1], stamps[-1] - t0


def drive(defer, in_q, out_q, items, timeout=60.0):
    """Feed `items` and collect as many results; returns (results, wall seconds)."""
    timeout = float(os.environ.get("ADAPT_BENCH_DRIVE_TIMEOUT_S", timeout))  # (profiler runs need minutes per step)
    t = threading.Thread(target=lambda: [in_q.put(x) for x in items], daemon=True)
    t0 = time.perf_counter()
    t.start()
    outs = []
    while len(outs) < len(items):
        try:
            outs.append(out_q.get(timeout=0.5))
        except queue.Empty:
            if defer.error:
                raise RuntimeError(defer.error)
            if time.perf_counter() - t0 > timeout:
                raise TimeoutError(f"bench: {len(outs)}/{len(items)} results; {defer.debug_state()}")
    dt = time.perf_counter() - t0
    t.join()
    return outs, dt


class Session:
    """One DEFER session of the bench (rank 0): dispatcher thread + queues."""

    def __init__(self, model, devices, part_at, batch, args, **kw):
        from adapt_b200.dispatcher import DEFER

        n = len(devices)
        self.model, self.part_at, self.B, self.n = model, list(part_at), batch, n
        # (a profiler replaying one kernel dozens of times stalls the stage far beyond the 50 ms liveness
        # window: profiling runs widen it with ADAPT_HEARTBEAT_TIMEOUT_S; numbers are never taken from those runs)
        self.defer = DEFER(devices, mode="bf16", max_batch=batch, warm_batches=[batch],
                           heartbeat_timeout=float(os.environ.get("ADAPT_HEARTBEAT_TIMEOUT_S", "0.05")),
                           max_inflight=max(2 * n + 4, 12), ingest_rails=args.ingest_rails, **kw)
        self.in_q, self.out_q = queue.Queue(2 * n + 8), queue.Queue()
        self.th = threading.Thread(target=self.defer.run_defer, args=(model, self.part_at, self.in_q, self.out_q), daemon=True)
        self.th.start()
        self.defer.wait_ready(600)

    def resident_window(self, K):
        """K steps on the preloaded input; (results, per-stage device ms over the window)."""
        from adapt_b200.dispatcher import ResidentBatch

        d = self.defer
        k0 = d.feed_index
        d.set_timing_window(k0, k0 + K)
        outs, wall = drive(d, self.in_q, self.out_q, [ResidentBatch(self.B)] * K)
        return outs, d.collect_timing(), wall

    def close(self):
        self.defer.shutdown()
        self.th.join(60)


def synthetic_batch(np, B, seed):
    rng = np.random.default_rng(seed)
    x = rng.uniform(0, 255, (B, 224, 224, 3)).astype(np.float32)
    return np.ascontiguousarray(x[..., ::-1] - np.array([103.939, 116.779, 123.68], dtype=np.float32))


def measure_small_config(model, devices, part_at, batch, args, K, W, np, torch) -> dict:
    """One of BASELINE.json's other configurations in a session of its own: resident-input rate (CUDA events per
    stage), end-to-end rate with pinned host inputs, per-stage times, values checked against the un-partitioned model."""
    from adapt_b200.dispatcher import ResidentBatch
    from adapt_b200.executor import LocalPredictor

    s = Session(model, devices, part_at, batch, args)
    try:
        x = synthetic_batch(np, batch, 1)
        s.defer.preload_input(x)
        drive(s.defer, s.in_q, s.out_q, [ResidentBatch(batch)] * W)
        outs, stage_ms, _ = s.resident_window(K)
        pinned = [torch.from_numpy(synthetic_batch(np, batch, 10 + i)).pin_memory() for i in range(3)]
        s.defer.latencies_ms.clear()
        e_outs, e_wall, _ = drive_steady(s.defer, s.in_q, s.out_q, [pinned[i % 3] for i in range(W + K)], W)
        s.defer.latencies_ms.clear()
        for i in range(min(10, K)):
            drive(s.defer, s.in_q, s.out_q, [pinned[i % 3]])
        lat = sorted(s.defer.latencies_ms)
    finally:
        s.close()
    local = LocalPredictor(model, mode="bf16", max_batch=batch)
    want = [local.run(p.numpy()) for p in pinned]
    want_res = local.run(x)
    local.close()
    bad = sum(0 if np.array_equal(o, want[i % 3]) else 1 for i, o in enumerate(e_outs))
    bad += sum(0 if np.array_equal(o, want_res) else 1 for o in outs)
    ms = max(stage_ms)
    return {"part_at": list(part_at), "micro_batch": batch, "value": round(K * batch / (ms / 1e3), 1), "unit": "img/s",
            "ms_per_step": round(ms / K, 4), "stage_ms_per_step": [round(m / K, 4) for m in stage_ms],
            "e2e": round(K * batch / e_wall, 1), "p50_latency_ms": round(lat[len(lat) // 2], 3) if lat else None,
            "mismatches_vs_unpartitioned": int(bad)}


def reference_style_ours(model, devices, part_at, np) -> dict:
    """The reference driver's own measurement (test/test.py:25-50) on the B200 path: wall clock started BEFORE
    DEFER is constructed (partitioning, weight upload and graph capture inside), 10 batch-1 items, stopped at the
    10th result."""
    from adapt_b200.dispatcher import DEFER

    x = synthetic_batch(np, 1, 0)
    t0 = time.perf_counter()
    d = DEFER(devices, mode="bf16", max_batch=1)
    in_q, out_q = queue.Queue(10), queue.Queue(10)
    th = threading.Thread(target=d.run_defer, args=(model, part_at, in_q, out_q), daemon=True)
    th.start()
    for _ in range(10):
        in_q.put(x)
    for _ in range(10):
        out_q.get(timeout=120)
    dt = time.perf_counter() - t0
    d.shutdown()
    th.join(60)
    return {"value": round(10 / dt, 2), "unit": "img/s", "items": 10, "batch": 1, "seconds": round(dt, 3),
            "clock": "started before DEFER() / partitioning / weight upload / graph capture, stopped at the 10th result "
                     "(reference test/test.py:25-37)"}


def replica_comparison(args, model, rank, world, want0, np, torch) -> dict:
    """SURVEY 8e's comparison line (every rank of a torchrun launch calls this after the pipeline sessions ended):
    the same N GPUs as R independent replicas of an S-stage pipeline, R x S = N, i.e. what the reference's pure
    pipeline design costs against data parallelism on this box.  Rank r with r % S == 0 leads a replica: an
    in-process DEFER session over GPUs r .. r + S - 1 (its own dispatcher, queues and ingest), the other ranks idle.
    All replicas start their timed window together (barrier); value = R * K * B / the slowest replica's device time."""
    import torch.distributed as dist

    n, B, K, W = args.gpus, args.batch, args.steps, args.warmup
    os.environ["ADAPT_MULTIPROCESS"] = "0"    # sessions below are in-process (worker threads), not rendezvous
    out = {}
    for S in sorted({1, n // 2}):
        R = n // S
        lead = rank % S == 0
        rec = None
        if lead:
            devices = [f"cuda:{rank + i}" for i in range(S)]
            part_at = part_at_for(model, S, args.cuts)
            ses = Session(model, devices, part_at, B, args)
            x = synthetic_batch(np, B, 0)
            ses.defer.preload_input(x)
            from adapt_b200.dispatcher import ResidentBatch

            drive(ses.defer, ses.in_q, ses.out_q, [ResidentBatch(B)] * W)
        dist.barrier()
        if lead:
            outs, stage_ms, _ = ses.resident_window(K)
            pinned = [torch.from_numpy(synthetic_batch(np, B, i)).pin_memory() for i in range(3)]
        dist.barrier()
        if lead:
            e_outs, e_wall, _ = drive_steady(ses.defer, ses.in_q, ses.out_q, [pinned[i % 3] for i in range(W + K)], W)
            ses.close()
            rec = {"ms": max(stage_ms), "e2e_wall": e_wall, "part_at": part_at, "last": outs[-1], "e2e_first": e_outs[0]}
        got = [None] * world
        dist.all_gather_object(got, rec)
        if rank == 0:
            reps = [g for g in got if g is not None]
            ms = max(g["ms"] for g in reps)
            wall = max(g["e2e_wall"] for g in reps)
            bad = sum(0 if (np.array_equal(g["last"], want0) and np.array_equal(g["e2e_first"], want0)) else 1 for g in reps)
            out[f"{R}x{S}"] = {"replicas": R, "stages": S, "part_at": reps[0]["part_at"], "value": round(R * K * B / (ms / 1e3), 1),
                               "ms_per_step_slowest_replica": round(ms / K, 4), "e2e": round(R * K * B / wall, 1), "unit": "img/s",
                               "replicas_with_wrong_values": int(bad)}
    os.environ.pop("ADAPT_MULTIPROCESS", None)
    if rank == 0:
        out["note"] = ("replicas x stages on the same GPUs; the headline `value` is the reference's design (1 x N: one "
                       "pipeline over all N GPUs); e2e = pinned fp32 host batches per replica, slowest replica's wall clock")
    return out


def run_ours(args) -> int:
    import numpy as np
    import torch

    import adapt_b200
    from adapt_b200 import _native as N
    from adapt_b200.dispatcher import ResidentBatch, end_sessions
    from adapt_b200.executor import LocalPredictor
    from adapt_b200.node import Node

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    N.require_gpu()  # no CPU fallback: fail loudly
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group(backend="cpu:gloo,cuda:nccl", init_method="env://")
        if world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    n = args.gpus
    B = args.batch
    model = adapt_b200.ResNet50(weights=None)
    part_at = part_at_for(model, n, args.cuts)

    if world > 1 and rank > 0:
        Node(model, device=local_rank).run()      # worker role: serves every session of the bench, then returns
        if not args.no_replica_comparison:
            replica_comparison(args, model, rank, world, None, np, torch)
        dist.barrier()
        dist.destroy_process_group()
        return 0

    devices = [f"cuda:{i}" for i in range(n)]
    K, W = args.steps, args.warmup
    peaks = measured_peaks()
    clocks = ClockSampler()
    main = Session(model, devices, part_at, B, args)
    defer, in_q, out_q = main.defer, main.in_q, main.out_q

    # ---------------- leg 1: device-resident inputs, CUDA-event timing on every stage ----------------
    x = synthetic_batch(np, B, 0)
    defer.preload_input(x)
    clocks.start()  # sampled from the warm-up through every timed leg of the main session
    # clock ramp: a few hundred steps before the W warm-up steps the contract asks for, so that the timed window does
    # not start on GPUs that were idle a few milliseconds earlier (W steps are 1.5-8 ms of work; the first window of an
    # 8-stage run measured 7 % below the following ones without this)
    n_ramp = int(args.ramp_s / 0.4e-3) if args.ramp_s > 0 else 0
    for i in range(0, n_ramp, 64):
        drive(defer, in_q, out_q, [ResidentBatch(B)] * min(64, n_ramp - i))
    drive(defer, in_q, out_q, [ResidentBatch(B)] * W)
    torch.cuda.synchronize()
    launches0 = defer.gpu_launches
    outs, stage_ms, wall_resident = main.resident_window(K)
    handoff_gbs = list(defer.handoff_gbs)
    stage_busy_ms = [round(v, 4) for v in defer.stage_busy_ms]
    torch.cuda.synchronize()
    local_launches = defer.gpu_launches - launches0
    ms_total = max(stage_ms)
    value = K * B / (ms_total / 1e3)
    resident_outs = list(outs)
    # spread: two more windows of K steps, and one seconds-long window (sustained clocks / power, enough
    # nvidia-smi samples under load)
    repeats = [round(value, 1)]
    for _ in range(2):
        o2, sm2, _ = main.resident_window(K)
        repeats.append(round(K * B / (max(sm2) / 1e3), 1))
        resident_outs += o2[-1:]
    k_long = max(K, int(args.sustained_s / (ms_total / K / 1e3)))
    o3, sm3, _ = main.resident_window(k_long)
    sustained = {"seconds": round(max(sm3) / 1e3, 3), "steps": k_long, "value": round(k_long * B / (max(sm3) / 1e3), 1), "unit": "img/s"}
    resident_outs += o3[-1:]

    # ---------------- leg 2: end to end through the public API, pinned host inputs ----------------
    host_inputs = [x, synthetic_batch(np, B, 1), synthetic_batch(np, B, 2)]   # three distinct batches, round robin
    pinned = [torch.from_numpy(h).pin_memory() for h in host_inputs]
    # W warm-up items and K timed items flow back to back: the timed region starts when the last warm-up result
    # arrives (pipeline full) and ends with the K-th timed result
    defer.latencies_ms.clear()
    e2e_outs, e2e_wall, e2e_wall_all = drive_steady(defer, in_q, out_q, [pinned[i % 3] for i in range(W + K)], W)
    e2e_value = K * B / e2e_wall
    clock_info = clocks.stop()
    lat = sorted(defer.latencies_ms)
    p50 = lat[len(lat) // 2] if lat else None
    p99 = lat[min(len(lat) - 1, int(len(lat) * 0.99))] if lat else None
    # unloaded latency: one micro-batch in flight at a time (queue put -> result get)
    defer.latencies_ms.clear()
    for i in range(min(10, K)):
        drive(defer, in_q, out_q, [pinned[i % 3]])
    lat1 = sorted(defer.latencies_ms)
    p50_unloaded = lat1[len(lat1) // 2] if lat1 else None

    # ---------------- roofline of the dominant kernel class, on rank 0's stage ----------------
    # (1) the conv launches of the stage as a CUDA graph of their own (PDL overlap kept), timed with one pair of
    # events over 20 graph launches: the duration of the class INSIDE a step; (2) the per-launch table (events
    # around single launches, no overlap) for the per-layer breakdown only.
    wk = defer.workers[0]
    conv_ms, n_conv = wk.executor.time_kind(0, B, (N.OP_CONV, N.OP_CONV_CHAIN, N.OP_STEM_POOL), wk.stream, reps=20)
    prof = wk.executor.profile(0, B, wk.stream, reps=5)
    conv = [p for p in prof if p["kind"] in (N.OP_CONV, N.OP_CONV_CHAIN, N.OP_STEM_POOL)]
    conv_flop = 2.0 * sum(p["macs"] for p in conv)
    iso_conv_ms = sum(p["ms"] for p in conv)
    iso_all_ms = sum(p["ms"] for p in prof)
    stage0_ms = stage_ms[0] / K
    achieved_tf = conv_flop / (conv_ms / 1e3) / 1e12
    traffic, traffic_note = ncu_traffic_per_launch(n_conv) if n == 1 else (None, "captured at N=1 only")
    roofline = {
        "bound": "tensor", "kernel": "conv_gemm_kernel + conv_chain_kernel + stem_pool_kernel", "achieved": round(achieved_tf, 2), "peak": peaks["bf16_burst"],
        "unit": "TFLOP/s", "frac": round(achieved_tf / peaks["bf16_burst"], 4), "traffic": traffic, "traffic_note": traffic_note,
        "peak_source": f"{peaks['source']} bf16 burst (the timed window lasts {ms_total:.0f} ms; sustained figure "
                       f"{peaks['bf16_sustained']} TFLOP/s -> frac {achieved_tf / peaks['bf16_sustained']:.4f})",
        "stage": 0, "launches_per_step": n_conv, "conv_ms_in_step": round(conv_ms, 5), "stage_ms_per_step": round(stage0_ms, 5),
        "avg_launch_ms": round(conv_ms / max(1, n_conv), 5), "kernel_share_of_step": round(conv_ms / stage0_ms, 4),
        "how": "conv launches of rank 0's stage replayed as their own CUDA graph (same launch attributes, programmatic "
               "dependent launch overlap kept), CUDA events around 20 graph launches on the stage's stream",
        "isolated_launch_sum_ms": round(iso_conv_ms, 5), "isolated_all_ops_ms": round(iso_all_ms, 5),
        "algorithmic_flop_per_launch_avg": conv_flop / max(1, n_conv),
        "algorithmic_hbm_gbs": round(50.5e6 * value / 1e9, 1) if n == 1 else None,
        "whole_pipeline_frac": round(value * FLOP_PER_IMAGE / 1e12 / (n * peaks["bf16_burst"]), 4),
        "whole_pipeline_frac_sustained_peak": round(value * FLOP_PER_IMAGE / 1e12 / (n * peaks["bf16_sustained"]), 4),
        "whole_pipeline_frac_4p1gflop": round(value * 4.1e9 / 1e12 / (n * peaks["bf16_burst"]), 4),
    }
    if args.dump_ops:
        os.makedirs(os.path.dirname(os.path.abspath(args.dump_ops)), exist_ok=True)
        json.dump({"batch": B, "stage": 0, "ops": prof, "conv_ms_in_graph": conv_ms}, open(args.dump_ops, "w"), indent=1)

    # ---------------- fail-over leg (needs a stage to lose) ----------------
    failover = None
    if n >= 2 and not args.no_failover:
        # BASELINE config 5: the kill lands mid-stream, after >= 100 micro-batches, with more queued behind it than
        # the pipeline can hold in flight (otherwise every result may already be out when the stage dies)
        n_before, n_after = 100, 6 * n + 40
        items = [pinned[i % 3] for i in range(n_before + n_after)]
        t = threading.Thread(target=lambda: [in_q.put(p) for p in items], daemon=True)
        t.start()
        fo_outs = []
        t_kill = None
        while len(fo_outs) < len(items):
            fo_outs.append(out_q.get(timeout=120))
            if len(fo_outs) == n_before and t_kill is None:
                defer.inject_kill(n // 2)
                t_kill = time.perf_counter()
        t.join()
        failover = {"failover_ms": defer.stats["failover_ms"][-1] if defer.stats["failover_ms"] else None,
                    "definition": "kill injection -> first result delivered by the re-sliced pipeline",
                    "detect_ms": defer.stats.get("detect_ms", [None])[-1],
                    "phases_ms": defer.stats.get("reconfig_phases_ms", [None])[-1],
                    "reconfig_ms": defer.stats.get("reconfig_ms", [None])[-1], "replayed": defer.stats["replayed"],
                    "delivered_exactly_once": len(fo_outs) == len(items) and out_q.empty(), "stages_after": len(defer.order),
                    "part_at_after": list(defer.part_at), "killed_after_micro_batches": n_before}

    from adapt_b200.dag_util import build_weight_table, compile_part

    _, windex = build_weight_table(model)
    launches_per_step = sum(len(compile_part(p, windex).ops) for p in defer._partition(model, part_at))
    assert world > 1 or local_launches == launches_per_step * K, (local_launches, launches_per_step, K)
    main.close()

    # ---------------- values: everything delivered above against the un-partitioned model on one GPU ----------------
    # (every cut the bench uses is a residual block's output or pool1_pool, i.e. a fused-group boundary: the
    # pipeline runs the same kernels on the same data as one stage holding the whole model -> bit for bit)
    local = LocalPredictor(model, mode="bf16", max_batch=B)
    want = [local.run(h) for h in host_inputs]
    local.close()
    diffs = [float(np.abs(o - want[0]).max()) for o in resident_outs]
    diffs += [float(np.abs(o - want[i % 3]).max()) for i, o in enumerate(e2e_outs)]
    mism = sum(0 if np.array_equal(o, want[0]) else 1 for o in resident_outs)
    mism += sum(0 if np.array_equal(o, want[i % 3]) else 1 for i, o in enumerate(e2e_outs))
    checks = {"max_abs_vs_unpartitioned": max(diffs), "mismatched_micro_batches": int(mism),
              "compared_micro_batches": len(resident_outs) + len(e2e_outs),
              "e2e_first_last_match": bool(np.array_equal(e2e_outs[0], want[0]) and np.array_equal(e2e_outs[-1], want[(W + K - 1) % 3])),
              "distinct_inputs_distinct_outputs": bool(not np.array_equal(want[0], want[1])),
              "probs_row_sum": float(np.float64(outs[-1]).sum()) / B, "wall_resident_img_s": round(K * B / wall_resident, 1),
              "value_repeat": {"windows": repeats, "min": min(repeats), "max": max(repeats)}, "sustained": sustained}
    if failover is not None:
        fo_bad = sum(0 if np.array_equal(o, want[i % 3]) else 1 for i, o in enumerate(fo_outs))
        failover["mismatched_micro_batches"] = int(fo_bad)   # across the kill, the re-slice and the replay

    # ---------------- BASELINE.json's other configurations, same run ----------------
    baseline_configs = {}
    if not args.no_baseline_configs:
        if n in (2, 4):
            key = "config2" if n == 2 else "config3"
            baseline_configs[key] = measure_small_config(model, devices, CONFIG_PART_AT[n], CONFIG_MICRO_BATCH[n], args, K, W, np, torch)
        if n == 1:
            c1 = measure_small_config(model, devices, [], 1, args, K, W, np, torch)
            c1["reference_style"] = reference_style_ours(model, devices, [], np)
            baseline_configs["config1_b200"] = c1

    # ---------------- extra leg: raw uint8 inputs, preprocessing fused into the ingest kernel ----------------
    # (not the reference's queue contract, hence not `e2e`: reported beside it)
    e2e_u8 = None
    if not args.no_uint8_leg:
        s8 = Session(model, devices, part_at, B, args, input_format="uint8")
        rng8 = np.random.default_rng(0)
        raw = [torch.from_numpy(rng8.integers(0, 256, (B, 224, 224, 3), dtype=np.uint8)).pin_memory() for _ in range(3)]
        o8, wall8, _ = drive_steady(s8.defer, s8.in_q, s8.out_q, [raw[i % 3] for i in range(W + K)], W)
        s8.close()
        l8 = LocalPredictor(model, mode="bf16", max_batch=B, input_format="uint8")
        w8 = [l8.run(r.numpy()) for r in raw]
        l8.close()
        e2e_u8 = {"value": round(K * B / wall8, 1), "unit": "img/s", "h2d_bytes_per_step": int(raw[0].numel()),
                  "d2h_bytes_per_step": B * 1000 * 4,
                  "mismatched_micro_batches": int(sum(0 if np.array_equal(o, w8[i % 3]) else 1 for i, o in enumerate(o8))),
                  "input": "raw uint8 RGB (B,224,224,3), caffe preprocess_input fused into the ingest kernel"}
    end_sessions()
    replicas = None
    if world > 1 and not args.no_replica_comparison:
        replicas = replica_comparison(args, model, rank, world, want[0], np, torch)

    # ---------------- CPU baseline on a bounded sample (N=1 only) ----------------
    cpu = None
    if n == 1 and not args.no_cpu_baseline:
        from oracle.cpu_pipeline import time_cpu_pipeline

        weights = {l.name: l.get_weights() for l in model.layers if l.get_weights()}
        rb = args.ref_batch or B
        r = time_cpu_pipeline(weights, [], rb, args.cpu_steps, 1)
        cpu = {"value": round(r["value"], 2), "unit": "img/s", "cores": r["cores"], "kind": "port",
               "sample": f"{args.cpu_steps} micro-batches of {rb} images through the restated reference CPU "
                         f"pipeline (1 localhost worker, torch CPU fp32, {r['threads_per_stage']} threads), {r['seconds']:.1f} s",
               "reference_style": reference_style_cpu(weights, [])}

    line = {
        "metric": "images_per_sec", "value": round(value, 1), "unit": "img/s", "n_gpus": n, "steps": K, "warmup": W,
        "ms_per_step": round(ms_total / K, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": bench_config(n, B, part_at),
        "timing": {"l2": "per-step working set (activations of 53 convolutions x micro-batch, > 5 GB at B=128) far exceeds the "
                         "126 MB L2; input slots alternate", "handoff": "peer stores / copy engines over NVLink + stream memory ops",
                   "stage_ms_per_step": [round(m / K, 4) for m in stage_ms], "how": "CUDA events per stage stream, max over stages",
                   "untimed_before_window": f"{n_ramp} clock-ramp steps + {W} warm-up steps",
                   "stage_busy_ms": stage_busy_ms,
                   "stage_busy_how": "moving average of CUDA-event time around each stage's graph launch alone (no waits, no hand-off)",
                   "handoff_gbs": handoff_gbs,
                   "handoff_gbs_how": "per cut: bytes / CUDA-event time of the copy-engine hand-off of the window's last "
                                      "micro-batch (waits excluded); null = the last kernel stores straight into the peer's slot"},
        "clocks": clock_info,
        "e2e": {"value": round(e2e_value, 1), "unit": "img/s", "h2d_bytes_per_step": int(x.nbytes),
                "d2h_bytes_per_step": B * 1000 * 4, "p50_latency_ms": p50_unloaded, "p50_latency_ms_saturated": p50, "p99_latency_ms_saturated": p99,
                "value_incl_fill_and_drain": round((W + K) * B / e2e_wall_all, 1),
                "ingest": ("round-robin over the PCIe links of all %d GPUs, then NVLink into stage 1" % n)
                          if n > 1 and args.ingest_rails == "auto" else "the first stage's own PCIe link",
                "timing": "wall clock from the arrival of the last warm-up result to the arrival of the K-th timed "
                          "result, W + K pinned host batches fed back to back (pipeline full); "
                          "value_incl_fill_and_drain = (W + K) batches over the whole drive"},
        "e2e_uint8": e2e_u8,
        "gpu_launches": int(launches_per_step * K),
        "roofline": roofline,
        "cpu_baseline": cpu,
        "failover": failover,
        "baseline_configs": baseline_configs,
        "replica_comparison": replicas,
        "checks": checks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=128, help="micro-batch (images per step)")
    ap.add_argument("--cuts", choices=["balanced", "config"], default="balanced",
                    help="N>1: time-balanced cut points (default) or the split points BASELINE.json names for N=2/4")
    ap.add_argument("--ref-batch", type=int, default=0, help="micro-batch of the CPU reference runs (0 = the arm's own --batch)")
    ap.add_argument("--cpu-steps", type=int, default=16, help="micro-batches of the cpu_baseline sample (10-30 s of CPU work)")
    ap.add_argument("--sustained-s", type=float, default=1.5, help="length of the extra seconds-long resident window")
    ap.add_argument("--ingest-rails", choices=["auto", "off"], default="auto",
                    help="N>1: spread pinned input batches over every GPU's PCIe link (auto) or use the first stage's only")
    ap.add_argument("--ramp-s", type=float, default=0.15,
                    help="untimed clock-ramp steps (about this many seconds of resident micro-batches) before the W warm-up steps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-failover", action="store_true")
    ap.add_argument("--no-uint8-leg", action="store_true", help="skip the extra end-to-end leg with raw uint8 inputs")
    ap.add_argument("--no-baseline-configs", action="store_true", help="skip the sessions that measure BASELINE configs 1-3")
    ap.add_argument("--no-replica-comparison", action="store_true",
                    help="torchrun, N>1: skip the R replicas x S stages comparison sessions (SURVEY 8e)")
    ap.add_argument("--dump-ops", default=None, help="write the per-launch timing table (JSON) here")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
