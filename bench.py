#!/usr/bin/env python
"""Benchmark of the low-precision Linear hot path (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one pass of the hot-path operator over one batch of synthetic input.
Headline workload (N=1): MXFP8 1x32 rowwise + 32x1 columnwise quantization of the
Llama-3-8B MLP activation [8192, 4096] bf16 (BASELINE.json config 4), metric GB/s over
the algorithmic bytes of SURVEY.md section 8(d). Other workloads are reported under
"extra" on the same JSON line. Under torchrun each rank quantizes its own row shard
(weak scaling, no data-path collective); value = bytes of all ranks / max-over-ranks time.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS, COLS = 8192, 4096  # config 4 activation x: [tokens, in_features]


def mxfp8_bytes(rows, cols, elt=2):
    """Algorithmic bytes, rowwise + columnwise: read + 2 fp8 copies + 2 scale sets."""
    return rows * cols * elt + 2 * rows * cols + 2 * rows * cols // 32


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def run_reference(args):
    """The reference's own CPU implementation of the path (oracle/_ref `compute_ref`,
    tests/cpp/operator/test_cast_mxfp8.cu:39-166) on the box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import oracle as O
    from oracle import ref_host as R
    from tests.util import bits16, make_input
    rows = 2048  # bounded sample: a 2048-row slab of the [8192, 4096] activation per step
    x = bits16(make_input(rows, COLS, "randn", 0))
    if R.available():
        fn, kind = (lambda: R.mxfp8_compute_ref(x)), "reference"
    else:
        fn, kind = (lambda: O.mxfp8_quantize(x)), "port"
    for _ in range(args.warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    dt = (time.perf_counter() - t0) / args.steps
    gbs = mxfp8_bytes(rows, COLS) / dt / 1e9
    cores = O.num_threads()
    print(json.dumps({
        "impl": "reference", "metric": "mxfp8_quantize_GBps", "value": gbs, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16->e4m3",
        "data": "synthetic",
        "config": {"workload": f"MXFP8 1x32 rowwise+columnwise quantize, bf16 [{ROWS},{COLS}] "
                               f"(config 4 activation); CPU step = {rows}-row slab"},
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": kind,
                         "sample": f"{rows}x{COLS} bf16 slab, OpenMP over 32x32 tiles, "
                                   f"{torch.get_num_threads()} torch / {cores} omp threads"},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def cpu_baseline():
    from oracle import oracle as O
    from oracle import ref_host as R
    from tests.util import bits16, make_input
    rows = 4096  # BASELINE.json config 1: 4096x4096 bf16 on CPU
    x = bits16(make_input(rows, 4096, "randn", 0))
    fn, kind = ((lambda: R.mxfp8_compute_ref(x)), "reference") if R.available() else \
               ((lambda: O.mxfp8_quantize(x)), "port")
    fn()
    best = 1e30
    t_end = time.perf_counter() + 12.0
    n = 0
    while n < 3 or (time.perf_counter() < t_end and n < 10):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
        n += 1
    return {"value": mxfp8_bytes(rows, 4096) / best / 1e9, "unit": "GB/s", "cores": O.num_threads(),
            "kind": kind, "sample": f"MXFP8 rowwise+columnwise of 4096x4096 bf16 (config 1), best of {n}"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from transformerengine_b200 import cabi as C
    from transformerengine_b200 import ops

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    # Rotating buffer sets: > 2x L2 (126 MB) of distinct data between two uses of a set.
    nsets = 6
    g = torch.Generator(device="cpu").manual_seed(rank)
    host_x = torch.randn(ROWS, COLS, generator=g).to(torch.bfloat16).pin_memory()
    xs = [host_x.to(dev, non_blocking=True) * (1.0 + 0.01 * i) for i in range(nsets)]
    outs = [ops.mxfp8_quantize(xs[i]) for i in range(nsets)]
    ins = [C.plain(x) for x in xs]
    lib, stream = C.lib(), torch.cuda.current_stream()
    sp = stream.cuda_stream

    def step(i):
        C.check(lib.nvte_quantize_v2_status(ins[i % nsets].ptr, outs[i % nsets].handle().ptr, None, sp))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(args.warmup, 3)):
        step(i)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for i in range(args.steps):
        step(i)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None

    # end to end: host buffers in, host buffers out, through the public operator
    h_out = [torch.empty(ROWS, COLS, dtype=torch.uint8).pin_memory() for _ in range(2)]
    h_sc = [torch.empty_like(outs[0].rowwise_scale_inv, device="cpu").pin_memory(),
            torch.empty_like(outs[0].columnwise_scale_inv, device="cpu").pin_memory()]
    dx = torch.empty_like(xs[0])
    e2e_steps = max(3, min(args.steps, 10))

    def e2e_step(i):
        dx.copy_(host_x, non_blocking=True)
        q = ops.mxfp8_quantize(dx, out=outs[i % nsets])
        h_out[0].copy_(q.rowwise_data, non_blocking=True)
        h_out[1].copy_(q.columnwise_data, non_blocking=True)
        h_sc[0].copy_(q.rowwise_scale_inv, non_blocking=True)
        h_sc[1].copy_(q.columnwise_scale_inv, non_blocking=True)

    e2e_step(0)
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record(stream)
    for i in range(e2e_steps):
        e2e_step(i)
    e3.record(stream)
    barrier()
    e2e_ms = e2.elapsed_time(e3)

    t = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = t.tolist()
    if rank == 0:
        nbytes = mxfp8_bytes(ROWS, COLS)
        per_step_ms = ms / args.steps
        value = world * nbytes / (per_step_ms * 1e-3) / 1e9
        achieved = nbytes / (per_step_ms * 1e-3) / 1e9
        peak, peak_src = peaks()
        h2d = host_x.numel() * 2
        d2h = 2 * ROWS * COLS + h_sc[0].numel() + h_sc[1].numel()
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("mxfp8_quantize_8192x4096")
        line = {
            "metric": "mxfp8_quantize_GBps", "value": value, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": per_step_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16->e4m3",
            "data": "synthetic",
            "config": {"workload": f"MXFP8 1x32 rowwise+columnwise quantize, bf16 [{ROWS},{COLS}] per GPU "
                                   "(BASELINE config 4 activation)",
                       "l2": f"{nsets} rotating input/output sets, {nsets * 132} MB > 2x L2",
                       "algorithmic_bytes_per_step": nbytes, "parallelism": f"row-sharded x{world}"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "kernel": "teb::mxfp8::quantize_kernel<bf16,row,col,e4m3>"},
            "e2e": {"value": world * nbytes / (e2e_ms / e2e_steps * 1e-3) / 1e9, "unit": "GB/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps},
            "gpu_launches": args.steps,
            "clocks": clocks,
            "cpu_baseline": cpu_baseline() if world == 1 else None,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        args.steps = min(args.steps, 20)
        args.warmup = min(args.warmup, 2)
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
