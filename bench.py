#!/usr/bin/env python
"""Benchmark of the low-precision Linear hot path (contract: DESIGN.md section 5).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-extra]

A "step" is one pass of the hot-path quantizer over one batch of synthetic input: the MXFP8
quantization work of BASELINE.json config 4 (Llama-3-8B MLP Linear, 8192 tokens) -- x [8192,4096],
W [14336,4096] and dy [8192,14336], bf16 -> E4M3 rowwise + columnwise with E8M0 scales written in
the tcgen05 layout; 3 kernel launches, 851,968,000 algorithmic bytes.
Metric: GB/s over those bytes. The block-scaled GEMM TFLOP/s of configs 2-4 and the NVFP4+RHT
quantizer of config 3 are reported under "extra" on the same JSON line (N=1). Under torchrun each
rank processes its own copy of the workload (row-sharded quantization has no data-path collective):
weak scaling, value = bytes of all ranks / max-over-ranks time.
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

# BASELINE.json config 4: tokens 8192, in 4096, out 14336
TENSORS = (("x", 8192, 4096), ("W", 14336, 4096), ("dy", 8192, 14336))
WORKLOAD = ("MXFP8 1x32 rowwise+columnwise quantize (GEMM-swizzled E8M0 scales) of the config-4 "
            "Linear's tensors: x[8192,4096], W[14336,4096], dy[8192,14336] bf16, per GPU")
METRIC = "mxfp8_quantize_GBps"


def mxfp8_bytes(rows, cols, elt=2):
    """Algorithmic bytes, rowwise + columnwise: read + 2 fp8 copies + 2 scale sets."""
    return rows * cols * elt + 2 * rows * cols + 2 * rows * cols // 32


STEP_BYTES = sum(mxfp8_bytes(r, c) for _, r, c in TENSORS)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, \
        "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (NVML, every 2 ms from a thread;
    `nvidia-smi -lms` as the fallback when the NVML binding is missing)."""

    def __init__(self, index):
        self.index, self.sm, self.mx, self.reasons, self._stop, self._t = index, [], None, set(), False, None
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.reasons.update(n for n, b in bits.items() if r & b)
            except Exception:
                break
            time.sleep(0.002)

    def start(self):
        if self.nv is not None:
            try:
                self.mx = float(self.nv.nvmlDeviceGetMaxClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            except Exception:
                self.mx = None
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
            return
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms",
                                          "20", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.rows = []
            threading.Thread(target=lambda: [self.rows.append([c.strip() for c in ln.split(",")])
                                             for ln in self.proc.stdout], daemon=True).start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.nv is not None:
            self._stop = True
            if self._t is not None:
                self._t.join(timeout=1.0)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}
        if getattr(self, "proc", None) is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.1)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvidia-smi"}


# --------------------------------------------------------------------------------------------- #
# CPU arm: the reference's own host quantizer (oracle/_ref) or, if it was not built, the C port
# --------------------------------------------------------------------------------------------- #
SLAB_ROWS = 1024  # bounded sample: the first 1024 rows of each of the three tensors per CPU step
SLAB_BYTES = sum(mxfp8_bytes(SLAB_ROWS, c) for _, _, c in TENSORS)
SLAB_DESC = (f"first {SLAB_ROWS} rows of x, W and dy ({SLAB_BYTES} algorithmic bytes per step); "
             "reference compute_ref (tests/cpp/operator/test_cast_mxfp8.cu:39-166), OpenMP over 32x32 tiles")


def _cpu_fn():
    from oracle import oracle as O
    from oracle import ref_host as R
    from tests.util import bits16, make_input
    xs = [bits16(make_input(SLAB_ROWS, c, "randn", i)) for i, (_, _, c) in enumerate(TENSORS)]
    if R.available():
        return (lambda: [R.mxfp8_compute_ref(x) for x in xs]), "reference", O.num_threads()
    return (lambda: [O.mxfp8_quantize(x) for x in xs]), "port", O.num_threads()


def cpu_baseline(budget_s=12.0):
    fn, kind, cores = _cpu_fn()
    fn()
    best, n, t_end = 1e30, 0, time.perf_counter() + budget_s
    while n < 3 or (time.perf_counter() < t_end and n < 40):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
        n += 1
    return {"value": SLAB_BYTES / best / 1e9, "unit": "GB/s", "cores": cores, "kind": kind,
            "sample": SLAB_DESC + f", best of {n}"}


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    fn, kind, cores = _cpu_fn()
    steps, warm = min(args.steps, 50), min(args.warmup, 3)
    for _ in range(max(warm, 1)):
        fn()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    dt = (time.perf_counter() - t0) / steps
    gbs = SLAB_BYTES / dt / 1e9
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16->e4m3", "data": "synthetic",
        "config": {"workload": WORKLOAD, "cpu_step": SLAB_DESC},
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": kind, "sample": SLAB_DESC},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# --------------------------------------------------------------------------------------------- #
# GPU arm
# --------------------------------------------------------------------------------------------- #
def _time_loop(torch, fn, iters, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters  # ms


def extras(torch, ops, pk):
    """Other rows of the metric (GEMM TFLOP/s, NVFP4 quantize GB/s, config-2 step latency)."""
    import transformerengine_b200 as te
    from transformerengine_b200 import recipe
    out = {}
    hbm = pk["hbm_gbs"]
    bf16 = pk.get("bf16_tflops", 1590.0)
    dev = "cuda"

    def gemm(kind, n, m, k, layout, iters=40):
        ashape = (m, k) if layout[0] == "T" else (k, m)
        bshape = (k, n) if layout[1] == "T" else (n, k)
        a = (torch.randn(*ashape, device=dev) * 0.5).bfloat16()
        b = torch.randn(*bshape, device=dev).bfloat16()
        if kind == "mxfp8":
            qa, qb = ops.mxfp8_quantize(a, swizzled_scales=True), ops.mxfp8_quantize(b, swizzled_scales=True)
        elif kind == "fp8":
            s = torch.ones(1, device=dev)
            qa, qb = ops.fp8_quantize(a, s), ops.fp8_quantize(b, s)
        else:
            qa = ops.nvfp4_quantize(a, with_rht=False, swizzled_scales=True)
            qb = ops.nvfp4_quantize(b, with_rht=False, swizzled_scales=True)
        o = torch.empty(n, m, dtype=torch.bfloat16, device=dev)
        smp = ClockSampler(torch.cuda.current_device())
        smp.start()
        ms = _time_loop(torch, lambda i: ops.general_gemm(qa, qb, layout, out=o), iters)
        clk = smp.stop()
        tf = 2.0 * n * m * k / ms / 1e9
        nominal = 9000.0 if kind == "nvfp4" else 4500.0
        mult = 4.0 if kind == "nvfp4" else 2.0
        return {"tflops": round(tf, 1), "us": round(ms * 1e3, 1), "frac_of_nominal_dense_peak": round(tf / nominal, 3),
                f"frac_of_{int(mult)}x_measured_bf16": round(tf / (mult * bf16), 3),
                "sm_mhz_under_load": clk.get("sm_mhz"), "throttle_reasons": clk.get("reasons")}

    # config 3: NVFP4 quantize with RHT of 16384x8192 bf16 (two buffer sets > L2); the two launches
    # of the quantizer (amax pass, cast pass) timed through the C ABI with prebuilt handles
    from transformerengine_b200 import cabi as C
    R, Cc = 16384, 8192
    xs = [torch.randn(R, Cc, device=dev).bfloat16() for _ in range(2)]
    n = R * Cc
    outs = [ops.nvfp4_quantize(x, with_rht=True, swizzled_scales=True) for x in xs]
    ins = [C.plain(x) for x in xs]
    mat, cfg, lib, sp = C.plain(ops.rht_matrix(xs[0].device)), C.QuantConfig(), C.lib(), C.stream_ptr()
    amax_h = []
    for o in outs:
        h = C.TensorHandle(C.NVFP4_1D_SCALING)
        h.set(C.kAmax, o.amax_rowwise).set(C.kColumnwiseAmax, o.amax_columnwise)
        amax_h.append(h)

    def f_amax(i):
        C.check(lib.nvte_hadamard_transform_amax_status(ins[i % 2].ptr, amax_h[i % 2].ptr, 0, ops.RHT_SIGN_MASK, sp))

    def f_cast(i):
        C.check(lib.nvte_quantize_with_hadamard_transform_status(ins[i % 2].ptr, outs[i % 2].handle().ptr,
                                                                 mat.ptr, cfg.ptr, sp))

    def f_both(i):
        f_amax(i)
        f_cast(i)

    ms_amax, ms_cast, ms_e2e = _time_loop(torch, f_amax, 30), _time_loop(torch, f_cast, 30), _time_loop(torch, f_both, 30)
    b_amax, b_cast = n * 2, n * 2 + n + 2 * n // 16
    out["config3_nvfp4_rht_quantize_16384x8192"] = {
        "amax_pass": {"us": round(ms_amax * 1e3, 1), "GBps": round(b_amax / ms_amax / 1e6), "frac_hbm": round(b_amax / ms_amax / 1e6 / hbm, 3)},
        "cast_pass": {"us": round(ms_cast * 1e3, 1), "GBps": round(b_cast / ms_cast / 1e6), "frac_hbm": round(b_cast / ms_cast / 1e6 / hbm, 3)},
        "end_to_end": {"us": round(ms_e2e * 1e3, 1), "GBps": round((b_amax + b_cast) / ms_e2e / 1e6),
                       "frac_hbm": round((b_amax + b_cast) / ms_e2e / 1e6 / hbm, 3)},
        "kernel": "teb::nvfp4::nvfp4_tc_kernel (amax mode, cast mode)"}
    del xs, outs, ins, amax_h

    # RMSNorm / LayerNorm -> MXFP8 (rowwise + columnwise, swizzled scales) in ONE
    # kernel, against norm -> bf16 followed by the MXFP8 quantize kernel (what an unfused
    # LayerNormLinear does); 16384 x 8192 bf16 activations, two buffer sets > L2, C-ABI handles
    R, Cc = 16384, 8192
    n = R * Cc
    xs = [torch.randn(R, Cc, device=dev).bfloat16() for _ in range(2)]
    gamma, beta = torch.rand(Cc, device=dev).bfloat16(), torch.rand(Cc, device=dev).bfloat16()
    rs_shape, cs_shape = ops.mxfp8_scale_shapes(R, Cc)
    mx = [ops.MXFP8TensorStorage(
        rowwise_data=torch.empty(R, Cc, dtype=torch.uint8, device=dev),
        rowwise_scale_inv=torch.zeros(rs_shape, dtype=torch.uint8, device=dev),
        columnwise_data=torch.empty(R, Cc, dtype=torch.uint8, device=dev),
        columnwise_scale_inv=torch.zeros(cs_shape, dtype=torch.uint8, device=dev),
        fp8_dtype=torch.float8_e4m3fn, shape=(R, Cc), with_gemm_swizzled_scales=True) for _ in xs]
    ys = [torch.empty_like(x) for x in xs]
    hx, hy, hmx = [C.plain(x) for x in xs], [C.plain(y) for y in ys], [m.handle() for m in mx]
    hg, hb = C.plain(gamma), C.plain(beta)
    rsig, mu = torch.empty(R, device=dev), torch.empty(R, device=dev)
    hrs, hmu, hws = C.plain(rsig), C.plain(mu), C.plain(torch.empty(16, dtype=torch.uint8, device=dev))
    qcfg = C.QuantConfig()

    def norm_call(ln, z):
        def f(i):
            if ln:
                C.check(lib.nvte_layernorm_fwd_status(hx[i % 2].ptr, hg.ptr, hb.ptr, 1e-5, z[i % 2].ptr, hmu.ptr,
                                                      hrs.ptr, hws.ptr, 0, 0, sp))
            else:
                C.check(lib.nvte_rmsnorm_fwd_status(hx[i % 2].ptr, hg.ptr, 1e-5, z[i % 2].ptr, hrs.ptr, hws.ptr, 0, 0, sp))
        return f

    def unfused(ln):
        f = norm_call(ln, hy)

        def g(i):
            f(i)
            C.check(lib.nvte_quantize_v2_status(hy[i % 2].ptr, hmx[i % 2].ptr, qcfg.ptr, sp))
        return g

    norm = {}
    for name, ln in (("rmsnorm", False), ("layernorm", True)):
        ms_f, ms_h, ms_u = (_time_loop(torch, norm_call(ln, hmx), 30), _time_loop(torch, norm_call(ln, hy), 30),
                            _time_loop(torch, unfused(ln), 30))
        b_f, b_h = 4 * n + n // 16, 4 * n
        norm[name] = {
            "fused_to_mxfp8_row_col": {"us": round(ms_f * 1e3, 1), "GBps": round(b_f / ms_f / 1e6),
                                       "frac_hbm": round(b_f / ms_f / 1e6 / hbm, 3)},
            "to_bf16": {"us": round(ms_h * 1e3, 1), "GBps": round(b_h / ms_h / 1e6), "frac_hbm": round(b_h / ms_h / 1e6 / hbm, 3)},
            "unfused_norm_then_quantize_us": round(ms_u * 1e3, 1)}
    norm["kernel"] = "teb::norm::norm_fwd_kernel (cluster of CTAs per 32-row slab, st.async row statistics)"
    norm["algorithmic_bytes"] = "2 B read + 2 x 1 B written + scales per element (fused); 2 + 2 (bf16)"
    out["norm_fused_quantize_16384x8192"] = norm
    del xs, ys, mx, hx, hy, hmx

    # config 2: te.Linear 768->3072, 2048 tokens, DelayedScaling E4M3, fwd+bwd (README example)
    torch.manual_seed(0)
    model = te.Linear(768, 3072, bias=True)
    inp = torch.randn(2048, 768, device=dev)
    rec = recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.E4M3)

    def lin_step(_):
        with te.autocast(enabled=True, recipe=rec):
            o = model(inp)
        o.sum().backward()

    ms = _time_loop(torch, lin_step, 50, warm=5)
    # the same step captured once with te.make_graphed_callables (reference pytorch/graph.py)
    gmodel = te.Linear(768, 3072, bias=True)
    ginp = torch.randn(2048, 768, device=dev, requires_grad=True)
    gfn = te.make_graphed_callables(gmodel, (ginp,), num_warmup_iters=3, enabled=True, recipe=rec)

    def graphed_step(_):
        gmodel.zero_grad(set_to_none=True)
        gfn(ginp).sum().backward()

    ms_g = _time_loop(torch, graphed_step, 50, warm=5)
    out["config2_linear_768x3072_T2048_delayed_fwd_bwd"] = {
        "us_per_step": round(ms * 1e3, 1), "gemm_tflops": round(3 * 2.0 * 2048 * 768 * 3072 / ms / 1e9, 1),
        "us_per_step_cuda_graph": round(ms_g * 1e3, 1),
        "gemm_tflops_cuda_graph": round(3 * 2.0 * 2048 * 768 * 3072 / ms_g / 1e9, 1),
        "note": "3 casts + 3 GEMMs + bgrad + fused amax/scale update; the eager step is host-launch bound at "
                "this size, te.make_graphed_callables replays forward and backward as two CUDA graphs"}
    # GEMMs last: they drive the part into its power cap and the clocks stay low for a while
    out["config4_mxfp8_gemm"] = {"fprop_TN_8192x14336x4096": gemm("mxfp8", 8192, 14336, 4096, "TN"),
                                 "dgrad_NN_8192x4096x14336": gemm("mxfp8", 8192, 4096, 14336, "NN"),
                                 "wgrad_NT_14336x4096x8192": gemm("mxfp8", 14336, 4096, 8192, "NT")}
    out["config3_nvfp4_gemm_TN_16384x8192x8192"] = gemm("nvfp4", 16384, 8192, 8192, "TN")
    out["config5_fp8_gemm_TN_8192x28672x8192_unsharded"] = gemm("fp8", 8192, 28672, 8192, "TN", iters=10)

    # sec. 8(f) rank 4: a GroupedLinear's expert GEMMs (nvte_multi_tensor_gemm) as one persistent launch over the
    # concatenated tile list vs member-by-member launches replayed from a CUDA graph (no host time in either)
    def grouped(splits, k, n):
        q = lambda t: ops.mxfp8_quantize(t, swizzled_scales=True)
        gx = [q(torch.randn(m_, k, device=dev).bfloat16()) for m_ in splits]
        gw = [q((torch.randn(n, k, device=dev) * 0.5).bfloat16()) for _ in splits]
        go = [torch.empty(m_, n, dtype=torch.bfloat16, device=dev) for m_ in splits]
        fl = 2.0 * sum(splits) * k * n
        ms_one = _time_loop(torch, lambda i: ops.multi_tensor_gemm(gw, gx, "TN", out=go), 20)
        for w_, x_, o_ in zip(gw, gx, go):
            ops.general_gemm(w_, x_, "TN", out=o_)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for w_, x_, o_ in zip(gw, gx, go):
                ops.general_gemm(w_, x_, "TN", out=o_)
        ms_loop = _time_loop(torch, lambda i: gr.replay(), 20)
        return {"one_launch_us": round(ms_one * 1e3, 1), "one_launch_tflops": round(fl / ms_one / 1e9, 1),
                "member_by_member_cuda_graph_us": round(ms_loop * 1e3, 1),
                "member_by_member_tflops": round(fl / ms_loop / 1e9, 1)}
    try:
        out["grouped_mxfp8_gemm_moe"] = {
            "8_experts_ragged_8192tok_4096x14336": grouped([2304, 256, 1536, 128, 1792, 640, 1024, 512], 4096, 14336),
            "64_experts_x128tok_2048x1408": grouped([128] * 64, 2048, 1408)}
    except Exception as e:  # noqa: BLE001
        out["grouped_mxfp8_gemm_moe"] = {"error": str(e)[:200]}

    # the library kernel the reference would run here (cuBLASLt through torch._scaled_mm), same
    # shapes, same timing loop: a comparator, not on our path
    try:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import cublaslt_ref
        out["cublaslt_same_box"] = cublaslt_ref.run(verbose=False)
    except Exception as e:  # noqa: BLE001
        out["cublaslt_same_box"] = {"error": str(e)[:200]}

    return out


def config5_fprop_step(torch, dist, world, rank, iters=20):
    """BASELINE config 5: column-parallel Llama-3-70B FC1 (8192 -> 28672), 8192 tokens, FP8 current
    scaling, sequence parallel. One fprop step per rank = amax(x shard) -> all-reduce(MAX) over the
    TP group -> scale -> cast -> all-gather of the FP8 shard (NCCL over NVLink) -> GEMM against the
    rank's cached FP8 weight shard. Strong scaling: the problem is fixed, tp = world."""
    import transformerengine_b200 as te
    from transformerengine_b200 import recipe
    T, fin, fout = 8192, 8192, 28672
    group = dist.new_group(list(range(world))) if world > 1 else None
    torch.manual_seed(7)
    lin = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16,
                    parallel_mode="column" if world > 1 else None, sequence_parallel=world > 1, tp_group=group)
    x = torch.randn(T // world, fin, device="cuda", dtype=torch.bfloat16)
    rec = recipe.Float8CurrentScaling()

    def step(first):
        with torch.no_grad(), te.autocast(enabled=True, recipe=rec, amax_reduction_group=group):
            return lin(x, is_first_microbatch=first)

    def timed():
        step(True)
        for _ in range(3):
            step(False)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            step(False)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    ms = timed()

    def graphed():
        """The same step replayed from one CUDA graph (NCCL collectives captured): what the GPUs
        need when the host's per-launch cost is out of the picture."""
        step(True)
        for _ in range(3):
            step(False)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            step(False)
        torch.cuda.synchronize()
        for _ in range(3):
            g.replay()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    ms_graph = ms_graph_overlap = None
    try:
        ms_graph = graphed()
        if world > 1:   # all-gather in 2 pieces on a side stream, GEMM piece by piece behind them
            lin.overlap_ag_chunks = 2
            ms_graph_overlap = graphed()
            lin.overlap_ag_chunks = 1
    except Exception as e:  # noqa: BLE001
        ms_graph = ms_graph if ms_graph is not None else f"capture failed: {str(e)[:120]}"
    # symmetric_ag: the cast stores its bytes into every rank's gather buffer over NVLink (symmetric memory);
    # the all-gather collapses into that kernel plus one barrier
    symm = {}
    if world > 1:
        try:
            lin.symmetric_ag = True
            for chunks in (1, 2):   # 1: cast+push, barrier, one GEMM; 2: GEMM over piece 0 while piece 1 is on the wire
                lin.symmetric_ag_chunks, lin._symm = chunks, None
                symm[f"chunks{chunks}"] = {"ms_eager": round(timed(), 4), "ms_cuda_graph": round(graphed(), 4)}
        except Exception as e:  # noqa: BLE001
            symm["error"] = str(e)[:200]
        lin.symmetric_ag = False
    return {"tp": world, "ms_per_fprop_step": round(ms, 4), "symmetric_ag_cast_is_the_allgather": symm or None,
            "aggregate_tflops": round(2.0 * T * fin * fout / ms / 1e9, 1),
            "ms_per_fprop_step_cuda_graph": round(ms_graph, 4) if isinstance(ms_graph, float) else ms_graph,
            "ms_cuda_graph_with_allgather_gemm_overlap": round(ms_graph_overlap, 4) if ms_graph_overlap else None,
            "per_rank": {"x_shard": [T // world, fin], "w_shard": [fout // world, fin],
                         "allgather_bytes_received": (world - 1) * (T // world) * fin},
            "note": "te.Linear(parallel_mode='column', sequence_parallel=True) fprop under "
                    "Float8CurrentScaling, weight shard cast cached (is_first_microbatch=False). The eager "
                    "step is host-launch bound at high tp; the CUDA-graph replay is the device time. "
                    "ub_overlap_ag (opt-in): FP8 shard all-gathered in pieces on a side stream, GEMM piece "
                    "by piece behind them (rows land in place through the GEMM's output row blocks)"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from transformerengine_b200 import cabi as C
    from transformerengine_b200 import ops

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    warmup = max(args.warmup, 3)

    # Two rotating buffer sets; one set's inputs + outputs are ~830 MB, so every tensor is
    # evicted from the 126 MB L2 between two uses.
    nsets = 2
    g = torch.Generator(device="cpu").manual_seed(rank)
    host = [torch.randn(r, c, generator=g).to(torch.bfloat16).pin_memory() for _, r, c in TENSORS]
    xs = [[h.to(dev, non_blocking=True) * (1.0 + 0.01 * s) for h in host] for s in range(nsets)]
    outs = [[ops.mxfp8_quantize(x, swizzled_scales=True) for x in xs[s]] for s in range(nsets)]
    ins = [[C.plain(x) for x in xs[s]] for s in range(nsets)]
    lib, stream = C.lib(), torch.cuda.current_stream()
    sp = stream.cuda_stream
    launches = 0

    import ctypes
    nt = len(TENSORS)
    arr_in = [(ctypes.c_void_p * nt)(*[h.ptr for h in ins[s]]) for s in range(nsets)]
    arr_out = [(ctypes.c_void_p * nt)(*[o.handle().ptr for o in outs[s]]) for s in range(nsets)]
    one_launch = os.environ.get("BENCH_MXFP8_SEPARATE_LAUNCHES", "0") != "1"
    launches_per_step = 1 if one_launch else nt

    def step(i):
        # the three tensors of the Linear step in ONE persistent launch (nvte_multi_tensor_quantize walks the
        # concatenated tile list of x, W and dy); BENCH_MXFP8_SEPARATE_LAUNCHES=1: one nvte_quantize_v2 each
        nonlocal launches
        s = i % nsets
        if one_launch:
            C.check(lib.nvte_multi_tensor_quantize_status(arr_in[s], arr_out[s], None, nt, sp))
        else:
            for j in range(nt):
                C.check(lib.nvte_quantize_v2_status(ins[s][j].ptr, outs[s][j].handle().ptr, None, sp))
        launches += launches_per_step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(warmup):
        step(i)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches = 0
    e0.record(stream)
    for i in range(args.steps):
        step(i)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    timed_launches = launches
    clocks = sampler.stop() if rank == 0 else None

    # end to end through the public host-tensor operator (ops.MXFP8HostQuantizer): pinned host
    # inputs -> device -> nvte_quantize_v2 -> pinned host outputs, every byte inside the timed
    # region; row slabs pipeline H2D / quantize / D2H on three streams (PCIe is full duplex).
    h_out = [[torch.empty(r, c, dtype=torch.uint8).pin_memory() for _ in range(2)] for _, r, c in TENSORS]
    h_sc = [[torch.empty_like(outs[0][j].rowwise_scale_inv, device="cpu").pin_memory(),
             torch.empty_like(outs[0][j].columnwise_scale_inv, device="cpu").pin_memory()]
            for j in range(len(TENSORS))]
    pipes = [ops.MXFP8HostQuantizer(r, c, swizzled_scales=True) for _, r, c in TENSORS]
    e2e_steps = max(3, min(args.steps, 20))
    e2e_launches = sum(len(p.slabs) for p in pipes)

    def e2e_step(i):
        for j, p in enumerate(pipes):
            p(host[j], h_out[j][0], h_out[j][1], h_sc[j][0], h_sc[j][1])

    def e2e_join():
        for p in pipes:
            p.join()

    e2e_step(0)
    e2e_join()
    barrier()
    # the pipelined bytes must be the bytes of the one-shot device path
    ref = [ops.mxfp8_quantize(host[j].to(dev), swizzled_scales=True) for j in range(len(TENSORS))]
    for j in range(len(TENSORS)):
        assert torch.equal(h_out[j][0], ref[j].rowwise_data.cpu()), "e2e rowwise bytes differ"
        assert torch.equal(h_out[j][1], ref[j].columnwise_data.cpu()), "e2e columnwise bytes differ"
        assert torch.equal(h_sc[j][0], ref[j].rowwise_scale_inv.cpu()), "e2e rowwise scales differ"
        assert torch.equal(h_sc[j][1], ref[j].columnwise_scale_inv.cpu()), "e2e columnwise scales differ"
    del ref
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record(stream)
    for i in range(e2e_steps):
        e2e_step(i)
    e2e_join()
    e3.record(stream)
    barrier()
    e2e_ms = e2.elapsed_time(e3)

    t = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = t.tolist()
    c5 = None
    if not args.no_extra:
        del xs, outs, ins, pipes
        xs = outs = ins = pipes = None
        torch.cuda.empty_cache()
        try:
            c5 = config5_fprop_step(torch, dist, world, rank)
        except Exception as e:  # noqa: BLE001
            c5 = {"error": str(e)[:300]}
    if rank == 0:
        per_step_ms = ms / args.steps
        achieved = STEP_BYTES / (per_step_ms * 1e-3) / 1e9
        pk, peak_src = peaks()
        h2d = sum(h.numel() * 2 for h in host)
        d2h = sum(2 * r * c for _, r, c in TENSORS) + sum(s.numel() for pair in h_sc for s in pair)
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("mxfp8_quantize_config4_avg_per_launch")
        line = {
            "metric": METRIC, "value": world * achieved, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": warmup, "ms_per_step": per_step_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16->e4m3",
            "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "l2": f"{nsets} rotating buffer sets of ~830 MB each (>> 126 MB L2)",
                       "algorithmic_bytes_per_step": STEP_BYTES, "launches_per_step": launches_per_step,
                       "parallelism": f"independent row shards x{world} (no data-path collective)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved / pk["hbm_gbs"], "traffic": traffic, "peak_source": peak_src,
                         "bytes_per_launch_avg": STEP_BYTES / launches_per_step,
                         "us_per_launch_avg": per_step_ms * 1e3 / launches_per_step,
                         "kernel": "teb::mxfp8::quantize_tma_kernel<bf16, row, col, e4m3, swizzled>",
                         "traffic_note": "ncu dram bytes per launch, average of the 3 launches; below the "
                                         "algorithmic figure because output lines are still dirty in L2 at kernel end"},
            "e2e": {"value": world * STEP_BYTES / (e2e_ms / e2e_steps * 1e-3) / 1e9, "unit": "GB/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                    "launches_per_step": e2e_launches,
                    "how": "ops.MXFP8HostQuantizer: 16 MiB row slabs, H2D / nvte_quantize_v2 / D2H on "
                           "3 streams; outputs checked equal to the one-shot device path"},
            "gpu_launches": timed_launches,
            "clocks": clocks,
        }
        if world == 1:
            line["cpu_baseline"] = cpu_baseline()
            if not args.no_extra:
                line["extra"] = extras(torch, ops, pk)
        if c5 is not None:
            line.setdefault("extra", {})["config5_column_parallel_fprop"] = c5
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extra", action="store_true", help="skip the GEMM / NVFP4 / config-2 extras")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
