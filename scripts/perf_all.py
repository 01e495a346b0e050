"""Per-kernel timing table for the hot path (development aid, not the bench contract).

    python scripts/perf_all.py [quant] [gemm]

Prints one line per kernel: time per launch, GB/s (algorithmic bytes)
or TFLOP/s. Inputs rotate over several buffer sets so that L2 does not hold them.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from transformerengine_b200 import ops  # noqa: E402

PEAKS = {}
_p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(_p):
    PEAKS = json.load(open(_p))
HBM = PEAKS.get("hbm_gbs", 6650.0)


def timeit(fn, iters=20, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def report(name, ms, nbytes=None, flops=None):
    s = f"{name:58s} {ms * 1e3:9.1f} us"
    if nbytes:
        gbs = nbytes / ms / 1e6
        s += f"  {gbs:8.0f} GB/s  {gbs / HBM * 100:5.1f}% of measured HBM"
    if flops:
        s += f"  {flops / ms / 1e9:8.0f} TFLOP/s"
    print(s, flush=True)


def quant():
    for rows, cols in ((4096, 4096), (8192, 4096), (8192, 14336), (16384, 8192)):
        nset = max(2, int(600e6 // (rows * cols * 2)) + 1)
        nset = min(nset, 8)
        xs = [torch.randn(rows, cols, device="cuda").bfloat16() for _ in range(nset)]
        n = rows * cols
        # MXFP8 both directions
        outs = [ops.mxfp8_quantize(x) for x in xs]
        ms = timeit(lambda i: ops.mxfp8_quantize(xs[i % nset], out=outs[i % nset]))
        report(f"mxfp8 row+col {rows}x{cols}", ms, n * 2 + 2 * n + 2 * n // 32)
        outs = [ops.mxfp8_quantize(x, columnwise=False) for x in xs]
        ms = timeit(lambda i: ops.mxfp8_quantize(xs[i % nset], columnwise=False, out=outs[i % nset]))
        report(f"mxfp8 row     {rows}x{cols}", ms, n * 2 + n + n // 32)
        outs = [ops.mxfp8_quantize(x, swizzled_scales=True) for x in xs]
        ms = timeit(lambda i: ops.mxfp8_quantize(xs[i % nset], swizzled_scales=True, out=outs[i % nset]))
        report(f"mxfp8 row+col swizzled {rows}x{cols}", ms, n * 2 + 2 * n + 2 * n // 32)
        del outs
        # amax + FP8 per-tensor
        amax = torch.zeros(1, device="cuda")
        ms = timeit(lambda i: ops.compute_amax(xs[i % nset], amax))
        report(f"amax          {rows}x{cols}", ms, n * 2)
        scale = torch.ones(1, device="cuda")
        outs = [ops.fp8_quantize(x, scale) for x in xs]
        ms = timeit(lambda i: ops.fp8_quantize(xs[i % nset], scale, amax=amax, out=outs[i % nset]))
        report(f"fp8 cast      {rows}x{cols}", ms, n * 3)
        del outs
        # NVFP4
        ar, ac = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
        ms = timeit(lambda i: ops.hadamard_transform_amax(xs[i % nset], ar, ac))
        report(f"nvfp4 rht amax {rows}x{cols}", ms, n * 2)
        outs = [ops.nvfp4_quantize(x, with_rht=True) for x in xs]
        ms = timeit(lambda i: ops.nvfp4_quantize(xs[i % nset], with_rht=True, out=outs[i % nset]))
        report(f"nvfp4 rht amax+cast e2e {rows}x{cols}", ms, n * 2 + n * 2 + n + 2 * n // 16)
        outs = [ops.nvfp4_quantize(x, with_rht=False) for x in xs]
        ms = timeit(lambda i: ops.nvfp4_quantize(xs[i % nset], with_rht=False, out=outs[i % nset]))
        report(f"nvfp4 (no rht) amax+cast e2e {rows}x{cols}", ms, n * 2 + n * 2 + n + 2 * n // 16)
        del outs, xs
        torch.cuda.empty_cache()


def gemm():
    def run(kind, n, m, k, layout):
        ashape = (m, k) if layout[0] == "T" else (k, m)
        bshape = (k, n) if layout[1] == "T" else (n, k)
        a = (torch.randn(*ashape, device="cuda") * 0.5).bfloat16()
        b = torch.randn(*bshape, device="cuda").bfloat16()
        if kind == "mxfp8":
            qa = ops.mxfp8_quantize(a, swizzled_scales=True)
            qb = ops.mxfp8_quantize(b, swizzled_scales=True)
        elif kind == "fp8":
            s = torch.ones(1, device="cuda")
            qa, qb = ops.fp8_quantize(a, s), ops.fp8_quantize(b, s)
        else:
            qa = ops.nvfp4_quantize(a, with_rht=False, swizzled_scales=True)
            qb = ops.nvfp4_quantize(b, with_rht=False, swizzled_scales=True)
        out = torch.empty(n, m, dtype=torch.bfloat16, device="cuda")
        ms = timeit(lambda i: ops.general_gemm(qa, qb, layout, out=out), iters=20)
        report(f"gemm {kind} {layout} n={n} m={m} k={k}", ms, flops=2.0 * n * m * k)

    for kind in ("mxfp8", "fp8"):
        run(kind, 8192, 14336, 4096, "TN")
        run(kind, 8192, 4096, 14336, "NN")
        run(kind, 14336, 4096, 8192, "NT")
    run("fp8", 2048, 3072, 768, "TN")
    run("fp8", 8192, 28672, 8192, "TN")
    run("nvfp4", 16384, 8192, 8192, "TN")
    run("nvfp4", 8192, 14336, 4096, "TN")
    # cuBLAS bf16 on the same shapes, for context
    a = torch.randn(14336, 4096, device="cuda").bfloat16()
    b = torch.randn(8192, 4096, device="cuda").bfloat16()
    ms = timeit(lambda i: torch.matmul(b, a.t()), iters=20)
    report("torch.matmul bf16 8192x14336x4096", ms, flops=2.0 * 8192 * 14336 * 4096)


if __name__ == "__main__":
    which = sys.argv[1:] or ["quant", "gemm"]
    print(torch.cuda.get_device_name(0), "HBM peak used:", HBM)
    if "quant" in which:
        quant()
    if "gemm" in which:
        gemm()
