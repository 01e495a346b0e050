"""Grouped GEMM (nvte_multi_tensor_gemm): one launch vs member-by-member launches. python scripts/dev_grouped.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transformerengine_b200 import ops


def timeit(fn, iters=20, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def case(name, splits, k, n, kind="mxfp8"):
    q = (lambda t: ops.mxfp8_quantize(t, swizzled_scales=True)) if kind == "mxfp8" else \
        (lambda t: ops.nvfp4_quantize(t, with_rht=False, swizzled_scales=True))
    xs = [q(torch.randn(m, k, device="cuda").bfloat16()) for m in splits]
    ws = [q((torch.randn(n, k, device="cuda") * 0.5).bfloat16()) for _ in splits]
    outs = [torch.empty(m, n, dtype=torch.bfloat16, device="cuda") for m in splits]
    flops = 2.0 * sum(splits) * k * n
    t_g = timeit(lambda: ops.multi_tensor_gemm(ws, xs, "TN", out=outs))
    t_l = timeit(lambda: [ops.general_gemm(w, x, "TN", out=o) for w, x, o in zip(ws, xs, outs)])
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for w, x, o in zip(ws, xs, outs): ops.general_gemm(w, x, "TN", out=o)
    t_lg = timeit(g.replay)
    print(f"{name:44s} grouped {t_g*1e3:8.1f} us {flops/t_g/1e9:7.0f} TF | loop {t_l*1e3:8.1f} us {flops/t_l/1e9:7.0f} TF"
          f" | loop (CUDA graph) {t_lg*1e3:8.1f} us {flops/t_lg/1e9:7.0f} TF", flush=True)


torch.manual_seed(0)
# MoE MLP shapes: Llama-3-8B-like experts (4096 -> 14336), 8192 tokens routed over 8 / 16 / 64 experts
case("8 experts x 1024 tok, 4096->14336", [1024] * 8, 4096, 14336)
case("8 experts ragged, 4096->14336", [2304, 256, 1536, 128, 1792, 640, 1024, 512], 4096, 14336)
case("16 experts x 512 tok, 4096->14336", [512] * 16, 4096, 14336)
case("64 experts x 128 tok, 2048->1408", [128] * 64, 2048, 1408)
case("64 experts ragged, 2048->1408", [((i * 37) % 7) * 64 + 64 for i in range(64)], 2048, 1408)
case("8 experts x 1024 tok nvfp4, 4096->14336", [1024] * 8, 4096, 14336, "nvfp4")
