"""cuBLASLt (through torch._scaled_mm) on the config-4 / config-3 GEMM shapes: the library kernel the
reference would run on this box (gemm/cublaslt_gemm.cu:775), timed next to ours. Comparator only."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from transformerengine_b200 import ops  # noqa: E402


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def ceil_div(a, b):
    return (a + b - 1) // b


def run(verbose=True):
    res = {}
    dev = "cuda"
    for name, (n, m, k) in {"fprop_8192x14336x4096": (8192, 14336, 4096), "dgrad_8192x4096x14336": (8192, 4096, 14336),
                            "wgrad_14336x4096x8192": (14336, 4096, 8192)}.items():
        x = torch.randn(n, k, device=dev).bfloat16()
        w = (torch.randn(m, k, device=dev) * 0.5).bfloat16()
        fl = 2.0 * n * m * k
        r = {}
        # per-tensor FP8
        xa, wa = x.to(torch.float8_e4m3fn), w.to(torch.float8_e4m3fn)
        one = torch.ones((), device=dev)
        try:
            ms = timeit(lambda: torch._scaled_mm(xa, wa.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16))
            r["cublaslt_fp8_tflops"] = round(fl / ms / 1e9, 1)
        except Exception as e:  # noqa: BLE001
            r["cublaslt_fp8_error"] = str(e)[:120]
        # MXFP8 (1x32 E8M0 block scales in the swizzled layout)
        try:
            sa = torch.full((ceil_div(n, 128) * 128 * ceil_div(k // 32, 4) * 4,), 127, dtype=torch.uint8, device=dev).view(torch.float8_e8m0fnu)
            sb = torch.full((ceil_div(m, 128) * 128 * ceil_div(k // 32, 4) * 4,), 127, dtype=torch.uint8, device=dev).view(torch.float8_e8m0fnu)
            ms = timeit(lambda: torch._scaled_mm(xa, wa.t(), scale_a=sa, scale_b=sb, out_dtype=torch.bfloat16))
            r["cublaslt_mxfp8_tflops"] = round(fl / ms / 1e9, 1)
        except Exception as e:  # noqa: BLE001
            r["cublaslt_mxfp8_error"] = str(e)[:160]
        ms = timeit(lambda: torch.matmul(x, w.t()))
        r["cublas_bf16_tflops"] = round(fl / ms / 1e9, 1)
        # ours, same operands
        qa, qb = ops.mxfp8_quantize(w, swizzled_scales=True), ops.mxfp8_quantize(x, swizzled_scales=True)
        o = torch.empty(n, m, dtype=torch.bfloat16, device=dev)
        ms = timeit(lambda: ops.general_gemm(qa, qb, "TN", out=o))
        r["ours_mxfp8_tflops"] = round(fl / ms / 1e9, 1)
        s = torch.ones(1, device=dev)
        fa, fb = ops.fp8_quantize(w, s), ops.fp8_quantize(x, s)
        ms = timeit(lambda: ops.general_gemm(fa, fb, "TN", out=o))
        r["ours_fp8_tflops"] = round(fl / ms / 1e9, 1)
        res[name] = r
        if verbose:
            print(name, r, file=sys.stderr, flush=True)
    # NVFP4 via _scaled_mm (float4_e2m1fn_x2 operands, e4m3 1x16 scales)
    try:
        n, m, k = 16384, 8192, 8192
        a4 = torch.randint(0, 255, (n, k // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
        b4 = torch.randint(0, 255, (m, k // 2), dtype=torch.uint8, device=dev).view(torch.float4_e2m1fn_x2)
        sa = torch.full((ceil_div(n, 128) * 128 * ceil_div(k // 16, 4) * 4,), 0x38, dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
        sb = torch.full((ceil_div(m, 128) * 128 * ceil_div(k // 16, 4) * 4,), 0x38, dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
        ms = timeit(lambda: torch._scaled_mm(a4, b4.t(), scale_a=sa, scale_b=sb, out_dtype=torch.bfloat16))
        res["nvfp4_16384x8192x8192"] = {"cublaslt_nvfp4_tflops": round(2.0 * n * m * k / ms / 1e9, 1)}
    except Exception as e:  # noqa: BLE001
        res["nvfp4_16384x8192x8192"] = {"cublaslt_nvfp4_error": str(e)[:200]}
    if verbose:
        print(res["nvfp4_16384x8192x8192"], file=sys.stderr, flush=True)
    return res


if __name__ == "__main__":
    run()
