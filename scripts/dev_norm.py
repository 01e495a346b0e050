"""LayerNorm / RMSNorm timing (development aid): fused norm -> {bf16, FP8, MXFP8} and backward."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from transformerengine_b200 import cabi as C  # noqa: E402
from transformerengine_b200 import ops  # noqa: E402

HBM = 6457.1


def timeit(fn, iters=30, warm=3):
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


def mx_out(N, H, col, swz=True):
    rs, cs = ops.mxfp8_scale_shapes(N, H)
    dev = "cuda"
    return ops.MXFP8TensorStorage(
        rowwise_data=torch.empty(N, H, dtype=torch.uint8, device=dev),
        rowwise_scale_inv=torch.zeros(rs, dtype=torch.uint8, device=dev),
        columnwise_data=torch.empty(N, H, dtype=torch.uint8, device=dev) if col else None,
        columnwise_scale_inv=torch.zeros(cs, dtype=torch.uint8, device=dev) if col else None,
        fp8_dtype=torch.float8_e4m3fn, shape=(N, H), with_gemm_swizzled_scales=swz)


shapes = [(8192, 4096), (16384, 8192), (16384, 2048), (4096, 16384)] if len(sys.argv) < 2 else \
    [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
for N, H in shapes:
    n = N * H
    nset = min(8, max(2, int(600e6 // (n * 2)) + 1))
    xs = [torch.randn(N, H, device="cuda").bfloat16() for _ in range(nset)]
    gamma = torch.rand(H, device="cuda").bfloat16()
    beta = torch.rand(H, device="cuda").bfloat16()
    for ln in (False, True):
        name = "layernorm" if ln else "rmsnorm"
        b = beta if ln else None
        # bf16 out
        ys = [torch.empty_like(x) for x in xs]
        hs = [C.plain(y) for y in ys]
        ms = timeit(lambda i: ops.norm_fwd(ln, xs[i % nset], gamma, b, 1e-5, hs[i % nset]))
        nb = 4 * n
        print(f"{name} {N}x{H} -> bf16      : {ms * 1e3:7.1f} us {nb / ms / 1e6:6.0f} GB/s {nb / ms / 1e6 / HBM * 100:5.1f}%", flush=True)
        # fp8 delayed
        scale, sinv, amax = torch.ones(1, device="cuda"), torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
        ds = [torch.empty(N, H, dtype=torch.uint8, device="cuda") for _ in xs]
        hs = []
        for d in ds:
            z = C.TensorHandle(C.DELAYED_TENSOR_SCALING)
            z.set(C.kRowwiseData, d, C.kFloat8E4M3).set(C.kRowwiseScaleInv, sinv).set(C.kScale, scale).set(C.kAmax, amax)
            hs.append(z)
        ms = timeit(lambda i: ops.norm_fwd(ln, xs[i % nset], gamma, b, 1e-5, hs[i % nset]))
        nb = 3 * n
        print(f"{name} {N}x{H} -> fp8       : {ms * 1e3:7.1f} us {nb / ms / 1e6:6.0f} GB/s {nb / ms / 1e6 / HBM * 100:5.1f}%", flush=True)
        del ds, hs
        for col in (False, True):
            outs = [mx_out(N, H, col) for _ in xs]
            hs = [o.handle() for o in outs]
            ms = timeit(lambda i: ops.norm_fwd(ln, xs[i % nset], gamma, b, 1e-5, hs[i % nset]))
            nb = (4 if col else 3) * n + (n // 16 if col else n // 32)
            print(f"{name} {N}x{H} -> mxfp8 {'r+c' if col else 'row'}: {ms * 1e3:7.1f} us {nb / ms / 1e6:6.0f} GB/s "
                  f"{nb / ms / 1e6 / HBM * 100:5.1f}%", flush=True)
            del outs, hs
        # unfused comparison: norm -> bf16, then MXFP8 quantize of the bf16
        q = [ops.mxfp8_quantize(x, swizzled_scales=True) for x in xs]
        hy = [C.plain(y) for y in ys]

        def unfused(i):
            ops.norm_fwd(ln, xs[i % nset], gamma, b, 1e-5, hy[i % nset])
            ops.mxfp8_quantize(ys[i % nset], swizzled_scales=True, out=q[i % nset])
        ms = timeit(unfused)
        print(f"{name} {N}x{H} unfused norm->bf16->mxfp8 r+c: {ms * 1e3:7.1f} us", flush=True)
        del q
        # backward
        dzs = [torch.randn(N, H, device="cuda").bfloat16() for _ in xs]
        mu = torch.zeros(N, device="cuda") if ln else None
        rsig = torch.ones(N, device="cuda")
        ms = timeit(lambda i: ops.norm_bwd(dzs[i % nset], xs[i % nset], mu, rsig, gamma))
        nb = 6 * n
        print(f"{name} {N}x{H} backward       : {ms * 1e3:7.1f} us {nb / ms / 1e6:6.0f} GB/s {nb / ms / 1e6 / HBM * 100:5.1f}%", flush=True)
        del dzs, ys
    del xs
