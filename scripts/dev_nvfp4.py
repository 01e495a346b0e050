"""NVFP4 cast-pass timing (development aid): amax given, only the cast kernel is timed.

    python scripts/dev_nvfp4.py [rows cols]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from transformerengine_b200 import cabi as C  # noqa: E402
from transformerengine_b200 import ops  # noqa: E402

R, Cc = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (16384, 8192)
HBM = 6457.1
xs = [torch.randn(R, Cc, device="cuda").bfloat16() for _ in range(2)]
n = R * Cc


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


for rht in (True, False):
    for swz in (True, False):
        outs = [ops.nvfp4_quantize(x, with_rht=rht, swizzled_scales=swz) for x in xs]
        ins = [C.plain(x) for x in xs]
        m = C.plain(ops.rht_matrix(xs[0].device))
        cfg = C.QuantConfig()
        sp = C.stream_ptr()
        lib = C.lib()
        if rht:
            fn = lambda i: C.check(lib.nvte_quantize_with_hadamard_transform_status(
                ins[i % 2].ptr, outs[i % 2].handle().ptr, m.ptr, cfg.ptr, sp))
        else:
            fn = lambda i: C.check(lib.nvte_quantize_v2_status(ins[i % 2].ptr, outs[i % 2].handle().ptr, cfg.ptr, sp))
        ms = timeit(fn)
        b = n * 2 + n + 2 * n // 16
        print(f"nvfp4 cast rht={int(rht)} swz={int(swz)} {R}x{Cc}: {ms * 1e3:8.1f} us  {b / ms / 1e6:6.0f} GB/s  "
              f"{b / ms / 1e6 / HBM * 100:5.1f}% of measured HBM", flush=True)
ar, ac = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
ms = timeit(lambda i: ops.hadamard_transform_amax(xs[i % 2], ar, ac))
print(f"nvfp4 rht amax {R}x{Cc}: {ms * 1e3:8.1f} us  {n * 2 / ms / 1e6:6.0f} GB/s  {n * 2 / ms / 1e6 / HBM * 100:5.1f}%")
