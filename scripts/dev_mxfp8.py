"""MXFP8 quantize timing on the config-4 tensors (development aid)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

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


tot = 0.0
for rows, cols in ((8192, 4096), (14336, 4096), (8192, 14336), (16384, 8192), (4096, 4096)):
    nset = min(8, max(2, int(600e6 // (rows * cols * 2)) + 1))
    xs = [torch.randn(rows, cols, device="cuda").bfloat16() for _ in range(nset)]
    n = rows * cols
    for kw, nb in (({"swizzled_scales": True}, 4 * n + n // 16), ({}, 4 * n + n // 16),
                   ({"columnwise": False, "swizzled_scales": True}, 3 * n + n // 32),
                   ({"rowwise": False, "swizzled_scales": True}, 3 * n + n // 32)):
        outs = [ops.mxfp8_quantize(x, **kw) for x in xs]
        ms = timeit(lambda i: ops.mxfp8_quantize(xs[i % nset], out=outs[i % nset], **kw))
        print(f"mxfp8 {rows}x{cols} {kw}: {ms * 1e3:7.1f} us {nb / ms / 1e6:6.0f} GB/s {nb / ms / 1e6 / HBM * 100:5.1f}%",
              flush=True)
        if kw == {"swizzled_scales": True} and (rows, cols) in ((8192, 4096), (14336, 4096), (8192, 14336)):
            tot += ms
    del xs, outs
print(f"config-4 step: {tot * 1e3:.1f} us -> {851968000 / tot / 1e6:.0f} GB/s = {851968000 / tot / 1e6 / HBM * 100:.1f}%")
