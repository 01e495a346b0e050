"""Short subset for compute-sanitizer (memcheck / racecheck): one launch of every hand-rolled-protocol kernel
(mbarrier rings, in-place smem staging, relaxed cluster arrives) on small multi-tile problems, each checked against
the oracle so that a sanitizer-clean run is also a correct one."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402
from tests.util import bits16, make_input, u8  # noqa: E402
from transformerengine_b200 import ops  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"


def run(name, fn):
    if which in ("all", name):
        fn()
        torch.cuda.synchronize()
        print("ok", name, flush=True)


def mxfp8():
    x = make_input(384, 512, "randn").cuda()          # 12 tiles of the TMA kernel
    q = ops.mxfp8_quantize(x, swizzled_scales=False)
    ref = O.mxfp8_quantize(bits16(x))
    assert np.array_equal(u8(q.rowwise_data), ref["rowwise"]) and np.array_equal(u8(q.columnwise_data), ref["colwise"])
    xs = [make_input(r, 256, "randn", seed=r).cuda() for r in (128, 384, 256)]
    outs = [ops.mxfp8_quantize(torch.zeros_like(t), swizzled_scales=True) for t in xs]
    ops.multi_tensor_quantize(xs, outs)               # one launch over the list
    for t, o in zip(xs, outs):
        assert torch.equal(o.rowwise_data, ops.mxfp8_quantize(t, swizzled_scales=True).rowwise_data)


def nvfp4():
    x = make_input(256, 384, "randn").cuda()
    q = ops.nvfp4_quantize(x, with_rht=True)
    xb = bits16(x)
    qr, _ = O.nvfp4_quantize_1d(xb, O.amax(xb))
    assert np.array_equal(u8(q.rowwise_data), qr)


def norm():
    import transformerengine_b200 as te
    from transformerengine_b200.normalization import rmsnorm_fwd
    x = make_input(256, 512, "randn").cuda()
    g = (torch.rand(512) + 0.5).bfloat16().cuda()
    rmsnorm_fwd(x, g, 1e-5, quantizer=te.MXFP8Quantizer(torch.float8_e4m3fn))
    rmsnorm_fwd(x, g, 1e-5)


def gemm():
    for kind in ("mxfp8", "nvfp4", "fp8"):
        for (n, m, k) in ((512, 768, 256), (96, 640, 160)):   # pair kernel (6 tiles) and the single-CTA BN=256 path
            if kind == "nvfp4" and k % 32:
                continue
            a = (torch.randn(m, k) * 0.5).bfloat16().cuda()
            b = torch.randn(n, k).bfloat16().cuda()
            if kind == "mxfp8":
                qa, qb = ops.mxfp8_quantize(a, swizzled_scales=True, columnwise=False), ops.mxfp8_quantize(b, swizzled_scales=True, columnwise=False)
            elif kind == "nvfp4":
                qa = ops.nvfp4_quantize(a, swizzled_scales=True, columnwise=False)
                qb = ops.nvfp4_quantize(b, swizzled_scales=True, columnwise=False)
            else:
                s = torch.ones(1, device="cuda")
                qa, qb = ops.fp8_quantize(a, s), ops.fp8_quantize(b, s)
            y = ops.general_gemm(qa, qb, "TN", torch.float32)
            ref = ops.dequantize(qb, torch.float32).double() @ ops.dequantize(qa, torch.float32).double().t()
            assert (y.double() - ref).abs().max() <= 2e-5 * ref.abs().max(), (kind, n, m, k)


run("mxfp8", mxfp8)
run("nvfp4", nvfp4)
run("norm", norm)
run("gemm", gemm)
print("done")
