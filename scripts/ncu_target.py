"""Small single-op driver for `ncu` captures: python scripts/ncu_target.py <op> [iters]
ops: mxfp8_c4, nvfp4_rht, nvfp4_rht_swz, nvfp4_amax, mxfp8, gemm_mxfp8, gemm_nvfp4, gemm_fp8, fp8_cast, amax"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from transformerengine_b200 import ops  # noqa: E402

op = sys.argv[1]
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 4
torch.manual_seed(0)
if op.startswith("gemm"):
    kind = op.split("_")[1]
    n, m, k = (16384, 8192, 8192) if kind == "nvfp4" else (8192, 14336, 4096)
    a = (torch.randn(m, k, device="cuda") * 0.5).bfloat16()
    b = torch.randn(n, k, device="cuda").bfloat16()
    if kind == "mxfp8":
        qa, qb = ops.mxfp8_quantize(a, swizzled_scales=True), ops.mxfp8_quantize(b, swizzled_scales=True)
    elif kind == "fp8":
        s = torch.ones(1, device="cuda")
        qa, qb = ops.fp8_quantize(a, s), ops.fp8_quantize(b, s)
    else:
        qa = ops.nvfp4_quantize(a, with_rht=False, swizzled_scales=True)
        qb = ops.nvfp4_quantize(b, with_rht=False, swizzled_scales=True)
    out = torch.empty(n, m, dtype=torch.bfloat16, device="cuda")
    for _ in range(iters):
        ops.general_gemm(qa, qb, "TN", out=out)
elif op.startswith("norm"):
    # norm_<rms|ln>_<bf16|mx>: fused norm forward on a 16384 x 8192 activation
    from transformerengine_b200 import cabi as C
    N, H = 16384, 8192
    ln = "_ln" in op
    xs = [torch.randn(N, H, device="cuda").bfloat16() for _ in range(2)]
    gamma, beta = torch.rand(H, device="cuda").bfloat16(), torch.rand(H, device="cuda").bfloat16()
    if op.endswith("bf16"):
        ys = [torch.empty_like(x) for x in xs]
        hs = [C.plain(y) for y in ys]
    else:
        rs, cs = ops.mxfp8_scale_shapes(N, H)
        outs = [ops.MXFP8TensorStorage(
            rowwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
            rowwise_scale_inv=torch.zeros(rs, dtype=torch.uint8, device="cuda"),
            columnwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
            columnwise_scale_inv=torch.zeros(cs, dtype=torch.uint8, device="cuda"),
            fp8_dtype=torch.float8_e4m3fn, shape=(N, H), with_gemm_swizzled_scales=True) for _ in xs]
        hs = [o.handle() for o in outs]
    for i in range(iters):
        ops.norm_fwd(ln, xs[i % 2], gamma, beta if ln else None, 1e-5, hs[i % 2])
elif op == "mxfp8_c4_multi":
    # the bench's step as the bench launches it: the three config-4 tensors in ONE launch
    shapes = ((8192, 4096), (14336, 4096), (8192, 14336))
    sets = [[torch.randn(r, c, device="cuda").bfloat16() for r, c in shapes] for _ in range(2)]
    outs = [[ops.mxfp8_quantize(x, swizzled_scales=True) for x in st] for st in sets]
    for i in range(iters):
        ops.multi_tensor_quantize(sets[i % 2], outs[i % 2])
elif op == "mxfp8_c4":
    # the bench's step: config-4 tensors x, W, dy (rowwise + columnwise, swizzled scales)
    shapes = ((8192, 4096), (14336, 4096), (8192, 14336))
    sets = [[torch.randn(r, c, device="cuda").bfloat16() for r, c in shapes] for _ in range(2)]
    outs = [[ops.mxfp8_quantize(x, swizzled_scales=True) for x in st] for st in sets]
    for i in range(iters):
        for x, o in zip(sets[i % 2], outs[i % 2]):
            ops.mxfp8_quantize(x, swizzled_scales=True, out=o)
else:
    R, C = (16384, 8192) if op.startswith("nvfp4") else (8192, 14336)
    xs = [torch.randn(R, C, device="cuda").bfloat16() for _ in range(2)]
    if op in ("nvfp4_rht", "nvfp4_rht_swz"):
        outs = [ops.nvfp4_quantize(x, with_rht=True, swizzled_scales=op.endswith("swz")) for x in xs]
        for i in range(iters):
            ops.nvfp4_quantize(xs[i % 2], with_rht=True, swizzled_scales=op.endswith("swz"), out=outs[i % 2])
    elif op == "nvfp4_amax":
        ar, ac = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
        for i in range(iters):
            ops.hadamard_transform_amax(xs[i % 2], ar, ac)
    elif op == "mxfp8":
        outs = [ops.mxfp8_quantize(x, swizzled_scales=True) for x in xs]
        for i in range(iters):
            ops.mxfp8_quantize(xs[i % 2], swizzled_scales=True, out=outs[i % 2])
    elif op == "fp8_cast":
        s, am = torch.ones(1, device="cuda"), torch.zeros(1, device="cuda")
        outs = [ops.fp8_quantize(x, s) for x in xs]
        for i in range(iters):
            ops.fp8_quantize(xs[i % 2], s, amax=am, out=outs[i % 2])
    elif op == "amax":
        am = torch.zeros(1, device="cuda")
        for i in range(iters):
            ops.compute_amax(xs[i % 2], am)
torch.cuda.synchronize()
print("done", op)
