"""Out-of-bounds WRITE guards (compute-sanitizer is closed on this GPU pool, see profiles/r02_sanitizer.txt): every
output buffer of the hand-rolled-protocol kernels is a window inside a larger poisoned allocation; after the kernel
the poison on both sides must be intact and the window must hold the same bytes as a plain run. Shapes are the ragged /
multi-tile ones where an off-by-one tile or scale atom would land outside the tensor."""
import pytest
import torch

from tests.util import make_input

pytestmark = pytest.mark.gpu
GUARD = 4096
POISON = 0xA5


class Guarded:
    """uint8 window of `nbytes` (16-byte aligned) inside a poisoned buffer."""

    def __init__(self, shape, dtype=torch.uint8):
        self.shape, self.dtype = tuple(shape), dtype
        n = 1
        for d in shape:
            n *= d
        self.nbytes = n * torch.empty(0, dtype=dtype).element_size()
        self.raw = torch.full((self.nbytes + 2 * GUARD,), POISON, dtype=torch.uint8, device="cuda")
        self.view = self.raw[GUARD:GUARD + self.nbytes].view(dtype).view(*shape)

    def intact(self):
        return bool((self.raw[:GUARD] == POISON).all() and (self.raw[GUARD + self.nbytes:] == POISON).all())


def _ops():
    from transformerengine_b200 import ops
    return ops


@pytest.mark.parametrize("shape", [(128, 128), (384, 512), (200, 96), (993, 512), (1024, 2176)])
@pytest.mark.parametrize("swizzled", [False, True])
def test_mxfp8_quantize_writes_stay_inside_their_buffers(shape, swizzled):
    ops = _ops()
    x = make_input(*shape, "randn").cuda()
    ref = ops.mxfp8_quantize(x, swizzled_scales=swizzled)
    g = {k: Guarded(getattr(ref, k).shape) for k in ("rowwise_data", "columnwise_data", "rowwise_scale_inv", "columnwise_scale_inv")}
    for k in ("rowwise_scale_inv", "columnwise_scale_inv"):
        g[k].view.zero_()
    out = ops.MXFP8TensorStorage(rowwise_data=g["rowwise_data"].view, rowwise_scale_inv=g["rowwise_scale_inv"].view,
                                 columnwise_data=g["columnwise_data"].view, columnwise_scale_inv=g["columnwise_scale_inv"].view,
                                 fp8_dtype=torch.float8_e4m3fn, shape=shape, with_gemm_swizzled_scales=swizzled)
    ops.mxfp8_quantize(x, swizzled_scales=swizzled, out=out)
    torch.cuda.synchronize()
    for k, b in g.items():
        assert b.intact(), f"{k}: write outside the buffer"
        assert torch.equal(b.view, getattr(ref, k)), k


def test_multi_tensor_quantize_writes_stay_inside_their_buffers():
    ops = _ops()
    shapes = [(128, 256), (384, 256), (256, 1024), (128, 128)]
    xs = [make_input(r, c, "randn", seed=r + c).cuda() for r, c in shapes]
    refs = [ops.mxfp8_quantize(x, swizzled_scales=True) for x in xs]
    guards, outs = [], []
    for r in refs:
        g = {k: Guarded(getattr(r, k).shape) for k in ("rowwise_data", "columnwise_data", "rowwise_scale_inv", "columnwise_scale_inv")}
        guards.append(g)
        outs.append(ops.MXFP8TensorStorage(rowwise_data=g["rowwise_data"].view, rowwise_scale_inv=g["rowwise_scale_inv"].view,
                                           columnwise_data=g["columnwise_data"].view,
                                           columnwise_scale_inv=g["columnwise_scale_inv"].view,
                                           fp8_dtype=torch.float8_e4m3fn, shape=r.shape, with_gemm_swizzled_scales=True))
    ops.multi_tensor_quantize(xs, outs)
    torch.cuda.synchronize()
    for g, r in zip(guards, refs):
        for k, b in g.items():
            assert b.intact(), f"{k}: write outside the buffer"
            assert torch.equal(b.view, getattr(r, k)), k


@pytest.mark.parametrize("shape", [(128, 128), (256, 384), (304, 304), (1024, 2048)])
@pytest.mark.parametrize("with_rht", [False, True])
def test_nvfp4_quantize_writes_stay_inside_their_buffers(shape, with_rht):
    ops = _ops()
    x = make_input(*shape, "randn").cuda()
    ref = ops.nvfp4_quantize(x, with_rht=with_rht, swizzled_scales=True)
    keys = ("rowwise_data", "columnwise_data", "rowwise_scale_inv", "columnwise_scale_inv")
    g = {k: Guarded(getattr(ref, k).shape) for k in keys}
    for k in keys[2:]:
        g[k].view.zero_()
    am = Guarded((2,), torch.float32)
    am.view.zero_()
    out = ops.NVFP4TensorStorage(rowwise_data=g["rowwise_data"].view, rowwise_scale_inv=g["rowwise_scale_inv"].view,
                                 columnwise_data=g["columnwise_data"].view, columnwise_scale_inv=g["columnwise_scale_inv"].view,
                                 amax_rowwise=am.view[0:1], amax_columnwise=am.view[1:2], shape=shape,
                                 with_gemm_swizzled_scales=True)
    ops.nvfp4_quantize(x, with_rht=with_rht, swizzled_scales=True, out=out)
    torch.cuda.synchronize()
    assert am.intact()
    for k, b in g.items():
        assert b.intact(), f"{k}: write outside the buffer"
        assert torch.equal(b.view, getattr(ref, k)), k


@pytest.mark.parametrize("kind", ["mxfp8", "nvfp4", "fp8"])
@pytest.mark.parametrize("n,m,k", [(512, 768, 256), (96, 640, 160), (200, 136, 96), (1000, 520, 512), (128, 384, 128)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_gemm_output_writes_stay_inside_the_output(kind, n, m, k, out_dtype):
    """Pair kernel and both single-CTA tile widths, ragged rows / columns; (96, 640, 160) and (128, 384, 128) are the
    BN = 256 shapes with an odd number of 128-row scale blocks (the skipped scale copy)."""
    ops = _ops()
    if kind == "nvfp4" and (k % 32 or m % 16 or n % 16):
        pytest.skip("NVFP4 needs K % 32 == 0")
    a = (torch.randn(m, k, generator=torch.Generator().manual_seed(1)) * 0.5).bfloat16().cuda()
    b = torch.randn(n, k, generator=torch.Generator().manual_seed(2)).bfloat16().cuda()
    if kind == "mxfp8":
        qa, qb = (ops.mxfp8_quantize(t, swizzled_scales=True, columnwise=False) for t in (a, b))
    elif kind == "nvfp4":
        qa, qb = (ops.nvfp4_quantize(t, swizzled_scales=True, columnwise=False) for t in (a, b))
    else:
        s = torch.ones(1, device="cuda")
        qa, qb = ops.fp8_quantize(a, s), ops.fp8_quantize(b, s)
    ref = ops.general_gemm(qa, qb, "TN", out_dtype)
    g = Guarded((n, m), out_dtype)
    ops.general_gemm(qa, qb, "TN", out=g.view)
    torch.cuda.synchronize()
    assert g.intact(), "GEMM wrote outside its output"
    assert torch.equal(g.view, ref)
    assert torch.isfinite(ref.float()).all()
