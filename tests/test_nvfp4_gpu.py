"""GPU parity tests for NVFP4: quantization (1-D, 2-D, RHT), amax, swizzle, dequantize
and the NVFP4 GEMM, against the oracle (itself pinned to the reference's C++ host
reference and pure-torch NVFP4 quantizer, tests/test_oracle.py) and against the
committed golden fixtures of the reference's torch quantizer (tests/golden/).

Bit-exact bar (atol = rtol = 0) on data, scales and both amaxes, as in the reference's
tests/pytorch/nvfp4/test_nvfp4_rht_quantize_exact.py:133-151 and
test_nvfp4_quantize_exact.py; GEMM within atol = rtol = 8e-3
(tests/pytorch/nvfp4/test_nvfp4_gemm_exact.py:239).
"""
import glob
import os

import numpy as np
import pytest
import torch

from tests.util import bits16, make_input, u8

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")

ops = None
O = None


@pytest.fixture(scope="module", autouse=True)
def _mods():
    global ops, O
    from oracle import oracle as _o
    from transformerengine_b200 import ops as _ops
    ops, O = _ops, _o


def _ref(x, with_rht, two_d=False):
    """Oracle NVFP4: rowwise on x, columnwise on x^T (RHT'd when asked)."""
    xb = bits16(x)
    amax_r = O.amax(xb)
    fn = O.nvfp4_quantize_2d if two_d else O.nvfp4_quantize_1d
    d, s = fn(xb, amax_r)
    xt = O.rht_transposed_bf16(xb) if with_rht else O.transpose_bf16(xb)
    amax_c = O.amax(xt) if with_rht else amax_r
    dt, st = fn(xt, amax_c)
    return d, s, dt, st, amax_r, amax_c


SHAPES = [(128, 128), (256, 384), (512, 256), (64, 128), (16, 16), (304, 304), (1024, 2048),
          (48, 80)]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("kind", ["randn", "uniform", "outliers"])
@pytest.mark.parametrize("with_rht", [False, True])
def test_nvfp4_quantize_bit_exact(shape, kind, with_rht):
    rows, cols = shape
    x = make_input(rows, cols, kind).cuda()
    q = ops.nvfp4_quantize(x, with_rht=with_rht)
    d, s, dt, st, amax_r, amax_c = _ref(x, with_rht)
    assert q.amax_rowwise.item() == amax_r
    assert q.amax_columnwise.item() == amax_c
    assert np.array_equal(u8(q.rowwise_scale_inv)[:rows, :cols // 16], s)
    assert np.array_equal(u8(q.rowwise_data), d)
    assert np.array_equal(u8(q.columnwise_scale_inv)[:cols, :rows // 16], st)
    assert np.array_equal(u8(q.columnwise_data), dt)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "nvfp4_rht_*.npz"))),
                         ids=os.path.basename)
def test_nvfp4_rht_matches_reference_torch_golden(path):
    """Directly against outputs of the reference's NVFP4QuantizerRef (with_rht=True)."""
    g = np.load(path)
    rows, cols, kind, seed = os.path.basename(path)[len("nvfp4_rht_"):-4].replace("x", "_").split("_")
    rows, cols = int(rows), int(cols)
    x = make_input(rows, cols, kind, int(seed)).cuda()
    q = ops.nvfp4_quantize(x, with_rht=True)
    assert q.amax_rowwise.item() == g["amax_row"][0]
    assert q.amax_columnwise.item() == g["amax_col"][0]
    assert np.array_equal(u8(q.rowwise_data), g["data"])
    assert np.array_equal(u8(q.rowwise_scale_inv)[:rows, :cols // 16], g["scale"])
    assert np.array_equal(u8(q.columnwise_data), g["data_t"])
    assert np.array_equal(u8(q.columnwise_scale_inv)[:cols, :rows // 16], g["scale_t"])


@pytest.mark.parametrize("shape", [(128, 128), (256, 512), (304, 304)])
def test_nvfp4_2d_quantize_bit_exact(shape):
    rows, cols = shape
    x = (make_input(rows, cols, "randn") * 0.02).bfloat16().cuda()
    q = ops.nvfp4_quantize(x, with_2d_quantization=True)
    d, s, dt, st, _, _ = _ref(x, False, two_d=True)
    assert np.array_equal(u8(q.rowwise_scale_inv)[:rows, :cols // 16], s)
    assert np.array_equal(u8(q.rowwise_data), d)
    assert np.array_equal(u8(q.columnwise_scale_inv)[:cols, :rows // 16], st)
    assert np.array_equal(u8(q.columnwise_data), dt)


@pytest.mark.parametrize("rowwise,columnwise", [(True, False), (False, True)])
def test_nvfp4_single_direction_and_fp32_input(rowwise, columnwise):
    x = make_input(256, 192, "randn", dtype=torch.float32).cuda()
    q = ops.nvfp4_quantize(x, rowwise=rowwise, columnwise=columnwise)
    xf = x.cpu().numpy()
    amax = O.amax(xf, 0)
    if rowwise:
        d, s = O.nvfp4_quantize_1d(xf, amax, code=0)
        assert np.array_equal(u8(q.rowwise_data), d)
        assert np.array_equal(u8(q.rowwise_scale_inv)[:256, :12], s)
    if columnwise:
        d, s = O.nvfp4_quantize_1d(np.ascontiguousarray(xf.T), amax, code=0)
        assert np.array_equal(u8(q.columnwise_data), d)
        assert np.array_equal(u8(q.columnwise_scale_inv)[:192, :16], s)


@pytest.mark.parametrize("shape", [(128, 128), (256, 384), (304, 304)])
def test_nvfp4_fused_swizzle_and_roundtrip(shape):
    x = make_input(*shape, "randn").cuda()
    compact = ops.nvfp4_quantize(x, with_rht=True)
    fused = ops.nvfp4_quantize(x, with_rht=True, swizzled_scales=True)
    two_step = ops.swizzle_scales(compact)
    assert torch.equal(fused.rowwise_data, compact.rowwise_data)
    assert torch.equal(fused.columnwise_data, compact.columnwise_data)
    assert torch.equal(fused.rowwise_scale_inv.flatten(), two_step.rowwise_scale_inv.flatten())
    assert torch.equal(fused.columnwise_scale_inv.flatten(), two_step.columnwise_scale_inv.flatten())
    assert np.array_equal(u8(two_step.rowwise_scale_inv).ravel(),
                          O.swizzle_scales(u8(compact.rowwise_scale_inv)))
    back = ops.swizzle_scales(two_step, inverse=True)
    assert torch.equal(back.columnwise_scale_inv, compact.columnwise_scale_inv)


def test_nvfp4_dequantize_matches_oracle():
    x = make_input(256, 256, "randn").cuda()
    q = ops.nvfp4_quantize(x)
    got = ops.dequantize(q, torch.float32).cpu().numpy()
    ref = O.nvfp4_dequantize(u8(q.rowwise_data), u8(q.rowwise_scale_inv), q.amax_rowwise.item())
    assert np.array_equal(got, ref)


def test_unfused_hadamard_transform_matches_oracle():
    from transformerengine_b200 import cabi as C
    x = make_input(256, 128, "randn").cuda()
    out_t = torch.empty(128, 256, dtype=torch.bfloat16, device="cuda")
    h = C.TensorHandle(C.DELAYED_TENSOR_SCALING).set(C.kColumnwiseData, out_t)
    inp = C.plain(x)
    C.check(C.lib().nvte_hadamard_transform_status(inp.ptr, h.ptr, 0, ops.RHT_SIGN_MASK,
                                                   C.stream_ptr()))
    assert np.array_equal(bits16(out_t), O.rht_transposed_bf16(bits16(x)))


@pytest.mark.parametrize("layout", ["TN", "NN", "NT"])
@pytest.mark.parametrize("n,m,k", [(128, 128, 256), (256, 384, 512), (512, 1024, 1024),
                                   (144, 272, 96)])
def test_nvfp4_gemm(layout, n, m, k):
    """fprop (TN), dgrad (NN: weight columnwise), wgrad (NT: both columnwise, RHT along K
    cancels because the same orthogonal 16x16 rotation is applied to both operands)."""
    g = torch.Generator().manual_seed(4)
    sa = (m, k) if layout[0] == "T" else (k, m)
    sb = (k, n) if layout[1] == "T" else (n, k)
    if any(d % 16 for d in sa + sb):
        pytest.skip("NVFP4 columnwise needs multiples of 16")
    a = (torch.randn(*sa, generator=g) * 0.5).bfloat16().cuda()
    b = torch.randn(*sb, generator=g).bfloat16().cuda()
    rht = layout == "NT"
    qa = ops.nvfp4_quantize(a, with_rht=rht, swizzled_scales=True)
    qb = ops.nvfp4_quantize(b, with_rht=rht, swizzled_scales=True)
    out = ops.general_gemm(qa, qb, layout, torch.float32)
    ca, cb = layout[0] != "T", layout[1] == "T"
    qa_c, qb_c = ops.swizzle_scales(qa, inverse=True), ops.swizzle_scales(qb, inverse=True)

    def deq(q, colwise):
        if colwise:
            return O.nvfp4_dequantize(u8(q.columnwise_data), u8(q.columnwise_scale_inv),
                                      q.amax_columnwise.item())
        return O.nvfp4_dequantize(u8(q.rowwise_data), u8(q.rowwise_scale_inv), q.amax_rowwise.item())

    A = deq(qa_c, ca)   # [m, k] either way: columnwise NVFP4 data is stored transposed
    B = deq(qb_c, cb)   # [n, k]
    ref = O.gemm_tn_ref(A, B)
    got = out.cpu().numpy()
    assert np.abs(got - ref).max() <= 2e-5 * np.abs(ref).max()
    out_bf16 = ops.general_gemm(qa, qb, layout, torch.bfloat16)
    np.testing.assert_allclose(out_bf16.float().cpu().numpy(), ref, rtol=8e-3, atol=8e-3)


def test_nvfp4_full_size_properties():
    """BASELINE config 3 (16384x8192 bf16, RHT): properties + oracle on a slab."""
    x = torch.randn(16384, 8192, device="cuda", dtype=torch.bfloat16)
    q = ops.nvfp4_quantize(x, with_rht=True)
    q2 = ops.nvfp4_quantize(x * 2.0, with_rht=True)      # power-of-two linearity: same bytes
    assert torch.equal(q2.rowwise_data, q.rowwise_data)
    assert torch.equal(q2.columnwise_data, q.columnwise_data)
    assert torch.equal(q2.rowwise_scale_inv, q.rowwise_scale_inv)
    assert q2.amax_columnwise.item() == 2.0 * q.amax_columnwise.item()
    # the rowwise half equals a plain (no RHT) quantization
    q3 = ops.nvfp4_quantize(x, with_rht=False, columnwise=False)
    assert torch.equal(q3.rowwise_data, q.rowwise_data)
    # oracle on a 128-row slab, using the device amaxes (global quantities)
    xs = bits16(x[4096:4224])
    d, s = O.nvfp4_quantize_1d(xs, q.amax_rowwise.item())
    assert np.array_equal(u8(q.rowwise_data[4096:4224]), d)
    assert np.array_equal(u8(q.rowwise_scale_inv[4096:4224, :512]), s)
    dt, st = O.nvfp4_quantize_1d(O.rht_transposed_bf16(xs), q.amax_columnwise.item())
    assert np.array_equal(u8(q.columnwise_data[:, 2048:2112]), dt)
    assert np.array_equal(u8(q.columnwise_scale_inv[:8192, 256:264]), st)


@pytest.mark.parametrize("swizzled", [False, True])
@pytest.mark.parametrize("with_rht", [False, True])
def test_nvfp4_tensor_core_kernel_single_direction_variants(swizzled, with_rht):
    """bf16, 128-aligned shapes take the TMA/tensor-core kernel; its rowwise-only and
    columnwise-only instantiations must give the bytes of the two-direction one (which the tests
    above pin to the oracle). 512 tiles > 296 resident CTAs, so the persistent loop wraps."""
    x = make_input(2048, 4096, "outliers").cuda()
    both = ops.nvfp4_quantize(x, with_rht=with_rht, swizzled_scales=swizzled)
    row = ops.nvfp4_quantize(x, with_rht=with_rht, swizzled_scales=swizzled, columnwise=False)
    assert torch.equal(row.rowwise_data, both.rowwise_data)
    assert torch.equal(row.rowwise_scale_inv, both.rowwise_scale_inv)
    col = ops.nvfp4_quantize(x, with_rht=with_rht, swizzled_scales=swizzled, rowwise=False)
    assert torch.equal(col.columnwise_data, both.columnwise_data)
    assert torch.equal(col.columnwise_scale_inv, both.columnwise_scale_inv)
    # and a 128-row slab against the oracle (global amaxes taken from the device)
    d, s = O.nvfp4_quantize_1d(bits16(x[1024:1152]), both.amax_rowwise.item())
    assert np.array_equal(u8(both.rowwise_data[1024:1152]), d)
    if not swizzled:
        assert np.array_equal(u8(both.rowwise_scale_inv[1024:1152, :256]), s)


# --------------------------------------------------------------------------- #
# Stochastic rounding (reference tests/pytorch/nvfp4/test_nvfp4_sr_quantize.py)
# --------------------------------------------------------------------------- #
def _grid_tensor(rows, cols, seed=0):
    """Values that NVFP4 represents exactly: E2M1 grid points times a power-of-two scale per 16x16
    square, a +-6 on every square's diagonal (so every 1x16 and 16x1 block's amax is 6 * scale) and
    a global amax of exactly 6 * 448 (so the global encode scale is 1)."""
    g = np.random.default_rng(seed)
    grid = np.array([0, 0.5, 1, 1.5, 2, 3, 4, 6], np.float32)
    v = grid[g.integers(0, 8, (rows, cols))] * g.choice([-1.0, 1.0], (rows, cols)).astype(np.float32)
    i = np.arange(rows)[:, None] % 16 == np.arange(cols)[None, :] % 16
    v[i] = 6.0 * g.choice([-1.0, 1.0], int(i.sum()))
    sc = np.exp2(g.integers(-2, 9, (rows // 16, cols // 16))).astype(np.float32)
    sc[0, 0] = 448.0
    v *= np.kron(sc, np.ones((16, 16), np.float32))
    return torch.from_numpy(v).to(torch.bfloat16)


@pytest.mark.parametrize("shape", [(256, 384), (48, 80)])   # tensor-core kernel, generic kernel
def test_nvfp4_stochastic_rounding_is_exact_on_representable_values(shape):
    """Nothing to round => the stochastic cast must give the round-to-nearest bytes (pins the
    nibble order and the scale path of the SR instantiations)."""
    x = _grid_tensor(*shape).cuda()
    assert x.float().abs().max().item() == 2688.0
    rn = ops.nvfp4_quantize(x)
    sr = ops.nvfp4_quantize(x, stochastic_rounding=True)
    for name in ("rowwise_data", "rowwise_scale_inv", "columnwise_data", "columnwise_scale_inv"):
        assert torch.equal(getattr(sr, name), getattr(rn, name)), name
    back = O.nvfp4_dequantize(u8(sr.rowwise_data), u8(sr.rowwise_scale_inv)[:, :shape[1] // 16],
                              sr.amax_rowwise.item())
    assert np.array_equal(back, x.float().cpu().numpy())


@pytest.mark.parametrize("shape,dtype", [((256, 512), torch.bfloat16), ((80, 96), torch.bfloat16),
                                         ((128, 128), torch.float32)])
@pytest.mark.parametrize("with_rht,two_d", [(False, False), (True, False), (False, True)])
def test_nvfp4_stochastic_rounding_mean_beats_round_to_nearest(shape, dtype, with_rht, two_d):
    """The reference's criterion (test_nvfp4_sr_quantize.py:225-283): the mean of 50 stochastic
    casts is closer to x than the round-to-nearest cast, in both directions; plus: block scales do
    not depend on the rounding mode, every value lands on one of x's two grid neighbours, and
    successive casts use different random bits."""
    rows, cols = shape
    if with_rht and dtype == torch.float32:
        pytest.skip("the Hadamard path takes bf16 input")
    torch.manual_seed(1234)
    x = (torch.randn(rows, cols, device="cuda") * 2 - 1).to(dtype)
    if with_rht:
        y = (O.rht_transposed_bf16(bits16(x)).astype(np.uint32) << 16).view(np.float32)
    else:
        y = x.float().t().contiguous().cpu().numpy()
    xf = x.float().cpu().numpy()

    def deq(q):
        r = O.nvfp4_dequantize(u8(q.rowwise_data), u8(q.rowwise_scale_inv)[:rows, :cols // 16],
                               q.amax_rowwise.item())
        c = O.nvfp4_dequantize(u8(q.columnwise_data), u8(q.columnwise_scale_inv)[:cols, :rows // 16],
                               q.amax_columnwise.item())
        return r, c

    kw = dict(with_rht=with_rht, with_2d_quantization=two_d)
    rn = ops.nvfp4_quantize(x, **kw)
    r_rn, c_rn = deq(rn)
    n = 50
    acc_r, acc_c = np.zeros_like(xf, np.float64), np.zeros_like(y, np.float64)
    prev = None
    for _ in range(n):
        sr = ops.nvfp4_quantize(x, stochastic_rounding=True, **kw)
        assert torch.equal(sr.rowwise_scale_inv, rn.rowwise_scale_inv)
        assert torch.equal(sr.columnwise_scale_inv, rn.columnwise_scale_inv)
        if prev is not None:
            assert not torch.equal(prev, sr.rowwise_data)
        prev = sr.rowwise_data
        r, c = deq(sr)
        acc_r += r
        acc_c += c
    # the last cast: every value within one grid step of x (|err_sr| <= 2 * max|err_rn| per block
    # is implied by: SR picks a neighbour, RN the nearer neighbour; a neighbour pair spans at most
    # 2 (E2M1 steps) * block scale * global scale)
    step = 2.0 * np.repeat(O.nvfp4_dequantize(np.full((rows, cols // 2), 0x22, np.uint8),  # 1.0 * scale
                                              u8(rn.rowwise_scale_inv)[:rows, :cols // 16],
                                              rn.amax_rowwise.item())[:, ::16], 16, axis=1)
    assert np.all(np.abs(r - xf) <= step + 1e-6)
    rmse = lambda a, b: float(np.sqrt(np.mean((a - b) ** 2)))
    assert rmse(acc_r / n, xf) < 0.5 * rmse(r_rn, xf)
    assert rmse(acc_c / n, y) < 0.5 * rmse(c_rn, y)
    # a single stochastic cast is noisier than round-to-nearest (it really is rounding randomly)
    assert rmse(r, xf) > rmse(r_rn, xf)


def test_nvfp4_stochastic_rounding_needs_rng_state():
    from transformerengine_b200 import cabi as C
    x = make_input(128, 128, "randn").cuda()
    out = ops.nvfp4_quantize(x)
    cfg = C.QuantConfig(stochastic_rounding=True)
    with pytest.raises(RuntimeError, match="RNG state"):
        C.check(C.lib().nvte_quantize_v2_status(C.plain(x).ptr, out.handle().ptr, cfg.ptr, C.stream_ptr()))


@pytest.mark.parametrize("layout", ["TN", "NT"])
@pytest.mark.parametrize("k", [384, 640, 96])
def test_nvfp4_gemm_small_output_with_partial_last_k_block(layout, k):
    """K % 256 != 0 with a <= 128-row output: the last k-block's scale atoms reach past the row
    block's own (regression: the single-CTA kernel read them linearly past the buffer -> NaN)."""
    from transformerengine_b200 import ops
    torch.manual_seed(k)
    m, n = 128, 512
    a = torch.randn((n, k) if layout == "TN" else (k, n), device="cuda").bfloat16()
    b = torch.randn((m, k) if layout == "TN" else (k, m), device="cuda").bfloat16()
    qa = ops.nvfp4_quantize(a, with_rht=False, swizzled_scales=True)
    qb = ops.nvfp4_quantize(b, with_rht=False, swizzled_scales=True)
    for _ in range(3):   # neighbouring allocations differ between calls
        junk = torch.full((1 << 20,), 0xFF, dtype=torch.uint8, device="cuda")  # NaN-encoded scales nearby
        out = ops.general_gemm(qa, qb, layout, torch.float32)
        del junk
        da = ops.dequantize(qa, columnwise=layout[0] != "T").float()
        db = ops.dequantize(qb, columnwise=layout[1] == "T").float()
        ref = db @ da.t()
        assert torch.isfinite(out).all()
        assert ((out - ref).norm() / ref.norm()).item() < 1e-5


@pytest.mark.parametrize("use_row_a,use_row_b", [(True, True), (False, True), (True, False), (False, False)])
@pytest.mark.parametrize("alpha_in", [1.0, 0.375])
def test_nvfp4_compute_per_tensor_scale_matches_oracle(use_row_a, use_row_b, alpha_in):
    """`nvte_nvfp4_compute_per_tensor_scale` (recipe/nvfp4.cu:72-78) against the oracle's
    alpha, bit for bit, for every rowwise / columnwise amax selection."""
    a = make_input(128, 256, "randn", seed=7).cuda()
    b = make_input(256, 256, "outliers", seed=8).cuda()
    qa = ops.nvfp4_quantize(a, with_rht=True)
    qb = ops.nvfp4_quantize(b, with_rht=True)
    assert qa.amax_rowwise.item() != qa.amax_columnwise.item()
    out = ops.nvfp4_compute_per_tensor_scale(qa, use_row_a, qb, use_row_b, alpha_in)
    am_a = (qa.amax_rowwise if use_row_a else qa.amax_columnwise).item()
    am_b = (qb.amax_rowwise if use_row_b else qb.amax_columnwise).item()
    ref = O.lib().teo_nvfp4_alpha(alpha_in, am_a, am_b)
    assert np.float32(out.item()).tobytes() == np.float32(ref).tobytes(), (out.item(), ref)
