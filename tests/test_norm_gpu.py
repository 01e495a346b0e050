"""GPU parity tests of LayerNorm / RMSNorm through the C ABI.

The oracle (oracle/teo_norm.c) is pinned bit-for-bit to the reference's host references
(tests/test_oracle.py, tests/golden/norm_*.npz). Here:
  * the row statistics are sums in a different order => tolerance: rtol 1e-5 against the oracle's
    sequential fp32 sums, rtol 1e-6 against float64 statistics (the reference's own tests use
    1e-5 .. 5e-3, tests/cpp/operator/test_normalization.cu:205-240);
  * everything downstream of the DEVICE's statistics is elementwise => compared exactly: the
    16-bit output bits, the per-tensor FP8 bytes and amax, the MXFP8 bytes and E8M0 scales in both
    directions (the oracle quantizers applied to the oracle's fp32 y);
  * backward against the oracle within bf16 rounding.
"""
import numpy as np
import pytest
import torch

from tests.util import bits16, make_input, u8

pytestmark = pytest.mark.gpu

ops = C = O = None


@pytest.fixture(scope="module", autouse=True)
def _mods():
    global ops, C, O
    from oracle import oracle as _o
    from transformerengine_b200 import cabi as _c
    from transformerengine_b200 import ops as _ops
    ops, C, O = _ops, _c, _o


EPS = 1e-5
# N % 32 != 0 -> ghost rows in the last slab; H picks the cluster size: 1, 1 (24 sub-tiles), 4, 5, 8, 12, 16
SHAPES = [(1, 128), (32, 64), (48, 256), (33, 768), (1000, 512), (64, 1024), (32, 1536), (128, 4096), (96, 5120), (64, 8192),
          (32, 12288), (32, 16384)]


def _params(H, zc, seed, dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    gamma = (torch.rand(H, generator=g) * 2 - (1.0 if zc else 0.0)).to(dtype).cuda()
    beta = (torch.rand(H, generator=g) - 0.5).to(dtype).cuda()
    return gamma, beta


def _check_stats(x, mu, rs, ln):
    # against the oracle (sequential fp32 sums, itself ~sqrt(H) ulp off) ...
    mu_o, rs_o = O.norm_stats(bits16(x), EPS, ln)
    np.testing.assert_allclose(rs.cpu().numpy(), rs_o, rtol=1e-5)
    if ln:
        np.testing.assert_allclose(mu.cpu().numpy(), mu_o, rtol=1e-4, atol=1e-6)
    # ... and tighter against float64 statistics (the kernel's blocked sums are the better ones)
    x64 = x.double().cpu()
    m64 = x64.mean(1, keepdim=True) if ln else torch.zeros(x.shape[0], 1, dtype=torch.float64)
    rs64 = torch.rsqrt(((x64 - m64) ** 2).mean(1) + EPS)
    np.testing.assert_allclose(rs.cpu().numpy(), rs64.numpy(), rtol=1e-6)
    if ln:
        np.testing.assert_allclose(mu.cpu().numpy(), m64[:, 0].numpy(), rtol=1e-5, atol=1e-7)


def _y_oracle(x, gamma, beta, mu, rs, ln, zc):
    return O.norm_output(bits16(x), bits16(gamma), bits16(beta) if ln else None,
                         mu.cpu().numpy() if ln else None, rs.cpu().numpy(), zero_centered=zc)


def _bf16(a):
    return bits16(torch.from_numpy(np.ascontiguousarray(a)).to(torch.bfloat16))


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("ln", [True, False], ids=["layernorm", "rmsnorm"])
@pytest.mark.parametrize("zc", [False, True], ids=["gamma", "zero_centered"])
def test_norm_fwd_bf16_output_and_amax(shape, ln, zc):
    N, H = shape
    x = make_input(N, H, "randn", seed=N + H).cuda()
    gamma, beta = _params(H, zc, 1)
    y = torch.empty_like(x)
    amax = torch.zeros(1, device="cuda")
    z = C.plain(y).set(C.kAmax, amax)
    mu, rs = ops.norm_fwd(ln, x, gamma, beta if ln else None, EPS, z, zc)
    _check_stats(x, mu, rs, ln)
    ref = _bf16(_y_oracle(x, gamma, beta, mu, rs, ln, zc))
    assert np.array_equal(bits16(y), ref)
    assert amax.item() == y.float().abs().max().item()


@pytest.mark.parametrize("shape", [(48, 256), (64, 1024), (128, 4096), (64, 8192)])
@pytest.mark.parametrize("ln", [True, False], ids=["layernorm", "rmsnorm"])
def test_norm_fwd_fp8_delayed_scaling_output(shape, ln):
    N, H = shape
    x = make_input(N, H, "uniform", seed=N).cuda()
    gamma, beta = _params(H, False, 2)
    data = torch.empty(N, H, dtype=torch.uint8, device="cuda")
    scale = torch.tensor([3.7], device="cuda")
    scale_inv = torch.zeros(1, device="cuda")
    amax = torch.zeros(1, device="cuda")
    z = C.TensorHandle(C.DELAYED_TENSOR_SCALING)
    z.set(C.kRowwiseData, data, C.kFloat8E4M3).set(C.kRowwiseScaleInv, scale_inv)
    z.set(C.kScale, scale).set(C.kAmax, amax)
    mu, rs = ops.norm_fwd(ln, x, gamma, beta if ln else None, EPS, z)
    _check_stats(x, mu, rs, ln)
    yo = _y_oracle(x, gamma, beta, mu, rs, ln, False)
    q, _, am, sinv = O.fp8_quantize(yo, scale.item(), kind=0)
    assert np.array_equal(u8(data), q)
    assert amax.item() == am == np.abs(yo).max()
    assert scale_inv.item() == sinv


@pytest.mark.parametrize("shape", [(32, 64), (64, 256), (96, 768), (128, 4096), (64, 5120), (64, 8192),
                                   (32, 16384)])
@pytest.mark.parametrize("ln", [True, False], ids=["layernorm", "rmsnorm"])
@pytest.mark.parametrize("columnwise", [False, True], ids=["inference", "training"])
@pytest.mark.parametrize("swizzled", [False, True], ids=["compact", "swizzled"])
def test_norm_fwd_mxfp8_output(shape, ln, columnwise, swizzled):
    N, H = shape
    x = make_input(N, H, "randn", seed=H).cuda()
    gamma, beta = _params(H, True, 3)
    rs_shape, cs_shape = ops.mxfp8_scale_shapes(N, H)
    out = ops.MXFP8TensorStorage(
        rowwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
        rowwise_scale_inv=torch.zeros(rs_shape, dtype=torch.uint8, device="cuda"),
        columnwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda") if columnwise else None,
        columnwise_scale_inv=torch.zeros(cs_shape, dtype=torch.uint8, device="cuda") if columnwise else None,
        fp8_dtype=torch.float8_e4m3fn, shape=(N, H), with_gemm_swizzled_scales=swizzled)
    mu, rs = ops.norm_fwd(ln, x, gamma, beta if ln else None, EPS, out.handle(), True)
    _check_stats(x, mu, rs, ln)
    yo = _y_oracle(x, gamma, beta, mu, rs, ln, True)
    ref = O.mxfp8_quantize(yo, kind=0, rowwise=True, colwise=columnwise)
    assert np.array_equal(u8(out.rowwise_data), ref["rowwise"])
    pad_r = np.zeros(rs_shape, np.uint8)
    pad_r[:N, :H // 32] = ref["rowwise_scale"]
    got_r = u8(out.rowwise_scale_inv)
    assert np.array_equal(got_r.reshape(-1), O.swizzle_scales(pad_r)) if swizzled else np.array_equal(got_r, pad_r)
    if columnwise:
        assert np.array_equal(u8(out.columnwise_data), ref["colwise"])
        pad_c = np.zeros(cs_shape, np.uint8)
        pad_c[:N // 32, :H] = ref["colwise_scale"]
        got_c = u8(out.columnwise_scale_inv)
        if swizzled:
            assert np.array_equal(got_c.reshape(-1), O.swizzle_scales(pad_c, transposed_src=True))
        else:
            assert np.array_equal(got_c, pad_c)


def test_norm_fwd_fp16_input_and_fp32_gamma():
    N, H = 64, 512
    x = make_input(N, H, "randn", seed=5, dtype=torch.float16).cuda()
    gamma = torch.rand(H, device="cuda") + 0.5
    beta = torch.rand(H, device="cuda") - 0.5
    y = torch.empty_like(x)
    mu, rs = ops.norm_fwd(True, x, gamma, beta, EPS, C.plain(y))
    yo = O.norm_output(bits16(x), gamma.cpu().numpy(), beta.cpu().numpy(), mu.cpu().numpy(), rs.cpu().numpy(),
                       code=O.F16, wcode=O.F32)
    assert np.array_equal(bits16(y), bits16(torch.from_numpy(yo).to(torch.float16)))
    ref = torch.nn.functional.layer_norm(x.float(), (H,), gamma, beta, EPS)
    torch.testing.assert_close(y.float(), ref, rtol=2e-3, atol=2e-3)


def test_norm_fwd_is_deterministic_and_matches_oracle_at_full_size():
    """BASELINE config-4 activation shape; two launches give the same bytes; a slab against the oracle."""
    N, H = 8192, 4096
    x = make_input(N, H, "randn", seed=9).cuda()
    gamma, _ = _params(H, False, 4)
    outs = []
    for _ in range(2):
        rs_shape, cs_shape = ops.mxfp8_scale_shapes(N, H)
        out = ops.MXFP8TensorStorage(
            rowwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
            rowwise_scale_inv=torch.zeros(rs_shape, dtype=torch.uint8, device="cuda"),
            columnwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
            columnwise_scale_inv=torch.zeros(cs_shape, dtype=torch.uint8, device="cuda"),
            fp8_dtype=torch.float8_e4m3fn, shape=(N, H), with_gemm_swizzled_scales=True)
        _, rs = ops.norm_fwd(False, x, gamma, None, EPS, out.handle())
        outs.append((out, rs))
    (a, rs_a), (b, rs_b) = outs
    for n in ("rowwise_data", "rowwise_scale_inv", "columnwise_data", "columnwise_scale_inv"):
        assert torch.equal(getattr(a, n), getattr(b, n)), n
    assert torch.equal(rs_a, rs_b)
    sl = slice(4096, 4096 + 64)
    yo = O.norm_output(bits16(x[sl]), bits16(gamma), None, None, rs_a[sl].cpu().numpy())
    ref = O.mxfp8_quantize(yo, kind=0)
    assert np.array_equal(u8(a.rowwise_data[sl]), ref["rowwise"])
    assert np.array_equal(u8(a.columnwise_data[sl]), ref["colwise"])
    # the swizzled scales undo to the compact ones
    back = ops.swizzle_scales(a, inverse=True)
    assert np.array_equal(u8(back.rowwise_scale_inv[sl, :H // 32]), ref["rowwise_scale"])
    assert np.array_equal(u8(back.columnwise_scale_inv[4096 // 32:4096 // 32 + 2, :H]), ref["colwise_scale"])


@pytest.mark.parametrize("shape,ln", [((16384, 8192), True), ((16384, 8192), False), ((32768, 1024), True),
                                      ((8192, 12288), False), ((4096, 16384), True)])
def test_norm_fwd_pipeline_stress_repeatable_and_right_everywhere(shape, ln):
    """Many slabs per persistent cluster (the 3-stage pipeline wraps many times, cluster sizes 2 to
    16): 10 launches give identical bytes (a shared-memory race would not), and EVERY row's
    statistics and a stripe of rows from the start, middle and end match the oracle."""
    N, H = shape
    x = make_input(N, H, "randn", seed=3).cuda()
    gamma, beta = _params(H, False, 8)
    rs_shape, cs_shape = ops.mxfp8_scale_shapes(N, H)

    def run():
        out = ops.MXFP8TensorStorage(
            rowwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
            rowwise_scale_inv=torch.zeros(rs_shape, dtype=torch.uint8, device="cuda"),
            columnwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
            columnwise_scale_inv=torch.zeros(cs_shape, dtype=torch.uint8, device="cuda"),
            fp8_dtype=torch.float8_e4m3fn, shape=(N, H), with_gemm_swizzled_scales=False)
        mu, rs = ops.norm_fwd(ln, x, gamma, beta if ln else None, EPS, out.handle())
        return out, mu, rs

    first, mu, rs = run()
    for _ in range(9):
        again, mu2, rs2 = run()
        for n in ("rowwise_data", "rowwise_scale_inv", "columnwise_data", "columnwise_scale_inv"):
            assert torch.equal(getattr(first, n), getattr(again, n)), n
        assert torch.equal(rs, rs2) and (not ln or torch.equal(mu, mu2))
    x64 = x.double()
    m64 = x64.mean(1, keepdim=True) if ln else torch.zeros(N, 1, dtype=torch.float64, device="cuda")
    rs64 = torch.rsqrt(((x64 - m64) ** 2).mean(1) + EPS)
    torch.testing.assert_close(rs.double(), rs64, rtol=1e-6, atol=0)
    for r0 in (0, (N // 2) // 32 * 32, N - 64):
        sl = slice(r0, r0 + 64)
        yo = O.norm_output(bits16(x[sl]), bits16(gamma), bits16(beta) if ln else None,
                           mu[sl].cpu().numpy() if ln else None, rs[sl].cpu().numpy())
        ref = O.mxfp8_quantize(yo, kind=0)
        assert np.array_equal(u8(first.rowwise_data[sl]), ref["rowwise"])
        assert np.array_equal(u8(first.columnwise_data[sl]), ref["colwise"])
        assert np.array_equal(u8(first.rowwise_scale_inv[sl, :H // 32]), ref["rowwise_scale"])
        assert np.array_equal(u8(first.columnwise_scale_inv[r0 // 32:r0 // 32 + 2, :H]), ref["colwise_scale"])


@pytest.mark.parametrize("shape", [(32, 64), (100, 768), (256, 1024), (512, 4096), (300, 8192), (64, 16384)])
@pytest.mark.parametrize("mode", ["layernorm", "rmsnorm", "rmsnorm_add"])
@pytest.mark.parametrize("zc", [False, True], ids=["gamma", "zero_centered"])
def test_norm_bwd_matches_oracle(shape, mode, zc):
    N, H = shape
    ln = mode == "layernorm"
    x = make_input(N, H, "randn", seed=N).cuda()
    dz = make_input(N, H, "uniform", seed=N + 1).cuda()
    add = make_input(N, H, "randn", seed=N + 2).cuda() if mode == "rmsnorm_add" else None
    gamma, _ = _params(H, zc, 6)
    mu_o, rs_o = O.norm_stats(bits16(x), EPS, ln)
    mu = torch.from_numpy(mu_o).cuda() if ln else None
    rs = torch.from_numpy(rs_o).cuda()
    dx, dg, db = ops.norm_bwd(dz, x, mu, rs, gamma, zc, add=add)
    dx_o, dg_o, db_o = O.norm_backward(bits16(dz), bits16(x), mu_o, rs_o, bits16(gamma),
                                       add=bits16(add) if add is not None else None, zero_centered=zc)
    # bf16 outputs: half an ulp of rounding (2^-9 relative) plus fp32 summation-order noise
    np.testing.assert_allclose(dx.float().cpu().numpy(), dx_o, rtol=2.0 ** -8, atol=1e-3)
    scale = np.abs(dg_o).max()
    np.testing.assert_allclose(dg.float().cpu().numpy(), dg_o, rtol=2.0 ** -7, atol=2e-3 * scale)
    if ln:
        np.testing.assert_allclose(db.float().cpu().numpy(), db_o, rtol=2.0 ** -7, atol=2e-3 * np.abs(db_o).max())
    # deterministic
    dx2, dg2, db2 = ops.norm_bwd(dz, x, mu, rs, gamma, zc, add=add)
    assert torch.equal(dx, dx2) and torch.equal(dg, dg2) and (not ln or torch.equal(db, db2))


@pytest.mark.parametrize("ln", [True, False])
@pytest.mark.parametrize("shape,dtype", [((32, 96), torch.bfloat16), ((71, 229), torch.float32),
                                         ((64, 128), torch.float32), ((33, 2304), torch.float16)])
def test_norm_generic_path_for_shapes_and_dtypes_the_fused_kernel_does_not_take(ln, shape, dtype):
    """fp32 inputs, hidden sizes that are not a multiple of 64 (the reference's tests/cpp/operator/
    test_normalization.cu sweeps them): the generic kernel (csrc/norm_generic.cu), against float64."""
    N, H = shape
    g = torch.Generator().manual_seed(N + H)
    x = (torch.randn(N, H, generator=g) * 2 + 0.5).to(dtype).cuda()
    gamma = (torch.rand(H, generator=g) + 0.5).to(dtype).cuda()
    beta = torch.randn(H, generator=g).to(dtype).cuda() if ln else None
    z = torch.empty_like(x)
    mu, rs = ops.norm_fwd(ln, x, gamma, beta, EPS, C.plain(z))
    xd = x.double()
    m = xd.mean(1, keepdim=True) if ln else torch.zeros(N, 1, dtype=torch.float64, device="cuda")
    var = ((xd - m) ** 2).mean(1, keepdim=True)
    ref_rs = 1.0 / torch.sqrt(var + EPS)
    ref = (xd - m) * ref_rs * gamma.double() + (beta.double() if ln else 0.0)
    assert torch.allclose(rs.double(), ref_rs.squeeze(1), rtol=1e-5)
    if ln:
        assert torch.allclose(mu.double(), m.squeeze(1), rtol=1e-5, atol=1e-6)
    tol = 1e-5 if dtype == torch.float32 else 2.0 ** -7
    assert torch.allclose(z.double(), ref, rtol=tol, atol=tol)


def test_norm_generic_path_mxfp8_e5m2_output_equals_quantizer_bytes():
    """MXFP8 E5M2 output (generic path): the bytes are those of the MXFP8 quantizer applied to the fp32 y."""
    N, H = 64, 256
    x = make_input(N, H, "randn").cuda()
    gamma, beta = _params(H, False, 0)
    rs_shape, cs_shape = ops.mxfp8_scale_shapes(N, H)
    out = ops.MXFP8TensorStorage(
        rowwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
        rowwise_scale_inv=torch.zeros(rs_shape, dtype=torch.uint8, device="cuda"),
        columnwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
        columnwise_scale_inv=torch.zeros(cs_shape, dtype=torch.uint8, device="cuda"),
        fp8_dtype=torch.float8_e5m2, shape=(N, H), with_gemm_swizzled_scales=False)
    mu, rs = ops.norm_fwd(True, x, gamma, beta, EPS, out.handle())
    y = ((x.float() - mu[:, None]) * rs[:, None]) * gamma.float() + beta.float()
    ref = ops.mxfp8_quantize(y, torch.float8_e5m2)
    same = (out.rowwise_data == ref.rowwise_data).float().mean().item()
    assert same > 0.999, same    # y is recomputed here in another association order: a few boundary ties may differ
    assert torch.equal(out.rowwise_scale_inv, ref.rowwise_scale_inv) or same > 0.999


@pytest.mark.parametrize("ln", [True, False])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_norm_generic_path_mxfp8_small_32x32(ln, dtype):
    """32 x 32 (the smallest case of the reference's tests/cpp/operator/test_normalization_mxfp8.cu): generic path,
    padded [128, 4] / [4, 128] scales. Every row's dequantized output must match y (one E4M3 rounding: 2^-4)."""
    N = H = 32
    x = make_input(N, H, "randn", dtype=dtype).cuda()
    gamma, beta = _params(H, False, 0)
    gamma, beta = gamma.to(dtype), beta.to(dtype)
    rs_shape, cs_shape = ops.mxfp8_scale_shapes(N, H)
    out = ops.MXFP8TensorStorage(
        rowwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
        rowwise_scale_inv=torch.zeros(rs_shape, dtype=torch.uint8, device="cuda"),
        columnwise_data=torch.empty(N, H, dtype=torch.uint8, device="cuda"),
        columnwise_scale_inv=torch.zeros(cs_shape, dtype=torch.uint8, device="cuda"),
        fp8_dtype=torch.float8_e4m3fn, shape=(N, H), with_gemm_swizzled_scales=False)
    res = ops.norm_fwd(ln, x, gamma, beta if ln else None, EPS, out.handle())
    rs = res[-1]
    mu = res[0] if ln else torch.zeros_like(rs)
    y = ((x.float() - mu[:, None]) * rs[:, None]) * gamma.float() + (beta.float() if ln else 0.0)
    row = out.rowwise_data.view(torch.float8_e4m3fn).float() * torch.exp2(out.rowwise_scale_inv[:N, :1].float() - 127.0)
    col = out.columnwise_data.view(torch.float8_e4m3fn).float() * torch.exp2(out.columnwise_scale_inv[:1, :H].float() - 127.0)
    for name, d in (("rowwise", row), ("columnwise", col)):
        err = (d - y).abs().max().item()
        assert err <= 2.0 ** -4 * y.abs().max().item(), f"{name}: {err}"
        assert (d.abs().amax(dim=1) > 1e-3).all(), f"{name}: a whole row dequantizes to ~0"


@pytest.mark.parametrize("ln", [True, False])
@pytest.mark.parametrize("shape,dtype", [((96, 1000), torch.float32), ((64, 256), torch.float32),
                                         ((40, 1004), torch.bfloat16), ((8, 20480), torch.bfloat16)])
def test_norm_backward_generic_path(ln, shape, dtype):
    """Backward for what the fast kernel does not take (fp32 tensors, H % 8 != 0, H > 16384: the cases the reference's
    tests/cpp/operator/test_normalization.cu sweeps) against torch autograd in float64."""
    N, H = shape
    x = make_input(N, H, "randn", dtype=dtype).cuda()
    dz = (make_input(N, H, "randn", seed=3, dtype=dtype) * 0.1).cuda()
    g = torch.Generator().manual_seed(5)
    gamma = (torch.rand(H, generator=g) + 0.5).to(dtype).cuda()
    beta = torch.rand(H, generator=g).to(dtype).cuda()
    xd = x.double().requires_grad_(True)
    gd, bd = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    if ln:
        mu = xd.mean(1, keepdim=True)
        rs = torch.rsqrt(((xd - mu) ** 2).mean(1, keepdim=True) + EPS)
        y = (xd - mu) * rs * gd + bd
    else:
        mu = None
        rs = torch.rsqrt((xd ** 2).mean(1, keepdim=True) + EPS)
        y = xd * rs * gd
    y.backward(dz.double())
    dx, dgamma, dbeta = ops.norm_bwd(dz, x, mu.detach().float().reshape(N) if ln else None,
                                     rs.detach().float().reshape(N), gamma)
    tol = 2e-4 if dtype == torch.float32 else 3e-2
    scale = lambda t: t.abs().max().item()
    assert dx.dtype == dtype and (dx.double() - xd.grad).abs().max().item() <= tol * scale(xd.grad)
    assert (dgamma.double() - gd.grad).abs().max().item() <= tol * scale(gd.grad)
    if ln:
        assert (dbeta.double() - bd.grad).abs().max().item() <= tol * scale(bd.grad)


# --------------------------------------------------------------------------- #
# modules
# --------------------------------------------------------------------------- #
def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


@pytest.mark.parametrize("kind", ["LayerNorm", "RMSNorm"])
@pytest.mark.parametrize("zc", [False, True])
def test_norm_modules_match_torch(kind, zc):
    import transformerengine_b200 as te
    torch.manual_seed(0)
    H = 1024
    mod = getattr(te, kind)(H, eps=1e-5, zero_centered_gamma=zc, params_dtype=torch.bfloat16)
    with torch.no_grad():
        mod.weight.add_(torch.randn(H, device="cuda") * 0.1)
        if kind == "LayerNorm":
            mod.bias.add_(torch.randn(H, device="cuda") * 0.1)
    x = torch.randn(4, 96, H, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    dy = torch.randn_like(x)
    y = mod(x)
    y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    w = (mod.weight.detach().float() + (1.0 if zc else 0.0)).requires_grad_(True)
    if kind == "LayerNorm":
        b = mod.bias.detach().float().requires_grad_(True)
        yr = torch.nn.functional.layer_norm(xr, (H,), w, b, 1e-5)
    else:
        yr = xr * torch.rsqrt(xr.pow(2).mean(-1, keepdim=True) + 1e-5) * w
    yr.backward(dy.float())
    assert y.shape == x.shape and y.dtype == x.dtype
    assert _rel(y, yr) < 4e-3
    assert _rel(x.grad, xr.grad) < 6e-3
    assert _rel(mod.weight.grad, w.grad) < 6e-3
    if kind == "LayerNorm":
        assert _rel(mod.bias.grad, b.grad) < 6e-3


@pytest.mark.parametrize("name", ["none", "delayed", "current", "mxfp8", "nvfp4"])
@pytest.mark.parametrize("normalization", ["LayerNorm", "RMSNorm"])
def test_layernorm_linear_matches_fp32_reference(name, normalization):
    import transformerengine_b200 as te
    from transformerengine_b200 import recipe
    recipes = {"none": (None, 0.02), "delayed": (recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.HYBRID), 0.12),
               "current": (recipe.Float8CurrentScaling(), 0.12), "mxfp8": (recipe.MXFP8BlockScaling(), 0.08),
               "nvfp4": (recipe.NVFP4BlockScaling(disable_stochastic_rounding=True), 0.3)}
    rec, tol = recipes[name]
    torch.manual_seed(1)
    T, fin, fout = 512, 256, 384
    lin = te.LayerNormLinear(fin, fout, eps=1e-5, bias=True, normalization=normalization,
                             params_dtype=torch.bfloat16)
    with torch.no_grad():
        lin.bias.copy_(torch.randn(fout) * 0.1)
        lin.layer_norm_weight.add_(torch.randn(fin, device="cuda") * 0.1)
        if normalization == "LayerNorm":
            lin.layer_norm_bias.add_(torch.randn(fin, device="cuda") * 0.1)
    x = torch.randn(T, fin, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    dy = torch.randn(T, fout, device="cuda", dtype=torch.bfloat16) * 0.01
    steps = 3 if (rec is not None and rec.delayed()) else 1
    for _ in range(steps):
        lin.zero_grad()
        x.grad = None
        if rec is None:
            y = lin(x)
        else:
            with te.autocast(recipe=rec):
                y = lin(x)
        y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    g = lin.layer_norm_weight.detach().float().requires_grad_(True)
    wr = lin.weight.detach().float().requires_grad_(True)
    br = lin.bias.detach().float().requires_grad_(True)
    if normalization == "LayerNorm":
        b = lin.layer_norm_bias.detach().float().requires_grad_(True)
        n = torch.nn.functional.layer_norm(xr, (fin,), g, b, 1e-5)
    else:
        n = xr * torch.rsqrt(xr.pow(2).mean(-1, keepdim=True) + 1e-5) * g
    yr = torch.nn.functional.linear(n, wr, br)
    yr.backward(dy.float())
    assert _rel(y, yr) < tol
    assert _rel(x.grad, xr.grad) < tol
    assert _rel(lin.weight.grad, wr.grad) < tol
    assert _rel(lin.layer_norm_weight.grad, g.grad) < tol
    if normalization == "LayerNorm":
        assert _rel(lin.layer_norm_bias.grad, b.grad) < tol
    assert _rel(lin.bias.grad, br.grad) < 1e-2


def test_layernorm_linear_mxfp8_equals_hand_composed_operators():
    """The module's forward is exactly: fused norm -> MXFP8 (one kernel), weight cast, TN GEMM."""
    import transformerengine_b200 as te
    from transformerengine_b200 import recipe
    from transformerengine_b200.normalization import rmsnorm_fwd
    torch.manual_seed(2)
    T, fin, fout = 256, 512, 640
    lin = te.LayerNormLinear(fin, fout, bias=False, normalization="RMSNorm", params_dtype=torch.bfloat16)
    x = torch.randn(T, fin, device="cuda", dtype=torch.bfloat16)
    with te.autocast(recipe=recipe.MXFP8BlockScaling()), torch.no_grad():
        y = lin(x)
    q = te.MXFP8Quantizer(torch.float8_e4m3fn, rowwise=True, columnwise=False)
    q.optimize_for_gemm = True
    xq, _, _ = rmsnorm_fwd(x, lin.layer_norm_weight, 1e-5, quantizer=q)
    wq = ops.mxfp8_quantize(lin.weight.detach(), swizzled_scales=True)
    assert torch.equal(y, ops.general_gemm(wq, xq, "TN"))


def test_layernorm_linear_3d_input_and_row_count_not_multiple_of_32():
    """[batch, seq, hidden] input; 3 x 50 = 150 tokens: MXFP8 needs 32-row blocks for the columnwise
    copy, so the module falls back to norm -> bf16 -> quantize for the wgrad operand's sake."""
    import transformerengine_b200 as te
    from transformerengine_b200 import recipe
    torch.manual_seed(4)
    lin = te.LayerNormLinear(256, 128, normalization="RMSNorm", params_dtype=torch.bfloat16)
    x = torch.randn(3, 64, 256, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    with te.autocast(recipe=recipe.MXFP8BlockScaling()):
        y = lin(x)
    y.float().sum().backward()
    n = x.detach().float() * torch.rsqrt(x.detach().float().pow(2).mean(-1, keepdim=True) + 1e-5) * lin.layer_norm_weight.float()
    ref = torch.nn.functional.linear(n, lin.weight.float(), lin.bias.float())
    assert y.shape == (3, 64, 128) and _rel(y, ref) < 0.08
    assert x.grad.shape == x.shape and torch.isfinite(x.grad).all()


def test_layernorm_linear_return_layernorm_output():
    import transformerengine_b200 as te
    from transformerengine_b200 import recipe
    torch.manual_seed(3)
    lin = te.LayerNormLinear(256, 128, bias=False, return_layernorm_output=True, params_dtype=torch.bfloat16)
    x = torch.randn(64, 256, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    with te.autocast(recipe=recipe.MXFP8BlockScaling()):
        y, ln_out = lin(x)
    ref = torch.nn.functional.layer_norm(x.float(), (256,), lin.layer_norm_weight.float(), lin.layer_norm_bias.float())
    assert _rel(ln_out, ref) < 4e-3
    (y.float().sum() + ln_out.float().sum()).backward()
    assert torch.isfinite(x.grad).all() and lin.layer_norm_weight.grad is not None
