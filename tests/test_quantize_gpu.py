"""GPU parity tests: CUDA kernels (through the C ABI) vs the CPU oracle.

Bit-exact bar for every quantized byte and scale factor (tolerance 0), like the
reference's exactness tests (tests/cpp/operator/test_cast_mxfp8.cu:309-320,
tests/pytorch/test_float8_current_scaling_exact.py:786-809). Shapes follow the
reference's C++ test sizes (test_cast_mxfp8.cu:524-533) plus ragged ones.
"""
import numpy as np
import pytest
import torch

from tests.util import bits16, make_input, u8

pytestmark = pytest.mark.gpu

ops = None
O = None


@pytest.fixture(scope="module", autouse=True)
def _mods():
    global ops, O
    from oracle import oracle as _o
    from transformerengine_b200 import ops as _ops
    ops, O = _ops, _o
    assert torch.cuda.is_available()


def _np_in(x):
    if x.dtype == torch.float32:
        return x.cpu().numpy(), 0
    return bits16(x), (1 if x.dtype == torch.bfloat16 else 2)


MX_SHAPES = [(1, 32), (16, 48), (64, 96), (128, 128), (256, 272), (304, 304), (993, 512),
             (1024, 1024), (8, 32 * 1024)]


@pytest.mark.parametrize("shape", MX_SHAPES)
@pytest.mark.parametrize("kind", ["randn", "uniform", "outliers"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
def test_mxfp8_quantize_bit_exact(shape, kind, dtype):
    if dtype == torch.float16 and kind == "outliers":
        pytest.skip("outlier magnitudes overflow fp16")
    rows, cols = shape
    x = make_input(rows, cols, kind, dtype=dtype).cuda()
    q = ops.mxfp8_quantize(x)
    xin, code = _np_in(x)
    ref = O.mxfp8_quantize(xin, code=code)
    kb, rb = (cols + 31) // 32, (rows + 31) // 32
    assert np.array_equal(u8(q.rowwise_scale_inv)[:rows, :kb], ref["rowwise_scale"])
    assert np.array_equal(u8(q.columnwise_scale_inv)[:rb, :cols], ref["colwise_scale"])
    assert np.array_equal(u8(q.rowwise_data), ref["rowwise"])
    assert np.array_equal(u8(q.columnwise_data), ref["colwise"])


@pytest.mark.parametrize("rowwise,columnwise", [(True, False), (False, True)])
@pytest.mark.parametrize("fp8", [torch.float8_e4m3fn, torch.float8_e5m2])
def test_mxfp8_single_direction_and_e5m2(rowwise, columnwise, fp8):
    x = make_input(320, 384, "uniform").cuda()
    q = ops.mxfp8_quantize(x, fp8_dtype=fp8, rowwise=rowwise, columnwise=columnwise)
    ref = O.mxfp8_quantize(bits16(x), kind=0 if fp8 == torch.float8_e4m3fn else 1,
                           rowwise=rowwise, colwise=columnwise)
    if rowwise:
        assert np.array_equal(u8(q.rowwise_data), ref["rowwise"])
        assert np.array_equal(u8(q.rowwise_scale_inv)[:320, :12], ref["rowwise_scale"])
    if columnwise:
        assert np.array_equal(u8(q.columnwise_data), ref["colwise"])
        assert np.array_equal(u8(q.columnwise_scale_inv)[:10, :384], ref["colwise_scale"])


@pytest.mark.parametrize("shape", [(128, 128), (256, 1024), (993, 512), (300, 160)])
def test_mxfp8_fused_swizzle_equals_compact_then_swizzle(shape):
    """reference tests/pytorch/mxfp8/test_mxfp8_quantize_swizzle_fusion.py:72-89"""
    x = make_input(*shape, "randn").cuda()
    fused = ops.mxfp8_quantize(x, swizzled_scales=True)
    two_step = ops.swizzle_scales(ops.mxfp8_quantize(x))
    assert torch.equal(fused.rowwise_data, two_step.rowwise_data)
    assert torch.equal(fused.columnwise_data, two_step.columnwise_data)
    assert torch.equal(fused.rowwise_scale_inv.flatten(), two_step.rowwise_scale_inv.flatten())
    assert torch.equal(fused.columnwise_scale_inv.flatten(), two_step.columnwise_scale_inv.flatten())


@pytest.mark.parametrize("shape", [(128, 128), (384, 1024), (993, 512)])
def test_swizzle_matches_oracle_and_roundtrips(shape):
    """reference tests/cpp/operator/test_swizzle.cu:30-86"""
    x = make_input(*shape, "uniform").cuda()
    q = ops.mxfp8_quantize(x)
    sw = ops.swizzle_scales(q)
    assert np.array_equal(u8(sw.rowwise_scale_inv).ravel(), O.swizzle_scales(u8(q.rowwise_scale_inv)))
    assert np.array_equal(u8(sw.columnwise_scale_inv).ravel(),
                          O.swizzle_scales(u8(q.columnwise_scale_inv), transposed_src=True))
    back = ops.swizzle_scales(sw, inverse=True)
    assert torch.equal(back.rowwise_scale_inv, q.rowwise_scale_inv)
    assert torch.equal(back.columnwise_scale_inv, q.columnwise_scale_inv)


def test_mxfp8_noop_flag_skips_the_cast():
    x = make_input(128, 256, "randn").cuda()
    q = ops.mxfp8_quantize(x)
    before = q.rowwise_data.clone()
    y = make_input(128, 256, "uniform", seed=7).cuda()
    noop = torch.ones(1, dtype=torch.float32, device="cuda")
    ops.mxfp8_quantize(y, noop=noop, out=q)
    assert torch.equal(q.rowwise_data, before)
    noop.zero_()
    ops.mxfp8_quantize(y, noop=noop, out=q)
    assert not torch.equal(q.rowwise_data, before)


@pytest.mark.parametrize("shape", [(256, 256), (993, 512)])
@pytest.mark.parametrize("columnwise", [False, True])
def test_mxfp8_dequantize_matches_oracle(shape, columnwise):
    x = make_input(*shape, "randn").cuda()
    q = ops.mxfp8_quantize(x)
    got = ops.dequantize(q, torch.float32, columnwise=columnwise).cpu().numpy()
    rows, cols = shape
    if columnwise:
        ref = O.mxfp8_dequantize(u8(q.columnwise_data), u8(q.columnwise_scale_inv), colwise=True)
    else:
        ref = O.mxfp8_dequantize(u8(q.rowwise_data), u8(q.rowwise_scale_inv), colwise=False)
    assert np.array_equal(got, ref)


def test_mxfp8_full_size_properties():
    """BASELINE config 1 size (4096x4096): size-independent properties."""
    x = make_input(4096, 4096, "randn").cuda()
    q = ops.mxfp8_quantize(x)
    # (1) linearity in powers of two: scaling the input by 2^k shifts every E8M0 by k
    q4 = ops.mxfp8_quantize(x * 4.0)
    assert torch.equal(q4.rowwise_data, q.rowwise_data)
    assert torch.equal(q4.columnwise_data, q.columnwise_data)
    assert torch.equal(q4.rowwise_scale_inv[:, :128], q.rowwise_scale_inv[:, :128] + 2)
    # (2) round trip: |dequant(quant(x)) - x| <= |x|/16 (3 mantissa bits) + block_amax * 2^-17
    d = ops.dequantize(q, torch.float32)
    xf = x.float()
    bmax = xf.view(4096, 128, 32).abs().amax(-1, keepdim=True).expand(-1, -1, 32).reshape(4096, 4096)
    assert bool(((d - xf).abs() <= xf.abs() / 16 + bmax * 2.0 ** -17).all())
    # (3) the transpose symmetry: colwise(x) == rowwise(x^T)^T
    qt = ops.mxfp8_quantize(x.t().contiguous(), columnwise=False)
    assert torch.equal(qt.rowwise_data.t().contiguous(), q.columnwise_data)
    assert torch.equal(qt.rowwise_scale_inv[:, :128].t().contiguous(), q.columnwise_scale_inv[:128])
    # (4) oracle on a 256-row slab of the same tensor
    ref = O.mxfp8_quantize(bits16(x[1024:1280]))
    assert np.array_equal(u8(q.rowwise_data[1024:1280]), ref["rowwise"])
    assert np.array_equal(u8(q.columnwise_data[1024:1280]), ref["colwise"])


# ---- per-tensor FP8 -----------------------------------------------------------
@pytest.mark.parametrize("n", [(2048, 768), (3072, 768), (1, 7), (333, 129), (65536 + 8,)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("fp8", [torch.float8_e4m3fn, torch.float8_e5m2])
def test_fp8_delayed_cast_bit_exact(n, dtype, fp8):
    """reference tests/cpp/operator/test_cast_transpose.cu:24-40"""
    shape = n if len(n) == 2 else (1, n[0])
    x = make_input(*shape, "uniform", dtype=dtype).cuda()
    scale = torch.tensor([2.5], device="cuda")
    amax = torch.zeros(1, device="cuda")
    q = ops.fp8_quantize(x, scale, fp8, amax=amax)
    xin, code = _np_in(x)
    kind = 0 if fp8 == torch.float8_e4m3fn else 1
    ref_q, _, ref_amax, ref_sinv = O.fp8_quantize(xin, 2.5, kind=kind, code=code)
    assert np.array_equal(u8(q.data).reshape(ref_q.shape), ref_q)
    assert amax.item() == ref_amax
    assert q.scale_inv.item() == ref_sinv


def test_fp8_cast_with_transpose_buffer():
    x = make_input(200, 136, "randn", dtype=torch.bfloat16).cuda()
    scale = torch.tensor([7.0], device="cuda")
    q = ops.fp8_quantize(x, scale, with_transpose=True)
    ref_q, ref_t, _, _ = O.fp8_quantize(bits16(x), 7.0, transpose=True)
    assert np.array_equal(u8(q.data), ref_q)
    assert np.array_equal(u8(q.transpose), ref_t)


@pytest.mark.parametrize("shape", [(2048, 768), (17, 33), (8192, 1024)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("pow2", [False, True])
def test_fp8_current_scaling_amax_scale_cast(shape, dtype, pow2):
    """reference tests/cpp/operator/test_cast_current_scaling.cu and
    tests/pytorch/test_float8_current_scaling_exact.py:786-809"""
    x = (make_input(*shape, "randn", dtype=dtype) * 3.7).to(dtype).cuda()
    amax = ops.compute_amax(x)
    xin, code = _np_in(x)
    ref_amax = O.amax(xin, code)
    assert amax.item() == ref_amax
    scale = torch.empty(1, device="cuda")
    ops.compute_scale_from_amax(amax, scale, torch.float8_e4m3fn, force_pow_2_scales=pow2)
    ref_scale = O.scale_from_amax(ref_amax, 448.0, pow2, 0.0)
    assert scale.item() == ref_scale
    q = ops.fp8_quantize(x, scale)
    ref_q, _, _, ref_sinv = O.fp8_quantize(xin, ref_scale, code=code)
    assert np.array_equal(u8(q.data), ref_q)
    assert q.scale_inv.item() == ref_sinv


def test_scale_from_amax_edge_cases():
    for a, eps in [(0.0, 0.0), (float("inf"), 0.0), (1e-45, 0.0), (0.0, 1e-3), (3e38, 0.0)]:
        amax = torch.tensor([a], device="cuda")
        scale = torch.empty(1, device="cuda")
        ops.compute_scale_from_amax(amax, scale, torch.float8_e4m3fn, amax_epsilon=eps)
        assert scale.item() == O.scale_from_amax(a, 448.0, False, eps), (a, eps)


@pytest.mark.parametrize("algo", ["max", "most_recent"])
@pytest.mark.parametrize("hist_len,n", [(1, 3), (3, 3), (1024, 2), (40, 70)])
def test_delayed_scaling_update(algo, hist_len, n):
    """reference tests/pytorch/test_recipe.py:110-180 (closed-form expectations)"""
    g = torch.Generator().manual_seed(3)
    hist = torch.rand(hist_len, n, generator=g) * 10
    hist[0, 0] = 0.0  # a zero most-recent amax keeps the old scale under "most_recent"
    scale = torch.rand(n, generator=g) + 0.5
    h_ref, s_ref = O.delayed_scaling_update(hist.numpy(), scale.numpy(), algo=algo, margin=1.0,
                                            bulk=False)
    hd, sd = hist.cuda(), scale.cuda()
    ops.delayed_scaling_update(hd, sd, algo, margin=1.0)
    assert np.array_equal(hd.cpu().numpy(), h_ref)
    assert np.array_equal(sd.cpu().numpy(), s_ref)


@pytest.mark.parametrize("algo", ["max", "most_recent"])
def test_delayed_scaling_bulk_update_after_reduction(algo):
    g = torch.Generator().manual_seed(5)
    hists = [torch.rand(5, k, generator=g) for k in (3, 2, 3)]
    scales = [torch.rand(k, generator=g) + 0.5 for k in (3, 2, 3)]
    red = torch.rand(8, generator=g) * 4
    red[1] = 0.0
    hists[0][0, 1] = 0.0  # zero in both places: no roll, scale kept
    refs, off = [], 0
    for h, s in zip(hists, scales):
        k = h.shape[1]
        refs.append(O.delayed_scaling_update(h.numpy(), s.numpy(), red[off:off + k].numpy(),
                                             algo=algo, bulk=True))
        off += k
    hd = [h.cuda() for h in hists]
    sd = [s.cuda() for s in scales]
    ops.fused_amax_and_scale_update_after_reduction(red.cuda(), hd, sd, algo,
                                                    torch.float8_e4m3fn, 0.0)
    for (h_ref, s_ref), h, s in zip(refs, hd, sd):
        assert np.array_equal(h.cpu().numpy(), h_ref)
        assert np.array_equal(s.cpu().numpy(), s_ref)


def test_errors_are_raised_not_swallowed():
    x = make_input(64, 64, "randn").cuda()
    q = ops.mxfp8_quantize(x)
    bad = ops.MXFP8TensorStorage(q.rowwise_data, None, None, None, torch.float8_e4m3fn, (64, 64))
    bad.rowwise_scale_inv = torch.empty(0, dtype=torch.uint8, device="cuda")
    with pytest.raises(RuntimeError, match="Scaling tensor must be allocated"):
        ops.mxfp8_quantize(x, columnwise=False, out=bad)


@pytest.mark.parametrize("swizzled", [True, False])
@pytest.mark.parametrize("shape,slab_bytes", [((1024, 512), 128 * 512 * 2), ((640, 256), 256 * 256 * 2),
                                              ((384, 1024), 16 << 20)])
def test_mxfp8_host_pipeline_equals_one_shot_and_oracle(shape, slab_bytes, swizzled):
    """ops.MXFP8HostQuantizer (row slabs, 3 streams, pinned host in/out) gives the bytes of the
    one-shot device call and of the oracle; two back-to-back passes exercise the cross-pass events."""
    rows, cols = shape
    pipe = ops.MXFP8HostQuantizer(rows, cols, swizzled_scales=swizzled, slab_bytes=slab_bytes)
    assert len(pipe.slabs) == max(1, -(-rows // max(128, slab_bytes // (cols * 2) // 128 * 128)))
    rs, cs = ops.mxfp8_scale_shapes(rows, cols)
    h_row, h_col = (torch.empty(rows, cols, dtype=torch.uint8).pin_memory() for _ in range(2))
    h_rs = torch.empty(rs, dtype=torch.uint8).pin_memory()
    h_cs = torch.empty(cs, dtype=torch.uint8).pin_memory()
    for seed in (1, 2):
        x = make_input(rows, cols, "outliers" if seed == 2 else "randn", seed=seed).pin_memory()
        pipe(x, h_row, h_col, h_rs, h_cs)
    pipe.join()
    torch.cuda.synchronize()
    one = ops.mxfp8_quantize(x.cuda(), swizzled_scales=swizzled)
    assert torch.equal(h_row, one.rowwise_data.cpu()) and torch.equal(h_col, one.columnwise_data.cpu())
    assert torch.equal(h_rs, one.rowwise_scale_inv.cpu()) and torch.equal(h_cs, one.columnwise_scale_inv.cpu())
    ref = O.mxfp8_quantize(bits16(x))
    assert np.array_equal(h_row.numpy(), ref["rowwise"]) and np.array_equal(h_col.numpy(), ref["colwise"])


@pytest.mark.parametrize("shape", [(256, 384), (1024, 512), (96, 160), (33, 64)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("kind", ["mxfp8", "mxfp8_swz", "fp8"])
def test_quantize_dbias_matches_cast_and_column_sum(shape, dtype, kind):
    """nvte_quantize_dbias (reference tests/cpp/operator/test_cast_dbias.cu, test_cast_mxfp8.cu with
    IS_DBIAS): the cast bytes are those of nvte_quantize, dbias = column sum accumulated in fp32 and
    rounded once to the input dtype (tolerance: one rounding of the output dtype + fp32 sum order)."""
    rows, cols = shape
    x = make_input(rows, cols, "randn", seed=5, dtype=dtype).cuda()
    if kind == "fp8":
        scale = torch.full((1,), 3.0, device="cuda")
        amax = torch.zeros(1, device="cuda")
        q, db = ops.fp8_quantize(x, scale, amax=amax, with_dbias=True)
        ref = ops.fp8_quantize(x, scale)
        assert torch.equal(q.data, ref.data) and q.scale_inv.item() == ref.scale_inv.item()
        assert amax.item() == x.abs().max().item()
    else:
        swz = kind.endswith("swz")
        q, db = ops.mxfp8_quantize(x, swizzled_scales=swz, with_dbias=True)
        ref = ops.mxfp8_quantize(x, swizzled_scales=swz)
        assert torch.equal(q.rowwise_data, ref.rowwise_data) and torch.equal(q.columnwise_data, ref.columnwise_data)
        assert torch.equal(q.rowwise_scale_inv, ref.rowwise_scale_inv)
        assert torch.equal(q.columnwise_scale_inv, ref.columnwise_scale_inv)
    want = x.double().sum(0)
    tol = (2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -20) * (want.abs().max().item() + rows ** 0.5)
    assert db.dtype == dtype and (db.double() - want).abs().max().item() <= tol


@pytest.mark.parametrize("fp8", [torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("rowwise,columnwise", [(True, True), (True, False), (False, True)])
def test_mxfp8_tma_kernel_variants(fp8, dtype, rowwise, columnwise):
    """128-aligned 16-bit inputs take the TMA kernel (quantize_mxfp8_tma.cu): every instantiation
    (E4M3 / E5M2 output, bf16 / fp16 input, either or both directions, compact and swizzled scales)
    against the oracle; 640 x 768 = 30 tiles, 2560 x 2048 = 320 tiles (> 296 CTAs: the ring wraps)."""
    code = 1 if dtype == torch.bfloat16 else 2
    for rows, cols, kind in ((640, 768, "uniform"), (2560, 2048, "randn")):
        x = make_input(rows, cols, kind, seed=9, dtype=dtype).cuda()
        ref = O.mxfp8_quantize(bits16(x), kind=0 if fp8 == torch.float8_e4m3fn else 1, code=code,
                               rowwise=rowwise, colwise=columnwise)
        for swz in (False, True):
            q = ops.mxfp8_quantize(x, fp8_dtype=fp8, rowwise=rowwise, columnwise=columnwise, swizzled_scales=swz)
            if swz:
                q = ops.swizzle_scales(q, inverse=True)
            if rowwise:
                assert np.array_equal(u8(q.rowwise_data), ref["rowwise"])
                assert np.array_equal(u8(q.rowwise_scale_inv)[:rows, :cols // 32], ref["rowwise_scale"])
            if columnwise:
                assert np.array_equal(u8(q.columnwise_data), ref["colwise"])
                assert np.array_equal(u8(q.columnwise_scale_inv)[:rows // 32, :cols], ref["colwise_scale"])


def test_mxfp8_tma_kernel_noop_flag():
    x = make_input(256, 384, "randn").cuda()
    q = ops.mxfp8_quantize(x, swizzled_scales=True)
    before = (q.rowwise_data.clone(), q.columnwise_scale_inv.clone())
    noop = torch.ones(1, dtype=torch.float32, device="cuda")
    ops.mxfp8_quantize(x * 3.0, swizzled_scales=True, noop=noop, out=q)
    assert torch.equal(q.rowwise_data, before[0]) and torch.equal(q.columnwise_scale_inv, before[1])
    noop.zero_()
    ops.mxfp8_quantize(x * 3.0, swizzled_scales=True, noop=noop, out=q)
    assert not torch.equal(q.rowwise_data, before[0])


def test_multi_tensor_quantize_equals_individual_casts():
    """`nvte_multi_tensor_quantize` (MoE-style list of tensors) = the same bytes as one cast each."""
    shapes = [(128, 256), (384, 256), (64, 256), (1024, 256)]
    xs = [make_input(r, c, "randn", seed=r).cuda() for r, c in shapes]
    ref = [ops.mxfp8_quantize(x, swizzled_scales=True) for x in xs]
    outs = [ops.mxfp8_quantize(torch.zeros_like(x), swizzled_scales=True) for x in xs]
    ops.multi_tensor_quantize(xs, outs)
    for a, b in zip(outs, ref):
        for n in ("rowwise_data", "rowwise_scale_inv", "columnwise_data", "columnwise_scale_inv"):
            assert torch.equal(getattr(a, n), getattr(b, n)), n


@pytest.mark.parametrize("fp8_dtype", [torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("npeers", [1, 3, 7])
def test_fp8_cast_with_peer_destinations_writes_the_same_bytes_everywhere(fp8_dtype, npeers):
    """kNVTEQuantizationConfigB200PeerRowwiseData: the cast that is its own all-gather. On one GPU the
    'peers' are local buffers; every destination must hold exactly the bytes of the plain cast."""
    from transformerengine_b200 import cabi as C
    from transformerengine_b200 import ops
    torch.manual_seed(3)
    x = (torch.randn(384, 1024, device="cuda") * 3).bfloat16()
    scale = torch.full((1,), 7.5, device="cuda")
    amax_ref, amax = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
    ref = ops.fp8_quantize(x, scale, fp8_dtype, amax=amax_ref)
    peers = [torch.zeros(384 * 1024 + 64, dtype=torch.uint8, device="cuda") for _ in range(npeers)]
    cfg = C.QuantConfig(peer_rowwise=[p.data_ptr() + 32 for p in peers])
    got = ops.fp8_quantize(x, scale, fp8_dtype, amax=amax, peer_config=cfg)
    torch.cuda.synchronize()
    assert torch.equal(got.data, ref.data) and torch.equal(got.scale_inv, ref.scale_inv)
    assert amax.item() == amax_ref.item() == x.abs().max().item()
    for p in peers:
        assert torch.equal(p[32:32 + x.numel()].view(384, 1024), ref.data)
        assert not p[:32].any() and not p[32 + x.numel():].any()   # nothing outside the destination


def test_fp8_cast_with_peer_destinations_rejects_what_it_cannot_do():
    from transformerengine_b200 import cabi as C
    from transformerengine_b200 import ops
    scale = torch.ones(1, device="cuda")
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    cfg = C.QuantConfig(peer_rowwise=[buf.data_ptr()])
    with pytest.raises(RuntimeError, match="bf16 / fp16"):
        ops.fp8_quantize(torch.randn(16, 64, device="cuda"), scale, peer_config=cfg)
    with pytest.raises(RuntimeError, match="16-byte aligned"):
        ops.fp8_quantize(torch.randn(16, 64, device="cuda").bfloat16(), scale,
                         peer_config=C.QuantConfig(peer_rowwise=[buf.data_ptr() + 8]))


@pytest.mark.parametrize("shape", [(200, 96), (96, 640), (993, 512)])
def test_mxfp8_noop_flag_keeps_swizzled_scales_of_ragged_shapes(shape):
    """for ragged shapes the swizzled scale buffers are zero-filled before the cast; with the noop
    flag set (cached weight under CUDA graphs) that fill must be skipped too, or the cached operand keeps its
    FP8 bytes but loses every scale."""
    x = make_input(*shape, "randn").cuda()
    q = ops.mxfp8_quantize(x, swizzled_scales=True)
    before = [t.clone() for t in (q.rowwise_data, q.columnwise_data, q.rowwise_scale_inv, q.columnwise_scale_inv)]
    assert before[2].any() and before[3].any()
    noop = torch.ones(1, dtype=torch.float32, device="cuda")
    ops.mxfp8_quantize(x * 5.0, swizzled_scales=True, noop=noop, out=q)
    for a, b in zip((q.rowwise_data, q.columnwise_data, q.rowwise_scale_inv, q.columnwise_scale_inv), before):
        assert torch.equal(a, b)
    noop.zero_()
    ops.mxfp8_quantize(x * 5.0, swizzled_scales=True, noop=noop, out=q)
    fresh = ops.mxfp8_quantize(x * 5.0, swizzled_scales=True)
    assert torch.equal(q.rowwise_scale_inv, fresh.rowwise_scale_inv)
    assert torch.equal(q.columnwise_scale_inv, fresh.columnwise_scale_inv)
    assert torch.equal(q.rowwise_data, fresh.rowwise_data)
    # the quantizer's weight-cache path forwards the flag
    import transformerengine_b200 as te
    qz = te.MXFP8Quantizer(torch.float8_e4m3fn)
    qz.optimize_for_gemm = True
    noop.fill_(1.0)
    kept = q.rowwise_scale_inv.clone()
    qz.update_quantized(x * 0.125, q, noop_flag=noop)
    assert torch.equal(q.rowwise_scale_inv, kept)
