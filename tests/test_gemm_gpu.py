"""GPU tests of the tcgen05 GEMM (`nvte_cublas_gemm_v2`) against the oracle's
dequantise-then-fp64-matmul reference.

The reference delegates this contraction to cuBLASLt (closed source), so parity
is tolerance-based: bf16 outputs within atol = rtol = 8e-3 of
the fp64 reference (the reference's own NVFP4 bound,
tests/pytorch/nvfp4/test_nvfp4_gemm_exact.py:239); fp32 outputs, which expose the
fp32 TMEM accumulation directly, within 1e-5 relative.
"""
import numpy as np
import pytest
import torch

from tests.util import u8

pytestmark = pytest.mark.gpu

ops = None
O = None


@pytest.fixture(scope="module", autouse=True)
def _mods():
    global ops, O
    from oracle import oracle as _o
    from transformerengine_b200 import ops as _ops
    ops, O = _ops, _o


def _shapes(n, m, k, layout):
    # TE convention: TN: A[m,k], B[n,k]; NN: A[k,m], B[n,k]; NT: A[k,m], B[k,n]
    return ((m, k) if layout[0] == "T" else (k, m)), ((k, n) if layout[1] == "T" else (n, k))


def _operands(n, m, k, layout, seed=0):
    g = torch.Generator().manual_seed(seed)
    sa, sb = _shapes(n, m, k, layout)
    a = (torch.randn(*sa, generator=g) * 0.5).bfloat16().cuda()
    b = torch.randn(*sb, generator=g).bfloat16().cuda()
    return a, b


def _to_mk_nk(A, B, layout):
    A2 = A if layout[0] == "T" else np.ascontiguousarray(A.T)
    B2 = np.ascontiguousarray(B.T) if layout[1] == "T" else B
    return A2, B2


def _assert_close(got, ref, out_dtype):
    if out_dtype == torch.float32:
        assert np.abs(got - ref).max() <= 1e-5 * np.abs(ref).max()
    else:
        np.testing.assert_allclose(got, ref, rtol=8e-3, atol=8e-3)


MX_CASES = [(128, 128, 128), (256, 384, 1024), (200, 136, 96), (128, 256, 256), (512, 1024, 512),
            (384, 128, 2048), (96, 640, 160)]


@pytest.mark.parametrize("layout", ["TN", "NN", "NT"])
@pytest.mark.parametrize("n,m,k", MX_CASES)
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_mxfp8_gemm(layout, n, m, k, out_dtype):
    if layout != "TN" and (m % 16 or n % 16):
        pytest.skip("MN-major operands need 16-byte rows")
    a, b = _operands(n, m, k, layout)
    qa, qb = ops.mxfp8_quantize(a), ops.mxfp8_quantize(b)
    out = ops.general_gemm(ops.swizzle_scales(qa), ops.swizzle_scales(qb), layout, out_dtype)
    ca, cb = layout[0] != "T", layout[1] == "T"
    A = O.mxfp8_dequantize(u8(qa.columnwise_data if ca else qa.rowwise_data),
                           u8(qa.columnwise_scale_inv if ca else qa.rowwise_scale_inv), colwise=ca)
    B = O.mxfp8_dequantize(u8(qb.columnwise_data if cb else qb.rowwise_data),
                           u8(qb.columnwise_scale_inv if cb else qb.rowwise_scale_inv), colwise=cb)
    ref = O.gemm_tn_ref(*_to_mk_nk(A, B, layout))
    _assert_close(out.float().cpu().numpy(), ref, out_dtype)


def test_mxfp8_gemm_fused_swizzle_operands_and_e5m2_grad():
    """Operands quantized straight into the swizzled layout; E5M2 gradient operand (dgrad)."""
    n, m, k = 256, 512, 384
    w = (torch.randn(k, m) * 0.1).bfloat16().cuda()          # W [out=k, in=m]
    dy = (torch.randn(n, k) * 1e-2).bfloat16().cuda()        # dy [tokens, out]
    qw = ops.mxfp8_quantize(w, swizzled_scales=True)
    qdy = ops.mxfp8_quantize(dy, fp8_dtype=torch.float8_e5m2, swizzled_scales=True)
    out = ops.general_gemm(qw, qdy, "NN", torch.float32)
    W = O.mxfp8_dequantize(u8(ops.swizzle_scales(qw, inverse=True).columnwise_data),
                           u8(ops.swizzle_scales(qw, inverse=True).columnwise_scale_inv), colwise=True)
    qd = ops.swizzle_scales(qdy, inverse=True)
    DY = O.mxfp8_dequantize(u8(qd.rowwise_data), u8(qd.rowwise_scale_inv), colwise=False, kind=1)
    ref = O.gemm_tn_ref(np.ascontiguousarray(W.T), DY)
    _assert_close(out.cpu().numpy(), ref, torch.float32)


@pytest.mark.parametrize("layout", ["TN", "NN", "NT"])
@pytest.mark.parametrize("n,m,k", [(2048, 3072, 768), (128, 128, 128), (208, 144, 96)])
def test_fp8_tensor_scaled_gemm(layout, n, m, k):
    """Per-tensor FP8 (delayed / current scaling): config 2 shapes. Non-TN layouts use
    MN-major tensor-core operands, no transposed copies (reference
    transformer_engine.cpp:1131-1145, cublaslt_gemm.cu:129-157)."""
    a, b = _operands(n, m, k, layout, seed=1)
    sa, sb = torch.tensor([13.0], device="cuda"), torch.tensor([3.5], device="cuda")
    qa = ops.fp8_quantize(a, sa)
    qb = ops.fp8_quantize(b, sb, torch.float8_e5m2 if layout != "TN" else torch.float8_e4m3fn)
    out = ops.general_gemm(qa, qb, layout, torch.float32)
    A = torch.from_numpy(u8(qa.data)).view(torch.float8_e4m3fn).float().numpy() * qa.scale_inv.item()
    B = torch.from_numpy(u8(qb.data)).view(qb.fp8_dtype).float().numpy() * qb.scale_inv.item()
    ref = O.gemm_tn_ref(*_to_mk_nk(A, B, layout))
    _assert_close(out.cpu().numpy(), ref, torch.float32)


def test_gemm_bias_accumulate_alpha_beta():
    n, m, k = 256, 384, 256
    a, b = _operands(n, m, k, "TN", seed=2)
    qa = ops.mxfp8_quantize(a, swizzled_scales=True, columnwise=False)
    qb = ops.mxfp8_quantize(b, swizzled_scales=True, columnwise=False)
    bias = torch.randn(m).bfloat16().cuda()
    base = ops.general_gemm(qa, qb, "TN", torch.float32)
    with_bias = ops.general_gemm(qa, qb, "TN", torch.bfloat16, bias=bias)
    np.testing.assert_allclose(with_bias.float().cpu().numpy(),
                               (base + bias.float()).cpu().numpy(), rtol=8e-3, atol=8e-3)
    acc = torch.full((n, m), 2.0, dtype=torch.float32, device="cuda")
    ops.general_gemm(qa, qb, "TN", out=acc, accumulate=True)
    assert torch.allclose(acc, base + 2.0, rtol=1e-6, atol=1e-5)
    scaled = torch.full((n, m), 4.0, dtype=torch.float32, device="cuda")
    ops.general_gemm(qa, qb, "TN", out=scaled, alpha=0.5, beta=0.25)
    assert torch.allclose(scaled, 0.5 * base + 1.0, rtol=1e-6, atol=1e-5)


def test_gemm_rejects_compact_scales_and_mismatched_k():
    a, b = _operands(128, 128, 128, "TN")
    qa, qb = ops.mxfp8_quantize(a), ops.mxfp8_quantize(b)
    with pytest.raises(RuntimeError, match="GEMM-swizzled"):
        ops.general_gemm(qa, qb, "TN")
    qc = ops.mxfp8_quantize(torch.randn(128, 256).bfloat16().cuda(), swizzled_scales=True)
    with pytest.raises(RuntimeError, match="inner dimensions"):
        ops.general_gemm(ops.swizzle_scales(qa), qc, "TN")


def test_mxfp8_gemm_full_size_linearity():
    """Config 4 fprop size: exact linearity of the block-scaled GEMM in powers of two."""
    n, m, k = 8192, 14336, 4096
    x = torch.randn(n, k, device="cuda").bfloat16()
    w = (torch.randn(m, k, device="cuda") * 0.02).bfloat16()
    qx = ops.mxfp8_quantize(x, swizzled_scales=True, columnwise=False)
    qw = ops.mxfp8_quantize(w, swizzled_scales=True, columnwise=False)
    y = ops.general_gemm(qw, qx, "TN", torch.float32)
    qx2 = ops.mxfp8_quantize(x * 2.0, swizzled_scales=True, columnwise=False)
    y2 = ops.general_gemm(qw, qx2, "TN", torch.float32)
    assert torch.equal(y2, y * 2.0)
    # a 128-row slab against the fp32 torch matmul of the dequantised operands
    dx = ops.dequantize(qx, torch.float32)[:128]
    dw = ops.dequantize(qw, torch.float32)
    ref = dx.double() @ dw.double().t()
    assert (y[:128].double() - ref).abs().max() <= 1e-5 * ref.abs().max()


@pytest.mark.parametrize("kind", ["fp8", "mxfp8"])
def test_gemm_output_row_blocks_land_in_place(kind):
    """`kNVTEMatmulConfigB200OutputRowBlocks`: the GEMM over piece c of a chunked all-gather (rows
    of every rank's shard, rank-major) writes its rows where they belong in the full output."""
    torch.manual_seed(5)
    world, rows_local, chunks, k, m = 4, 512, 2, 512, 384
    rows_c = rows_local // chunks
    x = torch.randn(world * rows_local, k, device="cuda").bfloat16()
    w = (torch.randn(m, k, device="cuda") * 0.1).bfloat16()
    if kind == "fp8":
        s = torch.ones(1, device="cuda")
        q = lambda t: ops.fp8_quantize(t, s)
    else:
        q = lambda t: ops.mxfp8_quantize(t, swizzled_scales=True)
    wq = q(w)
    full = ops.general_gemm(wq, q(x), "TN")
    out = torch.full_like(full, float("nan"))
    xr = x.view(world, chunks, rows_c, k)
    for c in range(chunks):
        piece = xr[:, c].reshape(world * rows_c, k).contiguous()   # what all-gather piece c holds
        ops.general_gemm(wq, q(piece), "TN", out=out, out_row_blocks=(rows_c, rows_local, c * rows_c),
                         sm_count=100 if c == 0 else 0)
    assert torch.equal(out, full)
    with pytest.raises(RuntimeError, match="row blocks"):
        ops.general_gemm(wq, q(x[:300]), "TN", out=out, out_row_blocks=(128, 512, 0))


# --------------------------------------------------------------------------- #
# nvte_multi_tensor_gemm: one persistent launch over the members' concatenated tile list
# (reference GroupedLinear path, tests/pytorch/test_numerics.py test_grouped_gemm: the grouped
# result must equal the member-by-member result)
# --------------------------------------------------------------------------- #
GROUP_SPLITS = [[256, 128, 384], [200, 0, 56, 512], [128] * 8, [1024, 96]]


def _quant(kind, t):
    if kind == "mxfp8":
        return ops.mxfp8_quantize(t, swizzled_scales=True)
    if kind == "fp8":
        return ops.fp8_quantize(t, torch.ones(1, device="cuda"))
    # rowwise only: the columnwise NVFP4 copy needs rows % 16 == 0, the fprop operands do not
    return ops.nvfp4_quantize(t, with_rht=False, swizzled_scales=True, columnwise=False)


@pytest.mark.parametrize("kind", ["mxfp8", "fp8", "nvfp4"])
@pytest.mark.parametrize("splits", GROUP_SPLITS)
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_multi_tensor_gemm_fprop_matches_member_by_member(kind, splits, out_dtype):
    """fprop of a GroupedLinear: member i = x_i[m_i, k] . W_i[n, k]^T, ragged and empty m_i."""
    k, n = 512, 384
    g = torch.Generator().manual_seed(7)
    xs = [torch.randn(m, k, generator=g).bfloat16().cuda() for m in splits]
    ws = [(torch.randn(n, k, generator=g) * 0.5).bfloat16().cuda() for _ in splits]
    live = [i for i, m in enumerate(splits) if m > 0]
    qx = {i: _quant(kind, xs[i]) for i in live}
    qw = {i: _quant(kind, ws[i]) for i in live}
    bias = [(torch.randn(n, generator=g)).to(out_dtype).cuda() for _ in live]
    outs = ops.multi_tensor_gemm([qw[i] for i in live], [qx[i] for i in live], "TN", out_dtype, bias=bias)
    for j, i in enumerate(live):
        ref = ops.general_gemm(qw[i], qx[i], "TN", out_dtype, bias=bias[j])
        assert outs[j].shape == (splits[i], n)
        assert torch.equal(outs[j], ref), f"member {i}: max diff {(outs[j].float() - ref.float()).abs().max()}"


@pytest.mark.parametrize("layout", ["NN", "NT"])
def test_multi_tensor_gemm_mxfp8_backward_layouts_and_accumulate(layout):
    """dgrad (NN) and wgrad (NT, K = tokens of the member, fp32 accumulate into main_grad)."""
    splits, k_in, n_out = [256, 128, 384], 512, 256
    g = torch.Generator().manual_seed(11)
    A, B, ref_out, outs = [], [], [], []
    for m in splits:
        if layout == "NN":      # dgrad: dx[m, in] = dy[m, out] . W[out, in]
            a = (torch.randn(n_out, k_in, generator=g) * 0.5).bfloat16().cuda()
            b = torch.randn(m, n_out, generator=g).bfloat16().cuda()
            shape = (m, k_in)
        else:                   # wgrad: dW[out, in] = dy[m, out]^T . x[m, in]
            a = torch.randn(m, k_in, generator=g).bfloat16().cuda()
            b = torch.randn(m, n_out, generator=g).bfloat16().cuda()
            shape = (n_out, k_in)
        A.append(_quant("mxfp8", a))
        B.append(_quant("mxfp8", b))
        init = torch.randn(*shape, generator=g).cuda()
        outs.append(init.clone())
        ref_out.append(ops.general_gemm(A[-1], B[-1], layout, torch.float32, accumulate=True, out=init.clone()))
    ops.multi_tensor_gemm(A, B, layout, torch.float32, accumulate=True, out=outs)
    for o, r in zip(outs, ref_out):
        assert torch.equal(o, r)


def test_multi_tensor_gemm_against_oracle_large_members():
    """Members big enough that every CTA pair crosses member boundaries several times."""
    splits, k, n = [1536, 640, 2048, 384], 1024, 1024
    g = torch.Generator().manual_seed(3)
    xs = [torch.randn(m, k, generator=g).bfloat16().cuda() for m in splits]
    ws = [(torch.randn(n, k, generator=g) * 0.5).bfloat16().cuda() for _ in splits]
    qx = [ops.mxfp8_quantize(x) for x in xs]
    qw = [ops.mxfp8_quantize(w) for w in ws]
    outs = ops.multi_tensor_gemm([ops.swizzle_scales(q) for q in qw], [ops.swizzle_scales(q) for q in qx], "TN",
                                 torch.bfloat16)
    for x, w, o in zip(qx, qw, outs):
        A = O.mxfp8_dequantize(u8(w.rowwise_data), u8(w.rowwise_scale_inv), colwise=False)
        B = O.mxfp8_dequantize(u8(x.rowwise_data), u8(x.rowwise_scale_inv), colwise=False)
        _assert_close(o.float().cpu().numpy(), O.gemm_tn_ref(A, B), torch.bfloat16)


def test_multi_tensor_gemm_replays_from_a_cuda_graph():
    """The member table reaches the device through a copy node: its host source must stay valid
    and unchanged for every replay, also after other grouped calls went through in between."""
    splits, k, n = [256, 128, 384], 512, 256
    g = torch.Generator().manual_seed(5)
    xs = [torch.randn(m, k, generator=g).bfloat16().cuda() for m in splits]
    ws = [(torch.randn(n, k, generator=g) * 0.5).bfloat16().cuda() for _ in splits]
    qx = [ops.mxfp8_quantize(x, swizzled_scales=True) for x in xs]
    qw = [ops.mxfp8_quantize(w, swizzled_scales=True) for w in ws]
    outs = [torch.zeros(m, n, dtype=torch.bfloat16, device="cuda") for m in splits]
    ref = ops.multi_tensor_gemm(qw, qx, "TN", torch.bfloat16)
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            ops.multi_tensor_gemm(qw, qx, "TN", out=outs)
    torch.cuda.synchronize()
    for rep in range(3):
        for o in outs:
            o.zero_()
        # an unrelated grouped call in between reuses the same device workspace
        ops.multi_tensor_gemm(qw[:2], qx[:2], "TN", torch.bfloat16)
        graph.replay()
        torch.cuda.synchronize()
        for o, r in zip(outs, ref):
            assert torch.equal(o, r), rep
