"""GEMM parity AT THE SIZES THE BENCH TIMES: the persistent tile loop's steady state
(several tiles per CTA pair, accumulator phase wrap, the overlapping-accumulator protocol, relaxed
hand-offs) compared with an independent value -- the oracle's dequantise-then-fp64 matmul -- on

  * >= 64 full output rows spread over EVERY 128-row block of the output (varying lane / quadrant), and
  * >= 64 full output columns spread over every 128-column block,

for MXFP8 TN / NN / NT at config 4 (8192 x 14336 x 4096 and its dgrad / wgrad shapes), NVFP4 TN at config 3
(16384 x 8192 x 8192) and per-tensor FP8 TN at config 5 (8192 x 28672 x 8192); plus mid-size problems with
>= 4 tiles per CTA pair (4096 x 4096 x 1024: 256 tiles of 256 x 256 on 74 pairs) checked over the WHOLE output.

Tolerances: fp32 output <= 1e-5 of max|ref| (2e-5 for NVFP4, whose alpha folds two amaxes);
bf16 output rtol = atol = 8e-3 (reference tests/pytorch/nvfp4/test_nvfp4_gemm_exact.py:239).
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


class _Operand:
    """A quantized GEMM operand + host copies of its bytes, with `rows(idx)` = the dequantised rows `idx`
    of its [mn, k] view (whatever its storage orientation), computed by the ORACLE."""

    def __init__(self, kind, t, colwise, rht=False, seed_scale=None):
        self.kind, self.colwise = kind, colwise
        if kind == "mxfp8":
            q = ops.mxfp8_quantize(t, rowwise=not colwise, columnwise=colwise)
            self.gemm = ops.swizzle_scales(q)
            self.data = u8(q.columnwise_data if colwise else q.rowwise_data)
            self.sc = u8(q.columnwise_scale_inv if colwise else q.rowwise_scale_inv)
        elif kind == "nvfp4":
            q = ops.nvfp4_quantize(t, with_rht=rht, rowwise=not colwise, columnwise=colwise)
            self.gemm = ops.swizzle_scales(q)
            self.data = u8(q.columnwise_data if colwise else q.rowwise_data)   # [mn, k/2] either way
            self.sc = u8(q.columnwise_scale_inv if colwise else q.rowwise_scale_inv)
            self.amax = (q.amax_columnwise if colwise else q.amax_rowwise).item()
        else:
            self.gemm = ops.fp8_quantize(t, seed_scale)
            self.data = u8(self.gemm.data)           # [mn, k] rowwise, [k, mn] when used MN-major
            self.sinv = self.gemm.scale_inv.item()

    def rows(self, idx):
        idx = np.asarray(idx)
        if self.kind == "mxfp8":
            if self.colwise:     # data [k, mn], scales [k/32 (pad 4), mn (pad 128)]
                d = np.ascontiguousarray(self.data[:, idx])
                s = np.ascontiguousarray(self.sc[:, idx])
                return np.ascontiguousarray(O.mxfp8_dequantize(d, s, colwise=True).T)
            return O.mxfp8_dequantize(np.ascontiguousarray(self.data[idx]), np.ascontiguousarray(self.sc[idx]))
        if self.kind == "nvfp4":
            return O.nvfp4_dequantize(np.ascontiguousarray(self.data[idx]), np.ascontiguousarray(self.sc[idx]),
                                      self.amax)
        d = self.data[:, idx].T if self.colwise else self.data[idx]
        f = torch.from_numpy(np.ascontiguousarray(d)).view(torch.float8_e4m3fn).float().numpy()
        return f * np.float32(self.sinv)


def _spread(n, at_least=64):
    """One index per 128-block of [0, n), more per block when there are fewer than `at_least` blocks, at
    offsets that walk through all lanes / quadrants."""
    blocks = (n + 127) // 128
    per = max(1, -(-at_least // blocks))
    out = []
    for b in range(blocks):
        for j in range(per):
            i = b * 128 + (37 * (b * per + j) + 5) % 128
            if i < n:
                out.append(i)
    return np.array(sorted(set(out)))


def _check(got, ref, out_dtype, kind, what):
    got = got.float().cpu().numpy()
    assert np.isfinite(got).all(), what
    if out_dtype == torch.float32:
        tol = 2e-5 if kind == "nvfp4" else 1e-5
        err = np.abs(got - ref).max() / np.abs(ref).max()
        assert err <= tol, f"{what}: rel err {err:.3e} > {tol}"
    else:
        np.testing.assert_allclose(got, ref, rtol=8e-3, atol=8e-3, err_msg=what)


def _run_case(kind, layout, n, m, k, whole=False):
    # TE convention: TN: A[m,k], B[n,k]; NN: A[k,m], B[n,k]; NT: A[k,m], B[k,n]; D[n,m]
    ca, cb = layout[0] != "T", layout[1] == "T"
    g = torch.Generator(device="cuda").manual_seed(1234 + n + m + k)
    a = (torch.randn(*((k, m) if ca else (m, k)), generator=g, device="cuda") * 0.05).bfloat16()
    b = torch.randn(*((k, n) if cb else (n, k)), generator=g, device="cuda").bfloat16()
    rht = kind == "nvfp4" and layout == "NT"
    sa = torch.tensor([211.0], device="cuda") if kind == "fp8" else None
    sb = torch.tensor([37.5], device="cuda") if kind == "fp8" else None
    A = _Operand(kind, a, ca, rht, sa)
    B = _Operand(kind, b, cb, rht, sb)
    del a, b
    outs = {dt: ops.general_gemm(A.gemm, B.gemm, layout, dt) for dt in (torch.float32, torch.bfloat16)}
    torch.cuda.synchronize()
    assert outs[torch.float32].shape == (n, m)
    if whole:
        ref = O.gemm_tn_ref(A.rows(np.arange(m)), B.rows(np.arange(n)))
        for dt, o in outs.items():
            _check(o, ref, dt, kind, f"{kind} {layout} {n}x{m}x{k} whole output {dt}")
        return
    ridx, cidx = _spread(n), _spread(m)
    assert len(ridx) >= 64 and len(cidx) >= 64
    ref_rows = O.gemm_tn_ref(A.rows(np.arange(m)), B.rows(ridx))       # [len(ridx), m]
    for dt, o in outs.items():
        _check(o[torch.from_numpy(ridx).cuda()], ref_rows, dt, kind, f"{kind} {layout} {n}x{m}x{k} rows {dt}")
    del ref_rows
    ref_cols = O.gemm_tn_ref(A.rows(cidx), B.rows(np.arange(n)))       # [n, len(cidx)]
    for dt, o in outs.items():
        _check(o[:, torch.from_numpy(cidx).cuda()], ref_cols, dt, kind, f"{kind} {layout} {n}x{m}x{k} cols {dt}")


@pytest.mark.parametrize("layout,n,m,k", [("TN", 8192, 14336, 4096),     # fprop
                                          ("NN", 8192, 4096, 14336),     # dgrad
                                          ("NT", 14336, 4096, 8192)])    # wgrad
def test_mxfp8_gemm_config4_rows_and_columns_over_every_tile(layout, n, m, k):
    _run_case("mxfp8", layout, n, m, k)


def test_nvfp4_gemm_config3_rows_and_columns_over_every_tile():
    _run_case("nvfp4", "TN", 16384, 8192, 8192)


def test_fp8_gemm_config5_rows_and_columns_over_every_tile():
    _run_case("fp8", "TN", 8192, 28672, 8192)


@pytest.mark.parametrize("kind", ["mxfp8", "nvfp4", "fp8"])
@pytest.mark.parametrize("layout", ["TN", "NN", "NT"])
def test_gemm_midsize_whole_output_four_tiles_per_pair(kind, layout):
    """4096 x 4096 x 1024: 256 tiles on 74 CTA pairs (3.5 tiles per pair: every pair wraps its accumulator
    phase), every output element against the oracle."""
    _run_case(kind, layout, 4096, 4096, 1024, whole=True)


@pytest.mark.parametrize("kind", ["mxfp8", "nvfp4", "fp8"])
def test_gemm_many_tiles_small_k(kind):
    """K = 256 (two k-blocks per tile at most): the epilogue, not the MMA loop, is the critical path and tiles
    hand over back to back -- 8192 x 2048 x 256, whole output."""
    _run_case(kind, "TN", 8192, 2048, 256, whole=True)
