"""CPU tests (no GPU): the oracle restatement against the reference's own outputs.

1. tests/golden/host_ref_hashes.json — SHA-256 of the outputs of the reference's
   C++ host references (compiled in place, oracle/_ref) on seeded inputs;
2. tests/golden/nvfp4_rht_*.npz — outputs of the reference's pure-torch NVFP4
   quantizer with the random Hadamard transform;
3. live comparison with oracle/_ref when that library is present;
4. independent cross-checks against torch's own float8 casts.
Both golden sets are produced by oracle/gen_golden.py.
"""
import glob
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O
from oracle import ref_host as R
from oracle.gen_golden import HOST_CASES
from tests.util import bits16, make_input

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(GOLD, "host_ref_hashes.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("case", HOST_CASES, ids=lambda c: f"{c[0]}x{c[1]}_{c[2]}")
def test_oracle_matches_reference_host_golden(case, golden):
    rows, cols, kind, seed = case
    g = golden[f"{rows}x{cols}_{kind}_{seed}"]
    x = bits16(make_input(rows, cols, kind, seed))
    xf = make_input(rows, cols, kind, seed, dtype=torch.float32).numpy()
    for e5 in (False, True):
        r = O.mxfp8_quantize(x, kind=int(e5))
        assert {k: sha(v) for k, v in r.items()} == g[f"mxfp8_{'e5m2' if e5 else 'e4m3'}"]
    assert {k: sha(v) for k, v in O.mxfp8_quantize(xf).items()} == g["mxfp8_e4m3_fp32in"]
    amax = O.amax(x)
    scale = O.scale_from_amax(amax)
    assert [amax, scale] == g["amax_scale"][:2]
    q, qt, am, sinv = O.fp8_quantize(x, scale, transpose=True)
    assert sinv == g["amax_scale"][2] or amax == 0.0
    assert {"data": sha(q), "transpose": sha(qt), "amax": am} == g["fp8_e4m3"]
    q, qt, am, _ = O.fp8_quantize(xf, 3.0, kind=1, transpose=True)
    assert {"data": sha(q), "transpose": sha(qt), "amax": am} == g["fp8_e5m2_fp32in"]
    if "nvfp4_1d" in g:
        xt = O.transpose_bf16(x)
        one_d = lambda a, am: O.nvfp4_quantize_1d(a, am, host_twin=True)  # noqa: E731
        two_d = lambda a, am: O.nvfp4_quantize_2d(a, am, host_twin=True)  # noqa: E731
        for name, fn in (("nvfp4_1d", one_d), ("nvfp4_2d", two_d)):
            d, s = fn(x, amax)
            dt, st = fn(xt, amax)
            assert {"data": sha(d), "scale": sha(s), "data_t": sha(dt), "scale_t": sha(st)} == g[name]


def test_oracle_swizzle_matches_reference_host_golden(golden):
    rng = np.random.default_rng(11)
    for M, K in [(128, 4), (256, 8), (384, 36)]:
        sc = rng.integers(0, 255, (M, K), dtype=np.uint8)
        g = golden[f"swizzle_{M}x{K}"]
        sw = O.swizzle_scales(sc)
        assert sha(sw) == g["row"]
        assert sha(O.swizzle_scales(sc.T.copy(), transposed_src=True)) == g["col"]
        assert np.array_equal(O.unswizzle_scales(sw, M, K), sc)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "nvfp4_rht_*.npz"))),
                         ids=os.path.basename)
def test_oracle_matches_reference_python_nvfp4_rht(path):
    """reference quantization_ref_nvfp4.py (`NVFP4QuantizerRef`, with_rht=True)."""
    g = np.load(path)
    rows, cols, kind, seed = os.path.basename(path)[len("nvfp4_rht_"):-4].replace("x", "_").split("_")
    x = bits16(make_input(int(rows), int(cols), kind, int(seed)))
    amax_row = O.amax(x)
    assert amax_row == g["amax_row"][0]
    d, s = O.nvfp4_quantize_1d(x, amax_row)
    assert np.array_equal(d, g["data"]) and np.array_equal(s, g["scale"])
    rht_t = O.rht_transposed_bf16(x)
    assert sha(rht_t) == str(g["rht_t_sha"])
    amax_col = O.amax(rht_t)
    assert amax_col == g["amax_col"][0]
    dt, st = O.nvfp4_quantize_1d(rht_t, amax_col)
    assert np.array_equal(dt, g["data_t"]) and np.array_equal(st, g["scale_t"])


def test_rht_matrix_constants():
    """sign vector / Sylvester matrix of reference pytorch/tensor/nvfp4_tensor.py:48-110"""
    H = np.load(os.path.join(GOLD, "rht_matrix_16x16.npy"))
    eye = np.eye(16, dtype=np.float32)
    got = np.stack([
        (O.rht_transposed_bf16(bits16(torch.from_numpy(eye[i].reshape(16, 1).copy()).bfloat16()))
         .astype(np.uint32) << 16).view(np.float32)[0] for i in range(16)])
    assert np.array_equal(got, H)


@pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("shape,kind", [((96, 160), "randn"), ((257, 96), "uniform"),
                                        ((128, 512), "outliers")])
def test_oracle_matches_reference_host_live(shape, kind):
    x = bits16(make_input(*shape, kind, seed=9))
    r, o = R.mxfp8_compute_ref(x), O.mxfp8_quantize(x)
    for k in r:
        assert np.array_equal(r[k], o[k]), k
    amax = O.amax(x)
    if shape[0] % 16 == 0:
        one_d = lambda a, am: O.nvfp4_quantize_1d(a, am, host_twin=True)  # noqa: E731
        two_d_fn = lambda a, am: O.nvfp4_quantize_2d(a, am, host_twin=True)  # noqa: E731
        for two_d, fn in ((False, one_d), (True, two_d_fn)):
            rd, rs = R.nvfp4_quantize(x, amax, two_d)
            od, os_ = fn(x, amax)
            assert np.array_equal(rd, od) and np.array_equal(rs, os_)


# ---- independent cross-checks ----------------------------------------------------
def test_scalar_conversions_against_torch():
    vals = torch.cat([torch.linspace(-500, 500, 20001), torch.randn(20000) * 1e-2,
                      torch.tensor([0.0, 448.0, 464.0, 465.0, 1e9, -1e9, 2.0 ** -9, 2.0 ** -10,
                                    1.5 * 2.0 ** -10, 57344.0, 61439.0, 61440.0])])
    e4 = vals.clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    e5 = vals.clamp(-57344, 57344).to(torch.float8_e5m2).view(torch.uint8).numpy()
    bf = vals.to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    l = O.lib()
    for v, a, b, c in zip(vals.tolist(), e4, e5, bf):
        assert l.teo_cvt_e4m3(v) == a, v
        assert l.teo_cvt_e5m2(v) == b, v
        assert l.teo_cvt_bf16(v) == c, v


def test_e2m1_grid_and_ties():
    """ties-to-even on {0,.5,1,1.5,2,3,4,6} (reference quantization_ref_nvfp4.py:49-71)"""
    l = O.lib()
    expect = {0.0: 0, 0.25: 0, 0.26: 1, 0.75: 2, 1.25: 2, 1.75: 4, 2.5: 4, 3.5: 6, 5.0: 6,
              5.01: 7, 6.0: 7, 100.0: 7, -0.25: 8, -0.3: 9, -6.5: 15}
    for v, code in expect.items():
        assert l.teo_cvt_e2m1(v) == code, v


def test_e8m0_known_answers():
    """reference tests/cpp/test_common.h:480-513"""
    l = O.lib()
    assert l.teo_cvt_e8m0(0.0) == 0
    assert l.teo_cvt_e8m0(1.0) == 127
    assert l.teo_cvt_e8m0(1.0000001) == 128          # any mantissa rounds the exponent up
    assert l.teo_cvt_e8m0(float("inf")) == 0xFE
    assert l.teo_cvt_e8m0(float("nan")) == 0xFF
    assert l.teo_cvt_e8m0(3.0) == 129
    assert l.teo_cvt_e8m0(2.0 ** -127) == 0           # denormal with mantissa <= 0x400000
    assert l.teo_cvt_e8m0_rcp(127) == 1.0
    assert l.teo_cvt_e8m0_rcp(254) == 2.0 ** -127
    assert l.teo_cvt_e8m0_rcp(0) == 2.0 ** 127


def test_delayed_scaling_closed_form():
    """reference tests/pytorch/test_recipe.py:110-180"""
    hist = np.array([[4.0, 0.5], [1.0, 8.0], [2.0, 3.0]], np.float32)
    scale = np.array([1.0, 1.0], np.float32)
    h, s = O.delayed_scaling_update(hist, scale, algo="max", margin=0.0, bulk=False)
    assert np.array_equal(h, [[0.0, 0.0], [2.0, 3.0], [4.0, 0.5]])
    assert np.array_equal(s, np.float32(448.0) / np.array([4.0, 8.0], np.float32))
    h, s = O.delayed_scaling_update(hist, scale, algo="most_recent", margin=1.0, bulk=False)
    assert np.array_equal(s, np.float32(224.0) / np.array([4.0, 0.5], np.float32))


# --------------------------------------------------------------------------- #
# LayerNorm / RMSNorm oracle (oracle/teo_norm.c) against the reference's host references
# (tests/golden/norm_*.npz, made by oracle/gen_golden.py from tests/cpp/operator/test_normalization.h)
# --------------------------------------------------------------------------- #
def _bf16_bits(a):
    return bits16(torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "norm_*.npz"))), ids=os.path.basename)
def test_norm_oracle_matches_reference_host_golden(path):
    g = np.load(path)
    ln, zc = bool(g["layernorm"]), bool(g["zero_centered"])
    mu, rs = O.norm_stats(g["x"], float(g["eps"]), ln)
    assert np.array_equal(rs, g["rsigma"])            # same sequential fp32 sums: exact
    if ln:
        assert np.array_equal(mu, g["mu"])
    y = O.norm_output(g["x"], g["gamma"], g["beta"] if ln else None, mu, rs, zero_centered=zc)
    if ln:   # the oracle fuses y * g + beta as the reference's GPU kernel does: <= 1 rounding apart
        np.testing.assert_allclose(y, g["y"], rtol=0, atol=2.0 ** -22 * float(g["amax"]))
    else:
        assert np.array_equal(y, g["y"])
    dx, dg, db = O.norm_backward(g["dz"], g["x"], mu, rs, g["gamma"], add=None if ln else g["add"], zero_centered=zc)
    assert np.array_equal(_bf16_bits(dx), g["dx"])
    assert np.array_equal(_bf16_bits(dg), g["dgamma"])
    if ln:
        assert np.array_equal(_bf16_bits(db), g["dbeta"])


@pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("ln", [True, False])
def test_norm_oracle_matches_reference_host_live(ln):
    x = bits16(make_input(40, 320, "outliers", 7))
    gamma = bits16(make_input(1, 320, "uniform", 8)[0])
    beta = bits16(make_input(1, 320, "randn", 9)[0])
    mu_r, rs_r = R.norm_stats(x, 1e-5, ln)
    mu, rs = O.norm_stats(x, 1e-5, ln)
    assert np.array_equal(rs, rs_r) and (not ln or np.array_equal(mu, mu_r))
    y_r, amax = R.norm_output(x, gamma, beta, mu_r, rs_r, ln)
    y = O.norm_output(x, gamma, beta if ln else None, mu, rs)
    np.testing.assert_allclose(y, y_r, rtol=0, atol=2.0 ** -22 * amax)
