"""ctypes front-end of oracle/_build/libte_oracle.so (numpy in, numpy out).

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py. Inputs are numpy arrays;
bf16 tensors travel as uint16 bit patterns (`dtype_code` 1), fp16 as uint16
(`dtype_code` 2), fp32 as float32 (`dtype_code` 0).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libte_oracle.so")
F32, BF16, F16 = 0, 1, 2
_lib = None


def build():
    subprocess.run(["make", "-s", "-C", _HERE], check=True)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        l = ctypes.CDLL(_LIB)
        l.teo_amax.restype = ctypes.c_float
        l.teo_scale_from_amax.restype = ctypes.c_float
        l.teo_scale_from_amax.argtypes = [ctypes.c_float, ctypes.c_float, ctypes.c_int, ctypes.c_float]
        l.teo_nvfp4_global_encode_scale.restype = ctypes.c_float
        l.teo_nvfp4_global_encode_scale.argtypes = [ctypes.c_float]
        l.teo_nvfp4_alpha.restype = ctypes.c_float
        l.teo_nvfp4_alpha.argtypes = [ctypes.c_float] * 3
        l.teo_swizzle_index.restype = ctypes.c_size_t
        l.teo_swizzle_index.argtypes = [ctypes.c_size_t] * 3
        l.teo_cvt_e8m0_rcp.restype = ctypes.c_float
        for n in ("teo_cvt_e4m3", "teo_cvt_e5m2", "teo_cvt_e2m1", "teo_cvt_e8m0"):
            getattr(l, n).argtypes = [ctypes.c_float]
            getattr(l, n).restype = ctypes.c_uint8
        l.teo_cvt_bf16.argtypes = [ctypes.c_float]
        l.teo_cvt_bf16.restype = ctypes.c_uint16
        _lib = l
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _code(a):
    if a.dtype == np.float32:
        return F32
    assert a.dtype == np.uint16, "bf16/fp16 inputs travel as uint16 bit patterns"
    return BF16


def _sz(x):
    return ctypes.c_size_t(int(x))


def num_threads():
    return lib().teo_num_threads()


def set_num_threads(n: int) -> int:
    """OpenMP threads of the oracle and of oracle/_ref (same libgomp) from now on; returns the new team size."""
    lib().teo_set_num_threads(int(n))
    return num_threads()


def amax(x, code=None):
    x = np.ascontiguousarray(x)
    return float(lib().teo_amax(_p(x), code if code is not None else _code(x), _sz(x.size)))


def scale_from_amax(amax_v, fp8_max=448.0, force_pow2=False, eps=0.0):
    return float(lib().teo_scale_from_amax(amax_v, fp8_max, int(force_pow2), eps))


def fp8_quantize(x, scale, kind=0, amax_in=0.0, transpose=False, code=None):
    """Returns (q [rows, cols] u8, q_t or None, amax, scale_inv)."""
    x = np.ascontiguousarray(x)
    rows, cols = x.reshape(-1, x.shape[-1]).shape
    out = np.empty((rows, cols), np.uint8)
    out_t = np.empty((cols, rows), np.uint8) if transpose else None
    am = ctypes.c_float(amax_in)
    sinv = ctypes.c_float(0)
    lib().teo_fp8_quantize(_p(x), code if code is not None else _code(x), _sz(rows), _sz(cols),
                           ctypes.c_float(scale), kind, _p(out), _p(out_t), ctypes.byref(am),
                           ctypes.byref(sinv))
    return out, out_t, am.value, sinv.value


def delayed_scaling_update(history, scale, reduced_amax=None, algo="max", fp8_max=448.0,
                           margin=0.0, bulk=True):
    h = np.ascontiguousarray(history, np.float32).copy()
    s = np.ascontiguousarray(scale, np.float32).copy()
    r = np.ascontiguousarray(reduced_amax, np.float32) if reduced_amax is not None else None
    lib().teo_delayed_scaling_update(_p(h), _p(s), _p(r), _sz(h.shape[0]), _sz(h.shape[1]),
                                     1 if algo == "most_recent" else 0, ctypes.c_float(fp8_max),
                                     ctypes.c_float(margin), int(bulk))
    return h, s


def mxfp8_quantize(x, kind=0, rowwise=True, colwise=True, code=None, with_amax=False):
    """Compact scales: rowwise [rows, ceil(cols/32)], colwise [ceil(rows/32), cols]."""
    x = np.ascontiguousarray(x)
    rows, cols = x.reshape(-1, x.shape[-1]).shape
    kb, rb = (cols + 31) // 32, (rows + 31) // 32
    qr = np.zeros((rows, cols), np.uint8) if rowwise else None
    sr = np.zeros((rows, kb), np.uint8) if rowwise else None
    qc = np.zeros((rows, cols), np.uint8) if colwise else None
    sc = np.zeros((rb, cols), np.uint8) if colwise else None
    am = ctypes.c_float(0)
    lib().teo_mxfp8_quantize(_p(x), code if code is not None else _code(x), _sz(rows), _sz(cols),
                             kind, _p(qr), _p(sr), _sz(kb), _p(qc), _p(sc), _sz(cols),
                             ctypes.byref(am) if with_amax else None)
    res = {"rowwise": qr, "rowwise_scale": sr, "colwise": qc, "colwise_scale": sc}
    if with_amax:
        res["amax"] = am.value
    return res


def mxfp8_dequantize(q, sc, colwise=False, kind=0):
    rows, cols = q.shape
    out = np.empty((rows, cols), np.float32)
    lib().teo_mxfp8_dequantize(_p(q), _p(np.ascontiguousarray(sc)), _sz(sc.shape[1]), int(colwise),
                               _sz(rows), _sz(cols), kind, _p(out))
    return out


def nvfp4_quantize_1d(x, global_amax, code=None, host_twin=False):
    """x [rows, cols] -> (data [rows, cols/2] u8, scales [rows, cols/16] e4m3 bytes)."""
    x = np.ascontiguousarray(x)
    rows, cols = x.shape
    out = np.empty((rows, cols // 2), np.uint8)
    sc = np.empty((rows, cols // 16), np.uint8)
    lib().teo_nvfp4_quantize_1d(_p(x), code if code is not None else _code(x), _sz(rows), _sz(cols),
                                ctypes.c_float(global_amax), _p(out), _p(sc), _sz(cols // 16),
                                int(host_twin))
    return out, sc


def nvfp4_quantize_2d(x, global_amax, code=None, host_twin=False):
    """16x16-tile NVFP4: (data [rows, cols/2], scales [rows, cols/16] replicated per tile row)."""
    x = np.ascontiguousarray(x)
    rows, cols = x.shape
    out = np.empty((rows, cols // 2), np.uint8)
    sc = np.empty((rows, cols // 16), np.uint8)
    lib().teo_nvfp4_quantize_2d(_p(x), code if code is not None else _code(x), _sz(rows), _sz(cols),
                                ctypes.c_float(global_amax), _p(out), _p(sc), _sz(cols // 16),
                                int(host_twin))
    return out, sc


def nvfp4_dequantize(q, sc, global_amax):
    rows, cols = q.shape[0], q.shape[1] * 2
    out = np.empty((rows, cols), np.float32)
    lib().teo_nvfp4_dequantize(_p(q), _p(np.ascontiguousarray(sc)), _sz(sc.shape[1]), _sz(rows),
                               _sz(cols), ctypes.c_float(global_amax), _p(out))
    return out


def transpose_bf16(x):
    rows, cols = x.shape
    out = np.empty((cols, rows), np.uint16)
    lib().teo_transpose_bf16(_p(np.ascontiguousarray(x)), _sz(rows), _sz(cols), _p(out))
    return out


def rht_transposed_bf16(x, sign_mask=0xD7E8):
    """bf16(RHT(x^T)) as uint16 [cols, rows]."""
    rows, cols = x.shape
    out = np.empty((cols, rows), np.uint16)
    lib().teo_rht_transposed_bf16(_p(np.ascontiguousarray(x)), _sz(rows), _sz(cols),
                                  ctypes.c_uint(sign_mask), _p(out))
    return out


def swizzle_scales(sc, transposed_src=False):
    """sc: padded compact scales. Row-major [M, K], or [K, M] when transposed_src."""
    sc = np.ascontiguousarray(sc)
    M, K = (sc.shape[1], sc.shape[0]) if transposed_src else sc.shape
    out = np.zeros(M * K, np.uint8)
    lib().teo_swizzle_scales(_p(sc), _sz(sc.shape[1]), int(transposed_src), _sz(M), _sz(K), _p(out))
    return out


def unswizzle_scales(sw, M, K, transposed_dst=False):
    shape = (K, M) if transposed_dst else (M, K)
    out = np.zeros(shape, np.uint8)
    lib().teo_unswizzle_scales(_p(np.ascontiguousarray(sw)), _sz(shape[1]), int(transposed_dst),
                               _sz(M), _sz(K), _p(out))
    return out


def gemm_tn_ref(A, B, alpha=1.0, beta=0.0, C=None, bias=None):
    """D[N, M] = alpha * B[N,K] @ A[M,K]^T (+ bias[M]) (+ beta*C), fp64 accumulate."""
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    M, K = A.shape
    N = B.shape[0]
    D = np.empty((N, M), np.float32)
    Cc = np.ascontiguousarray(C, np.float32) if C is not None else None
    bb = np.ascontiguousarray(bias, np.float32) if bias is not None else None
    lib().teo_gemm_tn_ref(_p(A), _p(B), _sz(M), _sz(N), _sz(K), ctypes.c_float(alpha),
                          ctypes.c_float(beta), _p(Cc), _p(bb), _p(D))
    return D


# --------------------------------------------------------------------------- #
# LayerNorm / RMSNorm (oracle/teo_norm.c)
# --------------------------------------------------------------------------- #
def norm_stats(x, eps, layernorm, code=None):
    """(mu or None, rsigma): sequential fp32 sums (reference test_normalization.h:39-63)."""
    x = np.ascontiguousarray(x)
    N, H = x.shape
    mu = np.zeros(N, np.float32)
    rs = np.zeros(N, np.float32)
    lib().teo_norm_stats(_p(x), code if code is not None else _code(x), _sz(N), _sz(H), ctypes.c_double(eps), int(layernorm), _p(mu), _p(rs))
    return (mu if layernorm else None), rs


def norm_output(x, gamma, beta, mu, rsigma, zero_centered=False, code=None, wcode=None):
    """fp32 y for GIVEN statistics (beta/mu None: RMSNorm); elementwise, exact."""
    x = np.ascontiguousarray(x)
    gamma = np.ascontiguousarray(gamma)
    N, H = x.shape
    layernorm = mu is not None
    beta = np.ascontiguousarray(beta) if layernorm else None
    y = np.empty((N, H), np.float32)
    lib().teo_norm_output(_p(x), code if code is not None else _code(x), _p(gamma), _p(beta),
                          wcode if wcode is not None else _code(gamma), _sz(N), _sz(H),
                          _p(np.ascontiguousarray(mu, np.float32)) if layernorm else None,
                          _p(np.ascontiguousarray(rsigma, np.float32)), int(layernorm), int(zero_centered), _p(y))
    return y


def norm_backward(dz, x, mu, rsigma, gamma, add=None, zero_centered=False):
    """(dx, dgamma, dbeta or None) in fp32 (reference test_normalization.h:140-195)."""
    dz, x, gamma = (np.ascontiguousarray(t) for t in (dz, x, gamma))
    N, H = x.shape
    layernorm = mu is not None
    dx = np.empty((N, H), np.float32)
    dg = np.empty(H, np.float32)
    db = np.empty(H, np.float32) if layernorm else None
    add = np.ascontiguousarray(add) if add is not None else None
    lib().teo_norm_backward(_p(dz), _p(x), _p(add), _code(x),
                            _p(np.ascontiguousarray(mu, np.float32)) if layernorm else None,
                            _p(np.ascontiguousarray(rsigma, np.float32)), _p(gamma), _code(gamma), _sz(N), _sz(H),
                            int(layernorm), int(zero_centered), _p(dx), _p(dg), _p(db))
    return dx, dg, db
