"""ctypes front-end of oracle/_ref/libte_ref_host.so: the REFERENCE's own host
reference functions (tests/cpp/operator/*.cu `compute_ref` & co.) compiled in
place by `make -C oracle ref`. TEST INFRASTRUCTURE ONLY.

Used to pin the C oracle restatement (tests/test_oracle.py, oracle/gen_golden.py)
and as the `"kind": "reference"` CPU baseline of bench.py.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_ref", "libte_ref_host.so")
_lib = None


def available():
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(LIB)
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _sz(x):
    return ctypes.c_size_t(int(x))


def mxfp8_compute_ref(x, e5m2=False, rowwise=True, colwise=True):
    """x: uint16 bf16 bits or float32, [rows, cols]. Compact scales."""
    x = np.ascontiguousarray(x)
    rows, cols = x.shape
    kb, rb = (cols + 31) // 32, (rows + 31) // 32
    qr = np.zeros((rows, cols), np.uint8)
    qc = np.zeros((rows, cols), np.uint8)
    sr = np.zeros((rows, kb), np.uint8)
    sc = np.zeros((rb, cols), np.uint8)
    if x.dtype == np.float32:
        assert not e5m2
        lib().ref_mxfp8_compute_ref_fp32(_p(x), int(rowwise), int(colwise), _p(qr), _p(qc), _p(sr),
                                         _p(sc), _sz(rows), _sz(cols), _sz(kb), _sz(cols))
    else:
        lib().ref_mxfp8_compute_ref_bf16(_p(x), int(e5m2), int(rowwise), int(colwise), _p(qr),
                                         _p(qc), _p(sr), _p(sc), _sz(rows), _sz(cols), _sz(kb),
                                         _sz(cols))
    return {"rowwise": qr, "rowwise_scale": sr, "colwise": qc, "colwise_scale": sc}


def nvfp4_quantize(x, global_amax, use_2d=False):
    x = np.ascontiguousarray(x)
    rows, cols = x.shape
    out = np.zeros((rows, cols // 2), np.uint8)
    sc = np.zeros((rows, cols // 16), np.uint8)
    lib().ref_nvfp4_quantize_bf16(_p(x), _p(out), _p(sc), _sz(rows), _sz(cols), _sz(cols // 16),
                                  ctypes.c_float(global_amax), int(use_2d))
    return out, sc


def nvfp4_compute_ref(x, amax, use_2d=False):
    """Rowwise + transposed columnwise NVFP4 of x (same amax both ways, no RHT)."""
    x = np.ascontiguousarray(x)
    rows, cols = x.shape
    out = np.zeros((rows, cols // 2), np.uint8)
    out_t = np.zeros((cols, rows // 2), np.uint8)
    sc = np.zeros((rows, cols // 16), np.uint8)
    sc_t = np.zeros((cols, rows // 16), np.uint8)
    lib().ref_nvfp4_compute_ref_bf16(_p(x), _p(out), _p(out_t), _p(sc), _p(sc_t),
                                     ctypes.c_float(amax), _sz(rows), _sz(cols), _sz(cols // 16),
                                     _sz(rows // 16), int(use_2d))
    return out, sc, out_t, sc_t


def fp8_cast_transpose(x, scale, e5m2=False):
    x = np.ascontiguousarray(x)
    rows, cols = x.shape
    oc = np.zeros((rows, cols), np.uint8)
    ot = np.zeros((cols, rows), np.uint8)
    amax = ctypes.c_float(0)
    lib().ref_fp8_cast_transpose(_p(x), int(x.dtype != np.float32), int(e5m2), _p(oc), _p(ot),
                                 _sz(rows), _sz(cols), ctypes.byref(amax), ctypes.c_float(scale))
    return oc, ot, amax.value


def amax_scale(x, max_fp8=448.0, epsilon=0.0):
    x = np.ascontiguousarray(x)
    a, s, si = ctypes.c_float(0), ctypes.c_float(0), ctypes.c_float(0)
    lib().ref_amax_scale(_p(x), int(x.dtype != np.float32), _sz(x.size), ctypes.byref(a),
                         ctypes.byref(s), ctypes.byref(si), ctypes.c_float(max_fp8),
                         ctypes.c_float(epsilon))
    return a.value, s.value, si.value


def swizzle(sc, M, K, row_scaling=True):
    out = np.zeros(M * K, np.uint8)
    lib().ref_swizzle(_p(np.ascontiguousarray(sc)), _p(out), _sz(M), _sz(K), int(row_scaling))
    return out


def unswizzle(sw, M, K, row_scaling=True):
    out = np.zeros(M * K, np.uint8)
    lib().ref_unswizzle(_p(np.ascontiguousarray(sw)), _p(out), _sz(M), _sz(K), int(row_scaling))
    return out


def norm_stats(x, eps, layernorm):
    """x: bf16 bits [N, H] -> (mu, rsigma) by the reference's compute_ref_stats."""
    x = np.ascontiguousarray(x)
    N, H = x.shape
    mu, rs = np.zeros(N, np.float32), np.zeros(N, np.float32)
    lib().ref_norm_stats(_p(x), int(layernorm), _sz(N), _sz(H), ctypes.c_double(eps), _p(mu), _p(rs))
    return mu, rs


def norm_output(x, gamma, beta, mu, rsigma, layernorm, zero_centered=False):
    """fp32 y and its amax by the reference's compute_ref_output (bf16 x / gamma / beta)."""
    x, gamma, beta = (np.ascontiguousarray(t) for t in (x, gamma, beta))
    N, H = x.shape
    y = np.zeros((N, H), np.float32)
    amax = ctypes.c_float(0)
    lib().ref_norm_output(_p(x), _p(gamma), _p(beta), int(layernorm), _sz(N), _sz(H), _p(mu), _p(rsigma),
                          int(zero_centered), _p(y), ctypes.byref(amax))
    return y, amax.value


def norm_backward(dz, x, mu, rsigma, gamma, layernorm, add=None, zero_centered=False):
    """bf16-bit dx, dgamma, dbeta by the reference's compute_ref_backward."""
    dz, x, gamma = (np.ascontiguousarray(t) for t in (dz, x, gamma))
    N, H = x.shape
    add = np.ascontiguousarray(add) if add is not None else np.zeros((N, H), np.uint16)
    dx = np.zeros((N, H), np.uint16)
    dg, db = np.zeros(H, np.uint16), np.zeros(H, np.uint16)
    lib().ref_norm_backward(_p(dz), _p(add), _p(x), int(layernorm), _sz(N), _sz(H), _p(mu), _p(rsigma), _p(gamma),
                            int(zero_centered), _p(dx), _p(dg), _p(db))
    return dx, dg, db
