"""`nvte_b200_quantize_dbias_static_scratch` -- the no-allocation twin of `nvte_quantize_dbias` (partial sums
in a static device array of the library instead of a caller-allocated workspace; no counterpart in the
reference, see include/transformerengine_b200.h) -- called through the C ABI and compared with the existing
entry point on the same inputs: cast bytes and scales equal, dbias equal (the same kernels run on the same
values, so the tolerance of the existing dbias test collapses to equality; the test still states it).

Two calls in a row per case, the second on a DIFFERENT and smaller input, so that a call that depended on
the scratch's previous contents (it keeps them) would show. This file is collected last on purpose
(`pytest -x`): it must pass alone in a fresh process and after every other test.
"""
import ctypes

import pytest
import torch

from tests.util import make_input

pytestmark = pytest.mark.gpu

LIMIT = 8 << 20    # NVTE_B200_DBIAS_STATIC_SCRATCH_BYTES


def _mods():
    from transformerengine_b200 import cabi as C
    from transformerengine_b200 import ops
    return C, ops


def _fp8_out(C, x, scale):
    data = torch.zeros(x.shape, dtype=torch.uint8, device="cuda")
    sinv = torch.zeros(1, dtype=torch.float32, device="cuda")
    amax = torch.zeros(1, dtype=torch.float32, device="cuda")
    h = C.TensorHandle(C.DELAYED_TENSOR_SCALING)
    h.set(C.kRowwiseData, data, C.kFloat8E4M3).set(C.kRowwiseScaleInv, sinv).set(C.kScale, scale).set(C.kAmax, amax)
    return h, (data, sinv, amax)


def _mx_out(ops, x, swizzled):
    rows, cols = x.shape
    rs, cs = ops.mxfp8_scale_shapes(rows, cols)
    st = ops.MXFP8TensorStorage(
        rowwise_data=torch.zeros(x.shape, dtype=torch.uint8, device="cuda"),
        rowwise_scale_inv=torch.zeros(rs, dtype=torch.uint8, device="cuda"),
        columnwise_data=torch.zeros(x.shape, dtype=torch.uint8, device="cuda"),
        columnwise_scale_inv=torch.zeros(cs, dtype=torch.uint8, device="cuda"),
        fp8_dtype=torch.float8_e4m3fn, shape=tuple(x.shape), with_gemm_swizzled_scales=swizzled)
    return st.handle(), (st.rowwise_data, st.rowwise_scale_inv, st.columnwise_data, st.columnwise_scale_inv)


def _both(kind, x, check_mem=False):
    """(buffers + dbias of nvte_quantize_dbias with a torch-allocated workspace,
        buffers + dbias of the static-scratch entry point) for one input."""
    C, ops = _mods()
    lib = C.lib()
    scale = torch.full((1,), 3.0, device="cuda")
    res = []
    for static in (False, True):
        h, bufs = _fp8_out(C, x, scale) if kind == "fp8" else _mx_out(ops, x, kind == "mxfp8_swz")
        if static:
            dbias = torch.zeros(x.shape[1], dtype=x.dtype, device="cuda")
            inp, hd = C.plain(x), C.plain(dbias)
            free0 = torch.cuda.mem_get_info()[0]
            C.check(lib.nvte_b200_quantize_dbias_static_scratch(inp.ptr, h.ptr, hd.ptr, C.stream_ptr()))
            torch.cuda.synchronize()
            assert not check_mem or torch.cuda.mem_get_info()[0] >= free0, "the call took device memory"
        else:
            dbias = ops.bgrad_quantize(x, h)
        res.append((bufs, dbias))
    torch.cuda.synchronize()
    return res


@pytest.mark.parametrize("kind", ["mxfp8", "mxfp8_swz", "fp8"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_entry_point_matches_quantize_dbias_twice_in_a_row(kind, dtype):
    # 1st call: config-4 sized columns, many slabs (the fused MXFP8 column pass for bf16; the slab kernels
    # otherwise); 2nd call: a smaller, ragged input that leaves most of the scratch holding the 1st call's sums
    for call, (rows, cols, seed) in enumerate([(2048, 4096, 3), (96, 160, 4)]):
        x = make_input(rows, cols, "randn", seed=seed, dtype=dtype).cuda()
        # (free device memory is compared across the 2nd call only: the 1st one may load kernel images)
        (ref_bufs, ref_db), (got_bufs, got_db) = _both(kind, x, check_mem=call == 1)
        for a, b in zip(ref_bufs, got_bufs):
            assert torch.equal(a, b), f"call {call}: quantized output differs from nvte_quantize_dbias"
        want = x.double().sum(0)
        tol = (2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -20) * (want.abs().max().item() + rows ** 0.5)
        assert got_db.dtype == dtype
        assert (got_db.double() - ref_db.double()).abs().max().item() <= tol
        assert (got_db.double() - want).abs().max().item() <= tol
        assert torch.equal(got_db, ref_db), f"call {call}: same kernels, same values: dbias must be equal"


def test_static_scratch_limit_is_enforced_with_a_clear_error():
    C, ops = _mods()
    lib = C.lib()
    # exactly at the limit: 256 slabs x 8192 columns x 4 bytes = 8 MiB -> accepted
    x = make_input(8192, 8192, "randn", seed=6, dtype=torch.bfloat16).cuda()
    (ref_bufs, ref_db), (got_bufs, got_db) = _both("mxfp8", x)
    assert all(torch.equal(a, b) for a, b in zip(ref_bufs, got_bufs)) and torch.equal(ref_db, got_db)
    # one slab more -> rejected before any launch, output untouched
    x = torch.ones(8192 + 32, 8192, dtype=torch.bfloat16, device="cuda")
    assert (8192 + 32) // 32 * 8192 * 4 > LIMIT
    h, bufs = _mx_out(ops, x, False)
    dbias = torch.full((8192,), 7.0, dtype=torch.bfloat16, device="cuda")
    inp, hd = C.plain(x), C.plain(dbias)
    rc = lib.nvte_b200_quantize_dbias_static_scratch(inp.ptr, h.ptr, hd.ptr, C.stream_ptr())
    assert rc != 0
    msg = lib.nvte_b200_last_error().decode()
    assert "static" in msg and str(LIMIT) in msg and "nvte_quantize_dbias" in msg, msg
    torch.cuda.synchronize()
    assert (dbias == 7.0).all() and not bufs[0].any()


def test_static_scratch_symbol_takes_no_workspace_argument():
    C, _ = _mods()
    fn = C.lib().nvte_b200_quantize_dbias_static_scratch
    assert fn.restype is ctypes.c_int and len(fn.argtypes) == 4
