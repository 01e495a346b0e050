"""world_size-2 gloo tests (CPU) of the tensor-parallel host logic: a row shard's quantization
equals the matching rows of the full tensor's quantization (given the all-reduced amax), and
`gather_along_first_dim` reassembles data + compact scales of every format into exactly the
full-tensor storage. Quantized bytes come from the oracle (test infrastructure); the code under
test is transformerengine_b200/distributed.py, which only moves bytes.
Mirrors reference tests/pytorch/distributed/run_numerics.py:396-455 (amax reduction),
:511-640 (quantized all-gather).
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.util import bits16, make_input

WORLD = 2
ROWS, COLS = 256, 384     # full tensor; each rank owns ROWS / WORLD rows


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _pad(a, shape):
    out = np.zeros(shape, np.uint8)
    out[: a.shape[0], : a.shape[1]] = a
    return torch.from_numpy(out)


def _mxfp8_storage(ops, xbits, shape):
    from oracle import oracle as O
    r = O.mxfp8_quantize(xbits)
    rs, cs = ops.mxfp8_scale_shapes(*shape)
    return ops.MXFP8TensorStorage(
        rowwise_data=torch.from_numpy(r["rowwise"].copy()), rowwise_scale_inv=_pad(r["rowwise_scale"], rs),
        columnwise_data=torch.from_numpy(r["colwise"].copy()), columnwise_scale_inv=_pad(r["colwise_scale"], cs),
        fp8_dtype=torch.float8_e4m3fn, shape=shape)


def _nvfp4_storage(ops, xbits, shape, amax_row, amax_col):
    from oracle import oracle as O
    qr, sr = O.nvfp4_quantize_1d(xbits, amax_row)
    qc, sc = O.nvfp4_quantize_1d(O.rht_transposed_bf16(xbits), amax_col)
    rs, cs = ops.nvfp4_scale_shapes(*shape)
    return ops.NVFP4TensorStorage(
        rowwise_data=torch.from_numpy(qr.copy()), rowwise_scale_inv=_pad(sr, rs),
        columnwise_data=torch.from_numpy(qc.copy()), columnwise_scale_inv=_pad(sc, cs),
        amax_rowwise=torch.tensor([amax_row]), amax_columnwise=torch.tensor([amax_col]), shape=shape)


def _worker(rank, port, errq):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=WORLD)
        from oracle import oracle as O
        from transformerengine_b200 import distributed as D
        from transformerengine_b200 import ops

        full = make_input(ROWS, COLS, "outliers", seed=7)
        n = ROWS // WORLD
        mine = full[rank * n:(rank + 1) * n].contiguous()
        fb, mb = bits16(full), bits16(mine)

        # --- amax all-reduce: local amaxes combine to the global one -----------------------
        local = torch.tensor([O.amax(mb), O.amax(O.rht_transposed_bf16(mb))])
        D.allreduce_amax(local, None)
        g_row, g_col = O.amax(fb), O.amax(O.rht_transposed_bf16(fb))
        assert local[0].item() == g_row and local[1].item() == g_col

        # --- plain tensor ---------------------------------------------------------------
        assert torch.equal(D.gather_along_first_dim(mine, None), full)

        # --- per-tensor FP8 (same scale everywhere after the amax reduction) -------------------
        scale = O.scale_from_amax(g_row)
        q_full = O.fp8_quantize(fb, scale)[0]
        q_mine = O.fp8_quantize(mb, scale)[0]
        st = ops.Float8TensorStorage(data=torch.from_numpy(q_mine.copy()), scale_inv=torch.tensor([1 / scale]),
                                     fp8_dtype=torch.float8_e4m3fn, shape=(n, COLS))
        g = D.gather_along_first_dim(st, None)
        assert g.shape == (ROWS, COLS) and np.array_equal(g.data.numpy(), q_full)

        # --- MXFP8: block scales are shard-local ------------------------------------------
        g = D.gather_along_first_dim(_mxfp8_storage(ops, mb, (n, COLS)), None)
        ref = _mxfp8_storage(ops, fb, (ROWS, COLS))
        for f in ("rowwise_data", "rowwise_scale_inv", "columnwise_data", "columnwise_scale_inv"):
            assert torch.equal(getattr(g, f).reshape(getattr(ref, f).shape), getattr(ref, f)), f
        assert g.shape == (ROWS, COLS)

        # --- NVFP4 + RHT: transposed column-wise payload gathers along dim 1 -------------------
        g = D.gather_along_first_dim(_nvfp4_storage(ops, mb, (n, COLS), g_row, g_col), None)
        ref = _nvfp4_storage(ops, fb, (ROWS, COLS), g_row, g_col)
        for f in ("rowwise_data", "rowwise_scale_inv", "columnwise_data", "columnwise_scale_inv"):
            assert torch.equal(getattr(g, f), getattr(ref, f)), f

        # --- reduce-scatter / all-reduce of partial results ---------------------------------
        part = torch.full((ROWS, 4), float(rank + 1))
        rs = D.reduce_scatter_along_first_dim(part, None)
        assert rs.shape == (n, 4) and torch.all(rs == 3.0)
        assert torch.all(D.allreduce(part.clone(), None) == 3.0)

        # --- swizzled scales must not be gathered ------------------------------------------
        bad = _mxfp8_storage(ops, mb, (n, COLS))
        bad.with_gemm_swizzled_scales = True
        try:
            D.gather_along_first_dim(bad, None)
            raise RuntimeError("gather of swizzled scales was not rejected")
        except AssertionError:
            pass
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        import traceback
        errq.put(f"rank {rank}: {e}\n{traceback.format_exc()}")
        raise


def test_tensor_parallel_gather_and_amax_reduction_gloo():
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, port, errq)) for r in range(WORLD)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    msgs = []
    while not errq.empty():
        msgs.append(errq.get())
    assert not msgs, "\n".join(msgs)
    assert all(p.exitcode == 0 for p in procs)


def test_recipes_and_quantizer_factories_on_cpu():
    """Host logic of the recipe surface (no GPU): defaults mirror the reference's."""
    from transformerengine_b200 import recipe
    d = recipe.DelayedScaling()
    assert d.margin == 0 and d.amax_history_len == 1024 and d.fp8_format == recipe.Format.HYBRID
    assert recipe.fp8_dtype_of(recipe.Format.HYBRID, True) == torch.float8_e4m3fn
    assert recipe.fp8_dtype_of(recipe.Format.HYBRID, False) == torch.float8_e5m2
    n = recipe.NVFP4BlockScaling()
    assert n.fp4_quant_fwd_inp.random_hadamard_transform and not n.fp4_quant_fwd_weight.random_hadamard_transform
    assert n.fp4_quant_fwd_weight.fp4_2d_quantization and n.fp4_quant_bwd_grad.stochastic_rounding
    with pytest.raises(AssertionError):
        recipe.MXFP8BlockScaling(fp8_format=recipe.Format.E5M2)
    from transformerengine_b200.quantization import FP8GlobalStateManager, autocast
    assert not FP8GlobalStateManager.is_fp8_enabled()
    with autocast(recipe=d):
        assert FP8GlobalStateManager.is_fp8_enabled() and FP8GlobalStateManager.get_fp8_recipe() is d
        assert FP8GlobalStateManager.is_first_fp8_module() and not FP8GlobalStateManager.is_first_fp8_module()
        with autocast(enabled=False):
            assert not FP8GlobalStateManager.is_fp8_enabled()
        assert FP8GlobalStateManager.is_fp8_enabled()
    assert not FP8GlobalStateManager.is_fp8_enabled() and FP8GlobalStateManager.AUTOCAST_DEPTH == 0


def test_grouped_linear_host_surface_on_cpu():
    """`te.GroupedLinear` host logic (no GPU): the reference's parameter names, split handling
    (ragged, empty) and the unquantized path outside autocast (reference grouped_linear.py:563-960)."""
    import transformerengine_b200 as te
    torch.manual_seed(0)
    g = te.GroupedLinear(3, 16, 8, bias=True, params_dtype=torch.float32, device="cpu")
    names = sorted(n for n, _ in g.named_parameters())
    assert names == ["bias0", "bias1", "bias2", "weight0", "weight1", "weight2"]
    with torch.no_grad():
        for i in range(3):
            getattr(g, f"bias{i}").normal_()
    x = torch.randn(10, 16, requires_grad=True)
    splits = [4, 0, 6]
    y = g(x, splits)
    assert y.shape == (10, 8)
    xs = torch.split(x, splits)
    ref = torch.cat([torch.nn.functional.linear(xs[i], getattr(g, f"weight{i}"), getattr(g, f"bias{i}"))
                     for i in range(3)])
    assert torch.equal(y, ref)
    y.sum().backward()
    assert g.weight1.grad is not None and not g.weight1.grad.any()   # the member without tokens
    assert torch.equal(g.weight0.grad, torch.ones(4, 8).t() @ xs[0].detach())
    with pytest.raises(AssertionError, match="Number of splits"):
        g(x, [10])


def test_symmetric_gather_layout_arithmetic():
    """Pure index arithmetic of `SymmetricFP8Gather`: piece buffers are rank-major, the GEMM's
    output row blocks put piece c of rank r back at row r * rows_local + c * rows_c."""
    world, rows_local, chunks = 4, 512, 2
    rows_c = rows_local // chunks
    seen = set()
    for c in range(chunks):
        size, stride, first = rows_c, rows_local, c * rows_c          # row_blocks(c)
        for r in range(world * rows_c):                               # row r of gathered piece c
            dst = (r // size) * stride + first + r % size
            rank, local = r // rows_c, c * rows_c + r % rows_c         # whose row, which local row
            assert dst == rank * rows_local + local
            seen.add(dst)
    assert seen == set(range(world * rows_local))
