"""torchrun worker: tensor-parallel `Linear` (column + sequence parallel, row + sequence parallel)
against the single-GPU module, per recipe. Launched by tests/test_tp_linear_gpu.py on >= 2 GPUs
(mirrors reference tests/pytorch/distributed/run_numerics.py:644-730)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import transformerengine_b200 as te  # noqa: E402
from transformerengine_b200 import recipe  # noqa: E402


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    group = dist.new_group(list(range(world)))
    T, fin, fout = 1024, 512, 768
    recipes = {
        "mxfp8": (recipe.MXFP8BlockScaling(), 0.0),
        "current": (recipe.Float8CurrentScaling(), 0.0),
        "nvfp4": (recipe.NVFP4BlockScaling(disable_stochastic_rounding=True), 0.0),
        "delayed": (recipe.DelayedScaling(fp8_format=recipe.Format.HYBRID, amax_history_len=4), 0.0),
    }
    for name, (rec, _) in recipes.items():
        torch.manual_seed(11)
        w = (torch.randn(fout, fin, device="cuda") * 0.05).bfloat16()
        x = torch.randn(T, fin, device="cuda").bfloat16()
        dy = (torch.randn(T, fout, device="cuda") * 0.01).bfloat16()
        for mode in ("column", "row"):
            ref = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16)
            tp = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16, parallel_mode=mode,
                           sequence_parallel=True, tp_group=group, ub_overlap_ag=(mode == "column"), symmetric_ag=False)
            n_t = T // world
            with torch.no_grad():
                ref.weight.copy_(w)
                if mode == "column":
                    n_o = fout // world
                    tp.weight.copy_(w[rank * n_o:(rank + 1) * n_o])
                else:
                    n_i = fin // world
                    tp.weight.copy_(w[:, rank * n_i:(rank + 1) * n_i])
            steps = 3 if rec.delayed() else 1
            for _ in range(steps):
                xr = x.clone().requires_grad_(True)
                ref.zero_grad()
                with te.autocast(recipe=rec):
                    yr = ref(xr)
                yr.backward(dy)
                tp.zero_grad()
                if mode == "column":
                    xs = x[rank * n_t:(rank + 1) * n_t].clone().requires_grad_(True)
                    with te.autocast(recipe=rec, amax_reduction_group=group):
                        ys = tp(xs)
                    ys.backward(dy[:, rank * n_o:(rank + 1) * n_o].contiguous())
                else:
                    xs = x[:, rank * n_i:(rank + 1) * n_i].clone().requires_grad_(True)
                    with te.autocast(recipe=rec, amax_reduction_group=group):
                        ys = tp(xs)
                    ys.backward(dy[rank * n_t:(rank + 1) * n_t].contiguous())
            # The TP module quantizes shards with their own amax (weights, grads), so it is compared
            # with the fp32 truth at the recipe's tolerance; MXFP8 block scales are shard-local, so
            # there fprop / wgrad must equal the single-GPU module bit for bit.
            tol = {"nvfp4": 0.3, "mxfp8": 0.06}.get(name, 0.12)
            y32 = torch.nn.functional.linear(x.float(), w.float())
            dx32 = dy.float() @ w.float()
            dw32 = dy.float().t() @ x.float()
            if mode == "column":
                y_ref, wg_ref, dx_ref = (yr[:, rank * n_o:(rank + 1) * n_o], ref.weight.grad[rank * n_o:(rank + 1) * n_o],
                                         xr.grad[rank * n_t:(rank + 1) * n_t])
                y_t, wg_t, dx_t = (y32[:, rank * n_o:(rank + 1) * n_o], dw32[rank * n_o:(rank + 1) * n_o],
                                   dx32[rank * n_t:(rank + 1) * n_t])
            else:
                y_ref, wg_ref, dx_ref = (yr[rank * n_t:(rank + 1) * n_t], ref.weight.grad[:, rank * n_i:(rank + 1) * n_i],
                                         xr.grad[:, rank * n_i:(rank + 1) * n_i])
                y_t, wg_t, dx_t = (y32[rank * n_t:(rank + 1) * n_t], dw32[:, rank * n_i:(rank + 1) * n_i],
                                   dx32[:, rank * n_i:(rank + 1) * n_i])
            assert ys.shape == y_ref.shape and xs.grad.shape == dx_ref.shape
            if mode == "column" and name in ("current", "delayed"):
                # per-tensor FP8: the input shard was gathered in pieces with the GEMM running behind
                # them; gathering first and running one GEMM must give the same bits
                plain = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16, parallel_mode=mode,
                                  sequence_parallel=True, tp_group=group, ub_overlap_ag=False, symmetric_ag=False)
                assert tp.overlap_ag_chunks > 1 and plain.overlap_ag_chunks == 1
                with torch.no_grad():
                    plain.weight.copy_(tp.weight)
                if rec.delayed():   # same scales as the module under test
                    plain._init_recipe_state(rec)
                    for d in (True, False):
                        plain._delayed[d].scale.copy_(tp._delayed[d].scale)
                        plain._delayed[d].amax_history.copy_(tp._delayed[d].amax_history)
                    with te.autocast(recipe=rec, amax_reduction_group=group), torch.no_grad():
                        y_over = tp(xs.detach())
                else:
                    y_over = ys
                with te.autocast(recipe=rec, amax_reduction_group=group), torch.no_grad():
                    y_plain = plain(xs.detach())
                assert torch.equal(y_over, y_plain), f"{name}/{mode}: chunked all-gather + GEMM differs"
                # the cast that pushes its bytes into every rank's gather buffer (symmetric memory,
                # no NCCL all-gather) must give the same bits too -- forward and backward, twice (the
                # two gather buffers alternate)
                symm = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16, parallel_mode=mode,
                                 sequence_parallel=True, tp_group=group, symmetric_ag=True,
                                 symmetric_ag_chunks=2 if name == "current" else 1)
                symm.symmetric_rs = False   # bit equality below: the dgrad reduce-scatter must be NCCL's on both sides
                with torch.no_grad():
                    symm.weight.copy_(tp.weight)
                for it in range(3):
                    if rec.delayed():
                        symm._init_recipe_state(rec)
                        plain._init_recipe_state(rec)
                        for d in (True, False):
                            for mod in (symm, plain):
                                mod._delayed[d].scale.copy_(tp._delayed[d].scale)
                                mod._delayed[d].amax_history.copy_(tp._delayed[d].amax_history)
                    outs = []
                    for mod in (symm, plain):
                        mod.zero_grad()
                        xi = (xs.detach() * (1.0 + 0.25 * it)).requires_grad_(True)
                        with te.autocast(recipe=rec, amax_reduction_group=group):
                            yi = mod(xi)
                        yi.backward(dy[:, rank * n_o:(rank + 1) * n_o].contiguous())
                        outs.append((yi.detach(), xi.grad, mod.weight.grad.clone()))
                    assert symm._symm is not None and symm._symm._step == it + 1
                    for a, b_ in zip(*outs):
                        assert torch.equal(a, b_), f"{name}/{mode}: symmetric cast+gather differs (iter {it})"
                if name == "current" and (T // world) % 256 == 0 and n_o > 128:
                    # the DEFAULT path (symmetric_ag=None -> auto): symmetric exchange, no collective and no
                    # barrier; the GEMM itself pushes the rank's rows to the peers and consumes the row chunks as
                    # their arrival counters come up ("gemm"); the older form where the cast kernel pushes and
                    # raises flags ("cast") stays covered. Same bits, forward and backward, over several steps
                    # (step counter / slot parity / buffer reuse), eager and replayed from a CUDA graph
                  for push in ("gemm", "cast"):
                    xchg = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16, parallel_mode=mode,
                                     sequence_parallel=True, tp_group=group)
                    xchg.symmetric_rs = False   # (as above)
                    if push == "cast":
                        from transformerengine_b200.distributed import SymmetricExchange
                        xchg._xch = SymmetricExchange(group, T // world, fin, x.device, push="cast")
                    with torch.no_grad():
                        xchg.weight.copy_(tp.weight)
                    for it in range(4):
                        outs = []
                        for mod in (xchg, plain):
                            mod.zero_grad()
                            xi = (xs.detach() * (1.0 + 0.5 * it) + 0.01 * rank).requires_grad_(True)
                            with te.autocast(recipe=rec, amax_reduction_group=group):
                                yi = mod(xi)
                            yi.backward(dy[:, rank * n_o:(rank + 1) * n_o].contiguous())
                            outs.append((yi.detach(), xi.grad, mod.weight.grad.clone()))
                        assert xchg._xch is not None and xchg._xch.push == push, "the symmetric exchange was not used"
                        assert xchg._xch.error_word.item() == 0, "a wait for a peer timed out"
                        assert xchg._xch.step_word.item() == it + 1
                        for a, b_ in zip(*outs):
                            assert torch.equal(a, b_), f"{name}/{mode}: symmetric exchange ({push}) differs (iter {it})"
                    if push == "gemm":   # modules of one shape share the gather buffers (one exchange, step after step)
                        from transformerengine_b200.distributed import SymmetricExchange as _SX
                        assert xchg._xch is _SX.shared(group, T // world, fin, x.device)
                    if rank == 0:
                        print(f"symmetric exchange push={push}: multicast={xchg._xch.multicast}", flush=True)
                    xg = (xs.detach() * 3.0).contiguous()
                    with te.autocast(recipe=rec, amax_reduction_group=group), torch.no_grad():
                        y_ref = plain(xg, is_first_microbatch=True)
                        xchg(xg, is_first_microbatch=True)
                        torch.cuda.synchronize()
                        dist.barrier()
                        g = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(g, capture_error_mode="thread_local"):
                            y_g = xchg(xg, is_first_microbatch=False)
                    for _ in range(3):
                        y_g.zero_()
                        g.replay()
                        torch.cuda.synchronize()
                        assert torch.equal(y_g, y_ref), f"{name}/{mode}: graph replay of the exchange ({push}) differs"
                    assert xchg._xch.error_word.item() == 0
            if name == "mxfp8" and mode == "column" and (T // world) % 256 == 0 and n_o > 128:
                # MXFP8 on the DEFAULT path: the shard is quantized into its window of the gather buffers and the GEMM
                # multicasts data + swizzled scales while it consumes the peers' chunks (SymmetricGatherMXFP8). Same bits
                # as gather-with-NCCL-then-GEMM, forward and backward, over several steps and from a CUDA graph
                mxd = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16, parallel_mode=mode,
                                sequence_parallel=True, tp_group=group)
                mxd.symmetric_rs = False   # (as above)
                mxn = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16, parallel_mode=mode,
                                sequence_parallel=True, tp_group=group, symmetric_ag=False)
                with torch.no_grad():
                    mxd.weight.copy_(tp.weight)
                    mxn.weight.copy_(tp.weight)
                for it in range(4):
                    outs = []
                    for mod in (mxd, mxn):
                        mod.zero_grad()
                        xi = (xs.detach() * (1.0 + 0.5 * it) + 0.01 * rank).requires_grad_(True)
                        with te.autocast(recipe=rec, amax_reduction_group=group):
                            yi = mod(xi)
                        yi.backward(dy[:, rank * n_o:(rank + 1) * n_o].contiguous())
                        outs.append((yi.detach(), xi.grad, mod.weight.grad.clone()))
                    used = mxd._mxg is not None and mxd._mxg[1] is not None
                    if not used:
                        break   # no NVSwitch multicast on this system: NCCL path, nothing to compare
                    assert mxd._mxg[1].error_word.item() == 0 and mxd._mxg[1].step_word.item() == it + 1
                    for a, b_ in zip(*outs):
                        assert torch.equal(a, b_), f"{name}/{mode}: in-GEMM MXFP8 gather differs (iter {it})"
                if used:
                    xg = (xs.detach() * 3.0).contiguous()
                    with te.autocast(recipe=rec, amax_reduction_group=group), torch.no_grad():
                        y_ref2 = mxn(xg, is_first_microbatch=True)
                        mxd(xg, is_first_microbatch=True)
                        torch.cuda.synchronize()
                        dist.barrier()
                        g = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(g, capture_error_mode="thread_local"):
                            y_g = mxd(xg, is_first_microbatch=False)
                    for _ in range(3):
                        y_g.zero_()
                        g.replay()
                        torch.cuda.synchronize()
                        assert torch.equal(y_g, y_ref2), f"{name}/{mode}: graph replay of the MXFP8 gather differs"
                    assert mxd._mxg[1].error_word.item() == 0
                if rank == 0:
                    print(f"in-GEMM MXFP8 all-gather: {'bit-equal to NCCL gather + GEMM' if used else 'not available'}", flush=True)
            if name == "mxfp8":
                assert torch.equal(tp.weight.grad, wg_ref), f"{name}/{mode}: wgrad differs"
                if mode == "column":
                    assert torch.equal(ys, y_ref), f"{name}/{mode}: fprop differs"
                else:
                    assert torch.equal(xs.grad, dx_ref), f"{name}/{mode}: dgrad differs"
            e_y, e_w, e_x = rel(ys, y_t), rel(tp.weight.grad, wg_t), rel(xs.grad, dx_t)
            assert e_y < tol and e_w < tol and e_x < tol, (name, mode, e_y, e_w, e_x)
            assert rel(yr, y32) < tol
            if rank == 0:
                print(f"{name:8s} {mode:6s} rel err vs fp32 truth: y {e_y:.2e} wgrad {e_w:.2e} dgrad {e_x:.2e}",
                      flush=True)
    # fused GEMM + reduce-scatter (SymmetricReduceScatter: the epilogue scatters partial rows over NVLink, the owner
    # sums them in rank order) against GEMM -> NCCL reduce-scatter: row-parallel fprop and column-parallel dgrad.
    # NCCL's ring rounds the running sum to bf16 at each of its world - 1 hops (2^-9 of the sum each), ours rounds once:
    # tolerance world x 2^-8 of the largest element (twice that bound); several calls, so both buffer sets are used
    from transformerengine_b200.distributed import SymmetricReduceScatter
    for mode, rs_rec in (("row", recipe.MXFP8BlockScaling()), ("column", recipe.MXFP8BlockScaling()),
                         ("row", recipe.Float8CurrentScaling()), ("column", recipe.Float8CurrentScaling())):
        # (per-tensor FP8 GEMMs scatter 128-byte row pieces through shared memory, the block-scaled kinds store directly)
        torch.manual_seed(21)
        w = (torch.randn(fout, fin, device="cuda") * 0.05).bfloat16()
        x = torch.randn(T, fin, device="cuda").bfloat16()
        dy = (torch.randn(T, fout, device="cuda") * 0.01).bfloat16()
        n_t, n_i, n_o = T // world, fin // world, fout // world
        mods = []
        for fused in (True, False):
            mod = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16, parallel_mode=mode,
                            sequence_parallel=True, tp_group=group, symmetric_ag=False)
            mod.symmetric_rs = fused
            with torch.no_grad():
                mod.weight.copy_(w[:, rank * n_i:(rank + 1) * n_i] if mode == "row" else w[rank * n_o:(rank + 1) * n_o])
            mods.append(mod)
        for it in range(3):
            res = []
            for mod in mods:
                mod.zero_grad()
                if mode == "row":
                    xi = (x[:, rank * n_i:(rank + 1) * n_i] * (1.0 + it)).clone().requires_grad_(True)
                    gi = dy[rank * n_t:(rank + 1) * n_t].contiguous()
                else:
                    xi = (x[rank * n_t:(rank + 1) * n_t] * (1.0 + it)).clone().requires_grad_(True)
                    gi = dy[:, rank * n_o:(rank + 1) * n_o].contiguous()
                with te.autocast(recipe=rs_rec, amax_reduction_group=group):
                    yi = mod(xi)
                yi.backward(gi)
                res.append((yi.detach().float(), xi.grad.float()))
            assert any(isinstance(v, SymmetricReduceScatter) for v in mods[0]._rs.values()), "fused reduce-scatter not used"
            assert not mods[1]._rs
            for name_, a, b_ in zip(("fprop", "dgrad"), res[0], res[1]):
                err = ((a - b_).abs().max() / b_.abs().max()).item()
                assert err <= world * 2.0 ** -8, f"{mode}: fused reduce-scatter {name_} differs from NCCL by {err}"
        if rank == 0:
            print(f"fused GEMM + reduce-scatter ({mode}-parallel, {type(rs_rec).__name__}): matches NCCL within "
                  "one bf16 rounding", flush=True)
    # row-parallel with bias: the bias is added after the reduction, outside `_Linear`, so its gradient
    # is the plain column sum of this rank's grad_output, counted once (with and without SP)
    for sp in (True, False):
        torch.manual_seed(12)
        n_i, n_t = fin // world, T // world
        tpb = te.Linear(fin, fout, bias=True, params_dtype=torch.bfloat16, parallel_mode="row",
                        sequence_parallel=sp, tp_group=group)
        x = torch.randn(T, fin, device="cuda").bfloat16()
        dy = (torch.randn(T, fout, device="cuda") * 0.01).bfloat16()
        xs = x[:, rank * n_i:(rank + 1) * n_i].clone().requires_grad_(True)
        with te.autocast(recipe=recipe.MXFP8BlockScaling(), amax_reduction_group=group):
            ys = tpb(xs)
        dy_l = dy[rank * n_t:(rank + 1) * n_t].contiguous() if sp else dy
        assert ys.shape == dy_l.shape
        ys.backward(dy_l)
        e_b = rel(tpb.bias.grad, dy_l.float().sum(0))
        assert e_b < 1e-2, f"row-parallel bias gradient (sp={sp}) off by {e_b}"
    dist.barrier()
    if rank == 0:
        print("TP LINEAR OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
