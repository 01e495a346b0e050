"""`te.GroupedLinear` (MoE experts) on the GPU: every recipe against (a) `num_gemms` separate
`te.Linear` layers holding the same weights (the reference's own criterion,
tests/pytorch/test_numerics.py `test_grouped_linear_accuracy`: outputs and gradients must agree)
and (b) the unquantized path; ragged and empty splits, bias, weight caching, fused wgrad accumulation.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _te():
    import transformerengine_b200 as te
    from transformerengine_b200 import recipe
    return te, recipe


def _recipes(recipe):
    return {
        "delayed": recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.HYBRID),
        "current": recipe.Float8CurrentScaling(),
        "mxfp8": recipe.MXFP8BlockScaling(),
        "nvfp4": recipe.NVFP4BlockScaling(disable_stochastic_rounding=True),
    }


def _pair(te, n, fin, fout, bias, dtype):
    torch.manual_seed(1)
    g = te.GroupedLinear(n, fin, fout, bias=bias, params_dtype=dtype)
    singles = [te.Linear(fin, fout, bias=bias, params_dtype=dtype) for _ in range(n)]
    with torch.no_grad():
        for i, s in enumerate(singles):
            s.weight.copy_(getattr(g, f"weight{i}"))
            if bias:
                getattr(g, f"bias{i}").normal_()
                s.bias.copy_(getattr(g, f"bias{i}"))
    return g, singles


@pytest.mark.parametrize("name", ["delayed", "current", "mxfp8", "nvfp4"])
@pytest.mark.parametrize("splits", [[256, 128, 384, 256], [512, 0, 128, 384], [64, 32, 896, 32]])
@pytest.mark.parametrize("bias", [True, False])
def test_grouped_linear_equals_separate_linears(name, splits, bias):
    te, recipe = _te()
    if name == "nvfp4" and any(s % 128 for s in splits):
        pytest.skip("NVFP4 + RHT columnwise needs 128-row members here")
    rec = _recipes(recipe)[name]
    n, fin, fout, dtype = len(splits), 512, 384, torch.bfloat16
    g, singles = _pair(te, n, fin, fout, bias, dtype)
    torch.manual_seed(2)
    x = torch.randn(sum(splits), fin, device="cuda", dtype=dtype, requires_grad=True)
    x2 = x.detach().clone().requires_grad_(True)
    dy = torch.randn(sum(splits), fout, device="cuda", dtype=dtype) * 0.1
    for _ in range(2):   # second pass: delayed scaling uses the scales the first one produced
        for mod in [g, *singles]:
            mod.zero_grad()
        x.grad = x2.grad = None
        with te.autocast(enabled=True, recipe=rec):
            y = g(x, splits)
        y.backward(dy)
        ys = []
        with te.autocast(enabled=True, recipe=rec):
            for xi, s, m in zip(torch.split(x2, splits), singles, splits):
                ys.append(s(xi) if m > 0 else xi.new_empty(0, fout))
        torch.cat(ys).backward(dy)
    torch.cuda.synchronize()
    # same quantizers, same kernels' arithmetic: bit-identical
    assert torch.equal(y, torch.cat(ys))
    assert torch.equal(x.grad, x2.grad)
    for i, s in enumerate(singles):
        wg = getattr(g, f"weight{i}").grad
        if splits[i] == 0:
            assert wg is not None and not wg.any()
            continue
        assert torch.equal(wg, s.weight.grad), f"wgrad {i}"
        if bias:
            assert torch.equal(getattr(g, f"bias{i}").grad, s.bias.grad), f"bgrad {i}"


def test_grouped_linear_close_to_unquantized_and_state_dict_names():
    te, recipe = _te()
    splits, fin, fout = [384, 128, 512], 256, 512
    g, _ = _pair(te, 3, fin, fout, True, torch.bfloat16)
    assert sorted(k for k in g.state_dict() if not k.endswith("_extra_state")) == \
        ["bias0", "bias1", "bias2", "weight0", "weight1", "weight2"]
    x = torch.randn(sum(splits), fin, device="cuda", dtype=torch.bfloat16)
    ref = g(x, splits)                      # outside autocast: plain torch
    with te.autocast(enabled=True, recipe=recipe.MXFP8BlockScaling()):
        out = g(x, splits)
    rel = ((out.float() - ref.float()).norm() / ref.float().norm()).item()
    assert rel < 0.06, rel


def test_grouped_linear_weight_cache_and_fused_wgrad_accumulation():
    te, recipe = _te()
    splits, fin, fout = [256, 256], 256, 256
    torch.manual_seed(5)
    g = te.GroupedLinear(2, fin, fout, bias=False, params_dtype=torch.bfloat16, fuse_wgrad_accumulation=True)
    for i in range(2):
        getattr(g, f"weight{i}").main_grad = torch.zeros(fout, fin, device="cuda", dtype=torch.float32)
    x = torch.randn(512, fin, device="cuda", dtype=torch.bfloat16)
    dy = torch.randn(512, fout, device="cuda", dtype=torch.bfloat16)
    rec = recipe.MXFP8BlockScaling()
    outs = []
    for first in (True, False):
        with te.autocast(enabled=True, recipe=rec):
            y = g(x, splits, is_first_microbatch=first)
        y.backward(dy)
        outs.append(y)
    assert torch.equal(outs[0], outs[1])
    one = getattr(g, "weight0").main_grad.clone()
    # two identical micro-batches accumulated: exactly twice one micro-batch
    g2 = te.GroupedLinear(2, fin, fout, bias=False, params_dtype=torch.bfloat16)
    with torch.no_grad():
        g2.weight0.copy_(g.weight0)
        g2.weight1.copy_(g.weight1)
    with te.autocast(enabled=True, recipe=rec):
        g2(x, splits).backward(dy)
    assert torch.allclose(one, 2 * g2.weight0.grad.float(), rtol=1e-2, atol=1e-2)


def test_grouped_linear_delayed_scaling_state_round_trips_through_state_dict():
    te, recipe = _te()
    rec = recipe.DelayedScaling(fp8_format=recipe.Format.HYBRID, amax_history_len=4)
    splits, fin, fout = [128, 256, 128], 256, 256
    g, _ = _pair(te, 3, fin, fout, False, torch.bfloat16)
    for i in range(3):
        x = (torch.randn(sum(splits), fin, device="cuda") * (i + 1)).bfloat16().requires_grad_(True)
        with te.autocast(enabled=True, recipe=rec):
            y = g(x, splits)
        y.backward(torch.ones_like(y) * 0.01)
    sd = g.state_dict()
    assert sd["_extra_state"].dtype == torch.uint8 and sd["_extra_state"].numel() > 0
    h, _ = _pair(te, 3, fin, fout, False, torch.bfloat16)
    h.load_state_dict(sd)
    assert len(h._delayed) == 3
    for da, db in zip(g._delayed, h._delayed):
        for d in (True, False):
            assert torch.equal(da[d].scale, db[d].scale) and torch.equal(da[d].amax_history, db[d].amax_history)
    # an equal recipe OBJECT keeps the state (it used to be rebuilt, losing the history)
    before = [d[True].amax_history.clone() for d in h._delayed]
    h._init_recipe_state(recipe.DelayedScaling(fp8_format=recipe.Format.HYBRID, amax_history_len=4))
    for b0, d in zip(before, h._delayed):
        assert torch.equal(b0, d[True].amax_history)
