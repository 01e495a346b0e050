"""`te.Linear` / `te.autocast` surface on the GPU: the README example, every recipe against
(a) the same pipeline composed by hand from the C-ABI operators (bit-identical) and (b) an fp32
torch reference (recipe-dependent tolerance), delayed-scaling bookkeeping against its closed form
(reference tests/pytorch/test_recipe.py:63-180), weight caching, backward overrides.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _te():
    import transformerengine_b200 as te
    from transformerengine_b200 import ops, recipe
    return te, ops, recipe


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


def _recipes(recipe):
    return {
        "delayed": (recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.E4M3), 0.06),
        "delayed_hybrid": (recipe.DelayedScaling(fp8_format=recipe.Format.HYBRID, amax_history_len=4), 0.12),
        "current": (recipe.Float8CurrentScaling(), 0.12),
        "mxfp8": (recipe.MXFP8BlockScaling(), 0.06),
        "nvfp4": (recipe.NVFP4BlockScaling(disable_stochastic_rounding=True), 0.25),
        "nvfp4_default": (recipe.NVFP4BlockScaling(), 0.3),   # stochastic rounding for gradients
        "nvfp4_plain": (recipe.NVFP4BlockScaling(disable_stochastic_rounding=True, disable_rht=True,
                                                 disable_2d_quantization=True), 0.25),
    }


def test_readme_example_runs():
    """reference README.rst:70-93, with `import transformerengine_b200 as te`."""
    te, _, recipe = _te()
    in_features, out_features, hidden = 768, 3072, 2048
    torch.manual_seed(0)
    model = te.Linear(in_features, out_features, bias=True)
    inp = torch.randn(hidden, in_features, device="cuda")
    fp8_recipe = recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.E4M3)
    for _ in range(3):
        with te.autocast(enabled=True, recipe=fp8_recipe):
            out = model(inp)
        loss = out.sum()
        loss.backward()
    torch.cuda.synchronize()
    assert out.shape == (hidden, out_features) and out.dtype == torch.float32
    ref = torch.nn.functional.linear(inp, model.weight, model.bias)
    assert _rel(out, ref) < 0.06
    assert model.weight.grad is not None and model.bias.grad is not None
    assert torch.isfinite(model.weight.grad).all()


@pytest.mark.parametrize("name", ["delayed", "delayed_hybrid", "current", "mxfp8", "nvfp4", "nvfp4_default",
                                  "nvfp4_plain"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_linear_matches_fp32_reference(name, dtype):
    te, _, recipe = _te()
    rec, tol = _recipes(recipe)[name]
    torch.manual_seed(1)
    T, fin, fout = 512, 256, 384
    lin = te.Linear(fin, fout, bias=True, params_dtype=dtype)
    with torch.no_grad():
        lin.bias.copy_(torch.randn(fout) * 0.1)
    x = torch.randn(T, fin, device="cuda", dtype=dtype, requires_grad=True)
    dy = torch.randn(T, fout, device="cuda", dtype=dtype) * 0.01
    steps = 3 if rec.delayed() else 1   # delayed scaling needs a populated amax history
    for _ in range(steps):
        lin.zero_grad()
        x.grad = None
        with te.autocast(recipe=rec):
            y = lin(x)
        y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    wr = lin.weight.detach().float().requires_grad_(True)
    br = lin.bias.detach().float().requires_grad_(True)
    yr = torch.nn.functional.linear(xr, wr, br)
    yr.backward(dy.float())
    assert y.dtype == dtype
    assert _rel(y, yr) < tol
    assert _rel(x.grad, xr.grad) < tol
    assert _rel(lin.weight.grad, wr.grad) < tol
    assert _rel(lin.bias.grad, br.grad) < 1e-2


def test_linear_equals_hand_composed_operators_mxfp8():
    te, ops, recipe = _te()
    torch.manual_seed(2)
    T, fin, fout = 256, 512, 640
    lin = te.Linear(fin, fout, bias=False, params_dtype=torch.bfloat16)
    x = torch.randn(T, fin, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    dy = torch.randn(T, fout, device="cuda", dtype=torch.bfloat16)
    with te.autocast(recipe=recipe.MXFP8BlockScaling()):
        y = lin(x)
    y.backward(dy)
    xq = ops.mxfp8_quantize(x.detach(), swizzled_scales=True)
    wq = ops.mxfp8_quantize(lin.weight.detach(), swizzled_scales=True)
    dq = ops.mxfp8_quantize(dy, swizzled_scales=True)
    assert torch.equal(y, ops.general_gemm(wq, xq, "TN"))
    assert torch.equal(x.grad, ops.general_gemm(wq, dq, "NN"))
    assert torch.equal(lin.weight.grad, ops.general_gemm(xq, dq, "NT"))


def test_delayed_scaling_bookkeeping_closed_form():
    te, _, recipe = _te()
    rec = recipe.DelayedScaling(margin=1, fp8_format=recipe.Format.HYBRID, amax_history_len=3)
    torch.manual_seed(3)
    lin = te.Linear(64, 128, bias=False, params_dtype=torch.bfloat16)
    amaxes_x, amaxes_dy = [], []
    for step in range(4):
        x = (torch.randn(32, 64, device="cuda") * (step + 1)).bfloat16().requires_grad_(True)
        dy = (torch.randn(32, 128, device="cuda") * 0.1 * (4 - step)).bfloat16()
        with te.autocast(recipe=rec):
            y = lin(x)
        y.backward(dy)
        amaxes_x.append(x.detach().abs().max().float().item())
        amaxes_dy.append(dy.abs().max().float().item())
        fwd, bwd = lin._delayed[True], lin._delayed[False]
        # the history rolls by -1 (reference delayed_scaling.cu:300-420): row 0 is cleared for the
        # next step, the last row holds this step's amax
        assert fwd.amax_history[0].abs().max().item() == 0.0
        assert fwd.amax_history[-1][0].item() == amaxes_x[-1]
        assert fwd.amax_history[-1][1].item() == lin.weight.detach().abs().max().float().item()
        assert bwd.amax_history[-1][0].item() == amaxes_dy[-1]
        # scale = fp8_max / max(last `amax_history_len` amaxes) / 2^margin ("max" algorithm)
        win_x = max(amaxes_x[-3:])
        win_dy = max(amaxes_dy[-3:])
        exp_x = torch.tensor(448.0 / 2.0, dtype=torch.float32) / torch.tensor(win_x, dtype=torch.float32)
        exp_dy = torch.tensor(57344.0 / 2.0, dtype=torch.float32) / torch.tensor(win_dy, dtype=torch.float32)
        assert fwd.scale[0].item() == exp_x.item()
        assert bwd.scale[0].item() == exp_dy.item()


def test_weight_cache_across_microbatches():
    te, _, recipe = _te()
    torch.manual_seed(4)
    lin = te.Linear(128, 256, bias=False, params_dtype=torch.bfloat16)
    x = torch.randn(64, 128, device="cuda", dtype=torch.bfloat16)
    rec = recipe.MXFP8BlockScaling()
    with te.autocast(recipe=rec):
        y0 = lin(x, is_first_microbatch=True)
        cached = lin._weight_q
        with torch.no_grad():
            lin.weight.add_(1.0)        # a stale cache must be what the second micro-batch uses
        y1 = lin(x, is_first_microbatch=False)
        assert lin._weight_q is cached
        y2 = lin(x)                      # None -> re-quantize
    assert torch.equal(y0, y1)
    assert not torch.equal(y0, y2)


@pytest.mark.parametrize("override", ["high_precision", "dequantized"])
def test_backward_override(override):
    te, _, recipe = _te()
    torch.manual_seed(5)
    lin = te.Linear(128, 256, bias=False, params_dtype=torch.bfloat16)
    x = torch.randn(64, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    dy = torch.randn(64, 256, device="cuda", dtype=torch.bfloat16)
    with te.autocast(recipe=recipe.MXFP8BlockScaling(backward_override=override)):
        y = lin(x)
    y.backward(dy)
    ref_dx = dy.float() @ lin.weight.detach().float()
    tol = 1e-2 if override == "high_precision" else 0.05
    assert _rel(x.grad, ref_dx) < tol


def test_disabled_autocast_is_plain_linear_and_sr_is_rejected():
    te, _, recipe = _te()
    lin = te.Linear(64, 64, bias=True, params_dtype=torch.bfloat16)
    x = torch.randn(32, 64, device="cuda", dtype=torch.bfloat16)
    with te.autocast(enabled=False):
        y = lin(x)
    assert torch.equal(y, torch.nn.functional.linear(x, lin.weight, lin.bias))


@pytest.mark.parametrize("name", ["delayed", "mxfp8", "current"])
def test_graphed_linear_matches_eager(name):
    """te.make_graphed_callables (reference pytorch/graph.py): the captured forward+backward of a
    Linear gives the eager path's outputs, gradients and delayed-scaling state, step after step."""
    te, _, recipe = _te()
    rec, _ = _recipes(recipe)[name]
    torch.manual_seed(3)
    eager = te.Linear(256, 512, bias=True, params_dtype=torch.bfloat16)
    graphed_mod = te.Linear(256, 512, bias=True, params_dtype=torch.bfloat16)
    graphed_mod.load_state_dict({k: v for k, v in eager.state_dict().items() if "extra_state" not in k}, strict=False)
    sample = torch.randn(384, 256, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    fn = te.make_graphed_callables(graphed_mod, (sample,), num_warmup_iters=3, enabled=True, recipe=rec)
    # warm-up / capture must leave no trace in the delayed-scaling state: both start fresh
    graphed_mod.zero_grad(set_to_none=True)
    for step in range(3):
        # as in a training loop: torch's graphed callables hand out their static gradient buffers
        eager.zero_grad(set_to_none=True)
        graphed_mod.zero_grad(set_to_none=True)
        xin = torch.randn(384, 256, device="cuda", dtype=torch.bfloat16) * (1.0 + step)
        xe, xg = xin.clone().requires_grad_(True), xin.clone().requires_grad_(True)
        with te.autocast(enabled=True, recipe=rec):
            ye = eager(xe)
        ye.backward(torch.ones_like(ye))
        yg = fn(xg)
        yg.backward(torch.ones_like(yg))
        torch.cuda.synchronize()
        assert torch.equal(ye, yg), f"step {step}: outputs differ"
        assert torch.equal(xe.grad, xg.grad), f"step {step}: dgrad differs"
        assert torch.equal(eager.weight.grad, graphed_mod.weight.grad), f"step {step}: wgrad differs"
        assert torch.equal(eager.bias.grad, graphed_mod.bias.grad), f"step {step}: bgrad differs"
    if name == "delayed":
        assert torch.equal(eager._delayed[True].scale, graphed_mod._delayed[True].scale)
        assert torch.equal(eager._delayed[False].amax_history, graphed_mod._delayed[False].amax_history)


@pytest.mark.parametrize("name", ["delayed", "mxfp8", "current"])
@pytest.mark.parametrize("layer", ["Linear", "LayerNormLinear"])
def test_return_bias_gradient_counted_once(name, layer):
    """`return_bias=True` (the Megatron setting): the caller adds the bias, so its gradient must come from
    that add alone (reference linear.py: `_Linear` gets `bias if apply_bias else None`)."""
    te, _, recipe = _te()
    rec, tol = _recipes(recipe)[name]
    torch.manual_seed(5)
    T, fin, fout = 256, 256, 384
    mod = getattr(te, layer)(fin, fout, bias=True, return_bias=True, params_dtype=torch.bfloat16)
    with torch.no_grad():
        mod.bias.copy_(torch.randn(fout) * 0.1)
    x = torch.randn(T, fin, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    dy = torch.randn(T, fout, device="cuda", dtype=torch.bfloat16) * 0.01
    for _ in range(3 if rec.delayed() else 1):
        mod.zero_grad()
        with te.autocast(recipe=rec):
            y, b = mod(x)
        assert b is mod.bias
        (y + b).backward(dy)
    ref = dy.float().sum(0)
    assert _rel(mod.bias.grad, ref) < 1e-2, "bias gradient is not dy.sum(0) (double counted?)"
    # and outside autocast
    mod.zero_grad()
    y, b = mod(x)
    (y + b).backward(dy)
    assert _rel(mod.bias.grad, ref) < 1e-2


def test_delayed_scaling_state_round_trips_through_state_dict_in_the_reference_format():
    """`_extra_state` is a uint8 tensor holding a pickled dict (reference module/base.py:1237-1355); loading it into
    a fresh module restores scales and amax histories, and the next step then produces the same bits."""
    te, _, recipe = _te()
    rec = recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.HYBRID, amax_history_len=4)
    torch.manual_seed(5)
    a = te.Linear(64, 128, bias=True, params_dtype=torch.bfloat16)
    xs = [(torch.randn(32, 64, device="cuda") * (i + 1)).bfloat16() for i in range(4)]
    dy = (torch.randn(32, 128, device="cuda") * 0.1).bfloat16()
    for x in xs[:3]:
        with te.autocast(recipe=rec):
            y = a(x.clone().requires_grad_(True))
        y.backward(dy)
    sd = a.state_dict()
    assert sd["_extra_state"].dtype == torch.uint8 and sd["_extra_state"].dim() == 1 and sd["_extra_state"].numel() > 0
    b = te.Linear(64, 128, bias=True, params_dtype=torch.bfloat16)
    b.load_state_dict(sd)
    for d in (True, False):
        assert torch.equal(a._delayed[d].scale, b._delayed[d].scale)
        assert torch.equal(a._delayed[d].amax_history, b._delayed[d].amax_history)
    outs = []
    for m in (a, b):
        m.zero_grad()
        xi = xs[3].clone().requires_grad_(True)
        with te.autocast(recipe=recipe.DelayedScaling(margin=0, fp8_format=recipe.Format.HYBRID, amax_history_len=4)):
            y = m(xi)
        y.backward(dy)
        outs.append((y.detach(), xi.grad, m.weight.grad))
    for u, v in zip(*outs):
        assert torch.equal(u, v)
    # a module that never ran FP8 stores the empty tensor, and loading it is a no-op
    c = te.Linear(64, 128, bias=False, params_dtype=torch.bfloat16)
    assert c.state_dict()["_extra_state"].numel() == 0
    te.Linear(64, 128, bias=False, params_dtype=torch.bfloat16).load_state_dict(c.state_dict())
