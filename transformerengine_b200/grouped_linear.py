"""`GroupedLinear`: `num_gemms` independent Linears (MoE experts) applied to consecutive row
blocks of one input, y_i = x_i W_i^T + b_i, under `autocast(recipe=...)`.

Mirrors the reference's `te.GroupedLinear` (pytorch/module/grouped_linear.py: `_GroupedLinear`
forward :62-330, backward :332-560, module :563-960): parameters `weight{i}` / `bias{i}`,
`forward(inp, m_splits, is_first_microbatch=None)`. The reference casts each split
(`tex.split_quantize`) and runs `general_grouped_gemm`, which loops cuBLASLt over its compute
streams. Here each direction's list of GEMMs is ONE persistent launch (`nvte_multi_tensor_gemm`,
csrc/gemm.cu `run_grouped`): the members' tiles form one list, so small or ragged experts do not
leave SMs idle between launches. Outputs are row slices of one preallocated tensor.

As in the reference, block-scaled recipes need every `m_splits[i]` to be a multiple of 32 (the
wgrad contraction runs over the member's tokens); pad with the reference's `Fp8Padding` convention.
"""
from typing import List, Optional

import torch

from . import ops
from .linear import _for_gemm
from .quantization import DelayedScalingState, FP8GlobalStateManager, make_quantizers


class _GroupedLinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, m_splits, module, weights_q, is_grad_enabled, *params):
        m = module
        n = m.num_gemms
        weights, biases = params[:n], params[n:]
        in_features, out_features = m.in_features, m.out_features
        inputmat = inp.reshape(-1, in_features)
        assert sum(m_splits) == inputmat.shape[0], "m_splits must add up to the number of input rows"
        live = [i for i in range(n) if m_splits[i] > 0]
        needs_wgrad = is_grad_enabled and weights[0].requires_grad
        needs_dgrad = is_grad_enabled and inp.requires_grad

        xs = torch.split(inputmat, list(m_splits))
        x_q = {}
        for i in live:
            q = m.quantizers_fwd[i][0]
            q.set_usage(rowwise=True, columnwise=needs_wgrad)
            q.optimize_for_gemm = True
            x_q[i] = q(xs[i])
        if weights_q is None:
            weights_q = []
            for i in range(n):
                q = m.quantizers_fwd[i][1]
                q.set_usage(rowwise=True, columnwise=needs_dgrad)
                q.optimize_for_gemm = True
                weights_q.append(q(weights[i]))

        out = torch.empty(inputmat.shape[0], out_features, dtype=inp.dtype, device=inp.device)
        outs = torch.split(out, list(m_splits))
        if live:
            ops.multi_tensor_gemm([_for_gemm(weights_q[i]) for i in live], [_for_gemm(x_q[i]) for i in live], "TN",
                                  out=[outs[i] for i in live],
                                  bias=[biases[i] for i in live] if m.use_bias else None)
        if is_grad_enabled:
            ctx.module, ctx.m_splits, ctx.live = m, list(m_splits), live
            ctx.x_q, ctx.weights_q, ctx.weights = x_q, weights_q, weights
            ctx.inp_shape, ctx.inp_dtype = inp.shape, inp.dtype
            ctx.needs_dgrad, ctx.needs_wgrad = needs_dgrad, needs_wgrad
            ctx.reduce_and_update_bwd = m.recipe.delayed() and FP8GlobalStateManager.is_first_fp8_module()
            ctx.fp8_group = FP8GlobalStateManager.get_fp8_group()
        return out.reshape(*inp.shape[:-1], out_features)

    @staticmethod
    def backward(ctx, grad_output):
        m, n, live, m_splits = ctx.module, ctx.module.num_gemms, ctx.live, ctx.m_splits
        weights = ctx.weights
        dy = grad_output.reshape(-1, m.out_features).contiguous()
        dys = torch.split(dy, m_splits)
        dy_q, grad_biases = {}, [None] * n
        for i in live:
            q = m.quantizers_bwd[i][0]
            q.set_usage(rowwise=ctx.needs_dgrad, columnwise=ctx.needs_wgrad)
            q.optimize_for_gemm = True
            if m.use_bias:   # bias gradient out of the cast's own pass (nvte_quantize_dbias)
                dy_q[i], gb = q.quantize_with_dbias(dys[i])
                grad_biases[i] = gb.to(weights[i].dtype)
            else:
                dy_q[i] = q(dys[i])
            dy_q[i] = _for_gemm(dy_q[i])
        if m.use_bias:
            for i in range(n):
                if grad_biases[i] is None:
                    grad_biases[i] = torch.zeros_like(weights[i][:, 0])

        dgrad = None
        if ctx.needs_dgrad:
            # dx_i[m_i, in] = dy_i[m_i, out] . W_i[out, in]
            dgrad = torch.empty(dy.shape[0], m.in_features, dtype=ctx.inp_dtype, device=dy.device)
            dxs = torch.split(dgrad, m_splits)
            if live:
                ops.multi_tensor_gemm([_for_gemm(ctx.weights_q[i]) for i in live], [dy_q[i] for i in live], "NN",
                                      out=[dxs[i] for i in live])
            dgrad = dgrad.reshape(ctx.inp_shape)

        wgrads = [None] * n
        if ctx.needs_wgrad:
            # dW_i[out, in] = dy_i[m_i, out]^T . x_i[m_i, in]; members without tokens get zeros
            fuse = m.fuse_wgrad_accumulation
            dst = [getattr(w, "main_grad") if fuse else torch.empty_like(w) for w in weights]
            if live:
                ops.multi_tensor_gemm([_for_gemm(ctx.x_q[i]) for i in live], [dy_q[i] for i in live], "NT",
                                      out=[dst[i] for i in live], accumulate=fuse)
            if not fuse:
                for i in range(n):
                    if i not in dy_q:
                        dst[i].zero_()
                wgrads = dst
        if ctx.reduce_and_update_bwd:
            FP8GlobalStateManager.FP8_GROUP, saved = ctx.fp8_group, FP8GlobalStateManager.FP8_GROUP
            FP8GlobalStateManager.reduce_and_update_fp8_tensors(forward=False)
            FP8GlobalStateManager.FP8_GROUP = saved
        return (dgrad, None, None, None, None, *wgrads, *(grad_biases if m.use_bias else []))


class GroupedLinear(torch.nn.Module):
    """Drop-in for `te.GroupedLinear` on the low-precision path (reference
    module/grouped_linear.py:563-960): `num_gemms`, `in_features`, `out_features`, `bias`,
    `params_dtype`, `fuse_wgrad_accumulation`, `init_method`, `device`; parameters are named
    `weight0..` / `bias0..` as in the reference, so state dicts are interchangeable."""

    def __init__(self, num_gemms: int, in_features: int, out_features: int, bias: bool = True, *,
                 params_dtype: Optional[torch.dtype] = None, fuse_wgrad_accumulation: bool = False,
                 init_method=None, device="cuda"):
        super().__init__()
        params_dtype = torch.get_default_dtype() if params_dtype is None else params_dtype
        self.num_gemms, self.in_features, self.out_features = num_gemms, in_features, out_features
        self.use_bias, self.fuse_wgrad_accumulation = bias, fuse_wgrad_accumulation
        init = init_method or (lambda w: torch.nn.init.normal_(w, mean=0.0, std=0.023))
        for i in range(num_gemms):
            w = torch.nn.Parameter(torch.empty(out_features, in_features, dtype=params_dtype, device=device))
            init(w)
            self.register_parameter(f"weight{i}", w)
            if bias:
                self.register_parameter(f"bias{i}", torch.nn.Parameter(
                    torch.zeros(out_features, dtype=params_dtype, device=device)))
        self.recipe = None
        self.quantizers_fwd = self.quantizers_bwd = None
        self._delayed = []
        self._weights_q = None

    def _init_recipe_state(self, recipe) -> None:
        if self.recipe is recipe and self.quantizers_fwd is not None:
            return
        dev = self.weight0.device
        key = (recipe.amax_history_len, recipe.fp8_format) if recipe.delayed() else None
        if key is None:
            self._delayed, self._delayed_key = [], None
        elif getattr(self, "_delayed_key", None) != key or not self._delayed:
            # one scaling state per member and direction (reference: num_gemms * 3 / * 2 meta slots)
            self._delayed = [{True: DelayedScalingState(recipe, True, 3, dev),
                              False: DelayedScalingState(recipe, False, 2, dev)} for _ in range(self.num_gemms)]
            self._delayed_key = key
        else:   # an equal recipe object: keep scales and histories
            for d in self._delayed:
                d[True].recipe = d[False].recipe = recipe
        st = lambda i, fwd: self._delayed[i][fwd] if self._delayed else None
        self.quantizers_fwd = [make_quantizers(recipe, True, dev, delayed_state=st(i, True))
                               for i in range(self.num_gemms)]
        self.quantizers_bwd = [make_quantizers(recipe, False, dev, delayed_state=st(i, False))
                               for i in range(self.num_gemms)]
        self.recipe = recipe
        self._weights_q = None

    # -- checkpointing of the delayed-scaling state: the reference's format, the members' blocks concatenated
    # (reference module/base.py:1237-1355, grouped_linear.py `_offsets`) --
    def get_extra_state(self):
        from . import checkpoint
        return checkpoint.pack(self.recipe, self._delayed)

    def set_extra_state(self, state):
        from . import checkpoint
        st = checkpoint.unpack(state)
        self._pending_extra_state = st
        rec = st.get("recipe") if st else None
        if rec is not None and hasattr(rec, "delayed") and rec.delayed():
            self._init_recipe_state(rec)
            self._load_pending()

    def _load_pending(self):
        from . import checkpoint
        st = getattr(self, "_pending_extra_state", None)
        if st and self._delayed:
            checkpoint.load_into(st, self._delayed, type(self).__name__)
            self._pending_extra_state = None

    def forward(self, inp: torch.Tensor, m_splits: List[int], is_first_microbatch: Optional[bool] = None):
        assert len(m_splits) == self.num_gemms, "Number of splits should match number of GEMMs."
        weights = [getattr(self, f"weight{i}") for i in range(self.num_gemms)]
        biases = [getattr(self, f"bias{i}") for i in range(self.num_gemms)] if self.use_bias else []
        if not FP8GlobalStateManager.is_fp8_enabled():
            xs = torch.split(inp.reshape(-1, self.in_features), list(m_splits))
            outs = [torch.nn.functional.linear(x, w, biases[i] if self.use_bias else None)
                    for i, (x, w) in enumerate(zip(xs, weights))]
            return torch.cat(outs).reshape(*inp.shape[:-1], self.out_features)
        recipe = FP8GlobalStateManager.get_fp8_recipe()
        self._init_recipe_state(recipe)
        self._load_pending()
        grad_on = torch.is_grad_enabled()
        if recipe.delayed() and self.training and grad_on:
            for d in self._delayed:
                FP8GlobalStateManager.add_fp8_tensors_to_global_buffer(d[True])
                FP8GlobalStateManager.add_fp8_tensors_to_global_buffer(d[False])
        weights_q = None
        if is_first_microbatch is False and self._weights_q is not None:
            weights_q = self._weights_q
        elif is_first_microbatch is True:
            weights_q = []
            for i, w in enumerate(weights):
                q = self.quantizers_fwd[i][1]
                q.set_usage(rowwise=True, columnwise=True)
                q.optimize_for_gemm = True
                weights_q.append(q(w.detach()))
            self._weights_q = weights_q
        inp = inp.to(weights[0].dtype) if inp.dtype != weights[0].dtype else inp
        return _GroupedLinear.apply(inp, list(m_splits), self, weights_q, grad_on, *weights, *biases)
