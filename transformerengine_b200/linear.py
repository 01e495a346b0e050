"""`Linear`: y = x W^T + b with quantized GEMMs under `autocast(recipe=...)`.

Mirrors the reference's `te.Linear` (pytorch/module/linear.py: `_linear_forward_impl` :252-725,
`_linear_backward` :727-1288, `_Linear` :1290, `Linear` :1374-2120) for the part of it that is
the low-precision hot path: quantize input / weight / grad_output with the recipe's quantizers,
fprop = TN GEMM, dgrad = NN GEMM, wgrad = NT GEMM, bias in the fprop epilogue, delayed-scaling
bookkeeping, weight-cast caching across micro-batches, and column / row tensor parallelism with
sequence parallelism (quantize the local shard -> all-gather the low-precision payload).
Every device operation is an `nvte_*` call (`ops`); outside `autocast` the layer is a plain
`torch.nn.functional.linear`, which is not part of this path.
"""
from typing import Optional

import os

import torch

from . import ops
from .distributed import (ChunkedFP8Gather, SymmetricExchange, SymmetricFP8Gather, SymmetricGatherMXFP8,
                          SymmetricReduceScatter, allreduce,
                          gather_along_first_dim,
                          get_distributed_world_size, reduce_scatter_along_first_dim)
from .quantizers import Float8CurrentScalingQuantizer, Float8Quantizer
from .quantization import DelayedScalingState, FP8GlobalStateManager, make_quantizers

_BLOCK_SCALED = (ops.MXFP8TensorStorage, ops.NVFP4TensorStorage)


def _for_gemm(t):
    """Block scales must be in the tcgen05 layout; quantizers write it directly unless the
    tensor had to be gathered first (then swizzle the gathered compact scales here)."""
    if isinstance(t, _BLOCK_SCALED) and not t.with_gemm_swizzled_scales:
        return ops.swizzle_scales(t)
    return t


def _dequantized(t, dtype, columnwise=False):
    return ops.dequantize(t, dtype=dtype, columnwise=columnwise)


class _Linear(torch.autograd.Function):
    """Quantized forward / backward of one Linear (reference module/linear.py:1290-1372)."""

    @staticmethod
    def forward(ctx, inp, weight, bias, module, weight_q, is_grad_enabled, ln_weight=None, ln_bias=None):
        m = module
        out_features, in_features = weight.shape
        inputmat = inp.reshape(-1, in_features)
        tp_group, tp_size = m.tp_group, m.tp_size
        sp_gather_in = m.parallel_mode == "column" and m.sequence_parallel and tp_size > 1
        q_in, q_w, _ = m.quantizers_fwd
        backward_needs_input = is_grad_enabled and weight.requires_grad
        needs_dgrad = inp.requires_grad or (ln_weight is not None and ln_weight.requires_grad)
        override = m.recipe.backward_override

        # ---- input cast (reference linear.py:301-364) ----
        q_in.set_usage(rowwise=True, columnwise=backward_needs_input and override is None)
        q_in.optimize_for_gemm = not sp_gather_in
        ln_out = mu = rsigma = x_q = symm_ag = arrival = None
        if ln_weight is not None:
            # LayerNormLinear: the norm kernel writes the quantized GEMM operand itself
            # (reference layernorm_linear.py:190-260 `apply_normalization` with the input quantizer)
            from .normalization import _norm_fwd
            x_q_local, ln_out, mu, rsigma = _norm_fwd(
                m.normalization == "LayerNorm", inputmat, ln_weight, ln_bias, m.eps, q_in, m.zero_centered_gamma,
                want_high_precision=m.return_layernorm_output or override == "high_precision")
        elif sp_gather_in and (xch := m._exchange_for(inputmat, q_in)) is not None:
            # current scaling: amax exchange, cast and gather are flag-carrying stores over NVLink and the GEMM
            # consumes the row chunks as their flags come up -- no collective, no barrier (SymmetricExchange)
            # (the local shard is copied out of the gather buffer after the GEMM, off its critical path)
            if weight_q is None:
                q_w.set_usage(rowwise=True, columnwise=is_grad_enabled and needs_dgrad and override is None)
                q_w.optimize_for_gemm = True
                weight_q = q_w(weight)
            x_q = xch.cast_and_gather(inputmat.contiguous(), q_in)
            x_q_local = None
            arrival = xch.arrival
        elif sp_gather_in and override is None and (mxg := m._mx_gather_for(inputmat, q_in)) is not None:
            # MXFP8: the shard is quantized into its window of the gather buffers (data + swizzled scales) and the
            # GEMM multicasts both while it consumes the peers' chunks (SymmetricGatherMXFP8)
            x_q, x_q_local = mxg.cast(inputmat.contiguous(), q_in.dtype,
                                      columnwise=backward_needs_input and override is None)
            arrival = mxg.arrival
            xch = None
        elif (sp_gather_in and m._symmetric_wanted() and isinstance(q_in, (Float8Quantizer, Float8CurrentScalingQuantizer))
              and inputmat.dtype in (torch.bfloat16, torch.float16) and inputmat.numel() % 16 == 0):
            # the cast writes straight into every rank's gather buffers over NVLink (no NCCL all-gather),
            # in row pieces, and the GEMM over piece c runs while piece c + 1 is on the wire
            symm_ag = m._symmetric_cast_and_gather(inputmat.contiguous(), q_in)
            x_q_local = symm_ag.local_copy() if backward_needs_input else None
        else:
            x_q_local = q_in(inputmat)
        # sequence-parallel input: all-gather the low-precision shard. Per-tensor FP8 shards are
        # gathered in pieces on a side stream and the GEMM runs piece by piece behind them
        # (reference `ub_overlap_ag`); other formats gather first.
        overlap = (sp_gather_in and m.overlap_ag_chunks > 1 and isinstance(x_q_local, ops.Float8TensorStorage)
                   and x_q_local.data.shape[0] % 128 == 0) or symm_ag is not None
        if not overlap and arrival is None:
            x_q = gather_along_first_dim(x_q_local, tp_group) if sp_gather_in else x_q_local

        # ---- weight cast, cached across micro-batches (reference linear.py:366-420) ----
        if weight_q is None:
            q_w.set_usage(rowwise=True, columnwise=is_grad_enabled and needs_dgrad and override is None)
            q_w.optimize_for_gemm = True
            weight_q = q_w(weight)

        # ---- fprop GEMM: y[T, out] = x[T, in] . W[out, in]^T (+ bias) ----
        gemm_bias = bias if (bias is not None and m.apply_bias_in_gemm) else None
        rs = None
        if overlap:
            ag = symm_ag or ChunkedFP8Gather(x_q_local, tp_group, m.overlap_ag_chunks, m._ag_cache).start()
            out = torch.empty(ag.rows_local * tp_size, out_features, dtype=inp.dtype, device=inp.device)
            for c in range(ag.chunks):
                # leave SMs to the NCCL kernel of the next piece (the GEMM grid is persistent and would
                # otherwise hold every SM); the last GEMM has nothing left to overlap with. The
                # symmetric-memory cast needs no such margin: its CTAs fit next to the GEMM's
                ops.general_gemm(weight_q, ag.piece(c), "TN", out=out, bias=gemm_bias,
                                 out_row_blocks=ag.row_blocks(c) if ag.chunks > 1 else None,
                                 sm_count=m.overlap_ag_gemm_sms if (c + 1 < ag.chunks and symm_ag is None) else 0)
        elif (rs := m._reduce_scatter_for("fwd", inputmat.shape[0], out_features, inp.dtype, inp.device,
                                          m.parallel_mode == "row" and gemm_bias is None,
                                          isinstance(q_in, (Float8Quantizer, Float8CurrentScalingQuantizer)))) is not None:
            # row-parallel + sequence-parallel: the GEMM's epilogue scatters the partial rows to their owner ranks
            # over NVLink, a barrier, then the owner sums its `tp` partial slabs (SymmetricReduceScatter)
            out = rs.gemm(_for_gemm(weight_q), _for_gemm(x_q), "TN")
        else:
            out = ops.general_gemm(_for_gemm(weight_q), _for_gemm(x_q), "TN", out_dtype=inp.dtype, bias=gemm_bias,
                                   arrival=arrival)
            if arrival is not None and xch is not None and backward_needs_input:
                x_q_local = xch.local_copy()

        # ---- row-parallel output reduction (reference linear.py:560-600) ----
        if m.parallel_mode == "row" and tp_size > 1 and rs is None:
            out = (reduce_scatter_along_first_dim(out, tp_group) if m.sequence_parallel
                   else allreduce(out, tp_group))

        if is_grad_enabled:
            if override == "high_precision":
                ctx.saved_input, ctx.saved_weight = (inputmat if ln_weight is None else ln_out), weight
            elif override == "dequantized":
                ctx.saved_input, ctx.saved_weight = x_q_local, weight_q
            else:
                ctx.saved_input, ctx.saved_weight = x_q_local, weight_q
            ctx.module, ctx.weight, ctx.use_bias = m, weight, bias is not None
            ctx.inp_shape, ctx.inp_dtype = inp.shape, inp.dtype
            ctx.inp_requires_grad = needs_dgrad
            ctx.sp_gather_in = sp_gather_in
            ctx.reduce_and_update_bwd = (m.recipe.delayed() and
                                         FP8GlobalStateManager.is_first_fp8_module())
            ctx.fp8_group = FP8GlobalStateManager.get_fp8_group()
            ctx.ln = ln_weight is not None
            if ctx.ln:
                ctx.ln_saved = (inputmat, ln_weight, mu, rsigma)
                ctx.ln_bias_given = ln_bias is not None
        if ln_weight is not None and m.return_layernorm_output:
            out = out.reshape(-1, *inp.shape[1:-1], out_features) if inp.dim() > 2 else out
            return out, ln_out.reshape(inp.shape)
        # leading dim changes under sequence parallelism (gathered / reduce-scattered tokens)
        return out.reshape(-1, *inp.shape[1:-1], out_features) if inp.dim() > 2 else out

    @staticmethod
    def backward(ctx, grad_output, grad_ln_out=None):
        m, weight = ctx.module, ctx.weight
        tp_group, tp_size = m.tp_group, m.tp_size
        out_features, in_features = weight.shape
        override = m.recipe.backward_override
        dy = grad_output.reshape(-1, out_features).contiguous()
        # with a quantized backward the bias gradient comes out of the cast's own pass over dy
        # (nvte_quantize_dbias, reference module/base.py:1648 `tex.bgrad_quantize`)
        grad_bias = dy.sum(dim=0).to(weight.dtype) if (ctx.use_bias and override is not None) else None
        sp_gather_dy = m.parallel_mode == "row" and m.sequence_parallel and tp_size > 1
        dgrad = wgrad = None

        if override is not None:
            # reference `backward_override`: backward GEMMs in high precision, on the saved
            # high-precision tensors or on dequantised copies of the forward operands
            if sp_gather_dy:
                dy = gather_along_first_dim(dy, tp_group)
            if override == "high_precision":
                x, w = ctx.saved_input, ctx.saved_weight
            else:
                x = _dequantized(ctx.saved_input, ctx.inp_dtype)
                w = _dequantized(ctx.saved_weight, weight.dtype)
            if ctx.sp_gather_in:
                x = gather_along_first_dim(x, tp_group)
            if ctx.inp_requires_grad:
                dgrad = dy.to(w.dtype) @ w
            if weight.requires_grad:
                wgrad = dy.t().to(x.dtype) @ x
        else:
            q_dy = m.quantizers_bwd[0]
            q_dy.set_usage(rowwise=ctx.inp_requires_grad, columnwise=weight.requires_grad)
            q_dy.optimize_for_gemm = not sp_gather_dy
            if ctx.use_bias:
                dy_q, grad_bias = q_dy.quantize_with_dbias(dy)
                grad_bias = grad_bias.to(weight.dtype)
            else:
                dy_q = q_dy(dy)
            if sp_gather_dy:
                dy_q = gather_along_first_dim(dy_q, tp_group)
            dy_q = _for_gemm(dy_q)
            rs = None
            if ctx.inp_requires_grad:
                # dgrad[T, in] = dy[T, out] . W[out, in]; column-parallel + sequence-parallel: the reduce-scatter of
                # dgrad starts in the GEMM's epilogue (SymmetricReduceScatter)
                rs = m._reduce_scatter_for("dgrad", dy.shape[0], weight.shape[1], ctx.inp_dtype, dy.device,
                                           m.parallel_mode == "column",
                                           isinstance(q_dy, (Float8Quantizer, Float8CurrentScalingQuantizer)))
                if rs is not None:
                    dgrad = rs.gemm(_for_gemm(ctx.saved_weight), dy_q, "NN")
                else:
                    dgrad = ops.general_gemm(_for_gemm(ctx.saved_weight), dy_q, "NN",
                                             out_dtype=ctx.inp_dtype)
            if weight.requires_grad:
                x_q = ctx.saved_input
                if ctx.sp_gather_in:  # re-gather the column-wise payload (reference linear.py:880-930)
                    if isinstance(x_q, _BLOCK_SCALED) and x_q.with_gemm_swizzled_scales:
                        x_q = ops.swizzle_scales(x_q, inverse=True)   # the gather concatenates compact scales
                    x_q = gather_along_first_dim(x_q, tp_group)
                # wgrad[out, in] = dy[T, out]^T . x[T, in]
                main_grad = getattr(weight, "main_grad", None) if m.fuse_wgrad_accumulation else None
                wgrad = ops.general_gemm(_for_gemm(x_q), dy_q, "NT", out_dtype=weight.dtype,
                                         out=main_grad, accumulate=main_grad is not None)
                if main_grad is not None:
                    wgrad = None

        if dgrad is not None and m.parallel_mode == "column" and tp_size > 1 and (override is not None or rs is None):
            dgrad = (reduce_scatter_along_first_dim(dgrad, tp_group) if m.sequence_parallel
                     else allreduce(dgrad, tp_group))
        dgamma = dbeta = None
        if ctx.ln and dgrad is not None:
            # backward of the normalization (reference layernorm_linear.py:1010-1060)
            x, ln_weight, mu, rsigma = ctx.ln_saved
            dz = dgrad.reshape(x.shape)
            if grad_ln_out is not None:
                dz = dz + grad_ln_out.reshape(x.shape)
            dgrad, dgamma, dbeta = ops.norm_bwd(dz, x, mu, rsigma, ln_weight, m.zero_centered_gamma)
        if dgrad is not None:
            dgrad = dgrad.reshape(ctx.inp_shape)
        if ctx.reduce_and_update_bwd:
            FP8GlobalStateManager.FP8_GROUP, saved = ctx.fp8_group, FP8GlobalStateManager.FP8_GROUP
            FP8GlobalStateManager.reduce_and_update_fp8_tensors(forward=False)
            FP8GlobalStateManager.FP8_GROUP = saved
        return dgrad, wgrad, grad_bias, None, None, None, dgamma, dbeta


class Linear(torch.nn.Module):
    """Drop-in for `te.Linear` on the low-precision path (reference module/linear.py:1374-1713).

    Parameters keep the reference's names and meaning: `in_features`, `out_features`, `bias`,
    `params_dtype`, `parallel_mode` (None | "column" | "row"), `sequence_parallel`, `tp_group`,
    `tp_size`, `fuse_wgrad_accumulation`, `return_bias`, `device`, `init_method`.
    `forward(inp, is_first_microbatch=None)` caches the quantized weight when
    `is_first_microbatch` is True and reuses it when False.
    """

    def __init__(self, in_features: int, out_features: int, bias: bool = True, *,
                 params_dtype: Optional[torch.dtype] = None, parallel_mode: Optional[str] = None,
                 sequence_parallel: bool = False, tp_group=None, tp_size: int = 1,
                 fuse_wgrad_accumulation: bool = False, return_bias: bool = False,
                 init_method=None, device="cuda", ub_overlap_ag: bool = False, overlap_ag_chunks: int = 4,
                 overlap_ag_sm_margin: int = 16, symmetric_ag: Optional[bool] = None, symmetric_ag_chunks: int = 1):
        super().__init__()
        # `symmetric_ag` (extension, per-tensor FP8 recipes, column-parallel + sequence-parallel): the input
        # cast stores its bytes into every rank's gather buffer through symmetric memory, replacing the
        # NCCL all-gather by one barrier (distributed.SymmetricFP8Gather)
        # `symmetric_ag=None` (default): ON whenever the symmetric-memory rendezvous over the TP group works (one
        # NVLink / NVSwitch domain, NCCL backend) and the recipe is per-tensor FP8; NCCL stays the fallback (and the
        # gloo path of the CPU tests). Under current scaling the flag-driven exchange is used (SymmetricExchange:
        # no barrier at all, the GEMM consumes row chunks as they arrive); `symmetric_ag_flags=False` selects the
        # round-1 cast + barrier path, which delayed scaling always uses.
        self.symmetric_ag, self.symmetric_ag_chunks = symmetric_ag, symmetric_ag_chunks
        self.symmetric_ag_flags = True
        self._symm = self._xch = self._mxg = None
        self._symm_failed = False
        # `symmetric_rs` (None = auto, like symmetric_ag): with sequence parallelism the reduce-scatter of the
        # row-parallel fprop output and of the column-parallel dgrad starts in the GEMM's epilogue
        # (distributed.SymmetricReduceScatter); NCCL reduce-scatter otherwise
        self.symmetric_rs = None
        self._rs = {}
        # `ub_overlap_ag` (off by default, as in the reference): column-parallel + sequence-parallel fprop
        # gathers the per-tensor FP8 input shard in this many pieces on a side stream and runs the GEMM
        # piece by piece behind them; otherwise gather first, one GEMM
        self.overlap_ag_chunks = overlap_ag_chunks if ub_overlap_ag else 1
        self.overlap_ag_gemm_sms = max(2, ops.sm_count() - overlap_ag_sm_margin)
        self._ag_cache = {}
        assert parallel_mode in (None, "column", "row"), f"parallel_mode {parallel_mode} not supported"
        params_dtype = torch.get_default_dtype() if params_dtype is None else params_dtype
        self.tp_group = tp_group
        self.tp_size = get_distributed_world_size(tp_group) if tp_group is not None else tp_size
        self.parallel_mode, self.sequence_parallel = parallel_mode, sequence_parallel
        self.in_features, self.out_features = in_features, out_features
        if parallel_mode == "column":
            assert out_features % self.tp_size == 0
            self.out_features = out_features // self.tp_size
        elif parallel_mode == "row":
            assert in_features % self.tp_size == 0
            self.in_features = in_features // self.tp_size
        self.use_bias, self.return_bias = bias, return_bias
        self.fuse_wgrad_accumulation = fuse_wgrad_accumulation
        # row-parallel bias is added after the reduction (reference linear.py:1620-1640)
        self.apply_bias_in_gemm = bias and not return_bias and not (parallel_mode == "row" and self.tp_size > 1)
        self.weight = torch.nn.Parameter(
            torch.empty(self.out_features, self.in_features, dtype=params_dtype, device=device))
        (init_method or (lambda w: torch.nn.init.normal_(w, mean=0.0, std=0.023)))(self.weight)
        if bias:
            self.bias = torch.nn.Parameter(torch.zeros(self.out_features, dtype=params_dtype, device=device))
        else:
            self.register_parameter("bias", None)
        self.recipe = None
        self.quantizers_fwd = self.quantizers_bwd = None
        self._delayed = {}
        self._weight_q = None

    # -- recipe state (reference module/base.py `init_fp8_metadata`, :1090-1180) --
    def _init_recipe_state(self, recipe) -> None:
        if self.recipe is recipe and self.quantizers_fwd is not None:
            return
        dev = self.weight.device
        if recipe.delayed():
            key = (recipe.amax_history_len, recipe.fp8_format)
            if self._delayed.get("key") != key:
                self._delayed = {"key": key,
                                 True: DelayedScalingState(recipe, True, 3, dev),
                                 False: DelayedScalingState(recipe, False, 2, dev)}
            else:
                for d in (True, False):
                    self._delayed[d].recipe = recipe
        kw = dict(tp_group=self.tp_group, sequence_parallel=self.sequence_parallel,
                  parallel_mode=self.parallel_mode)
        self.quantizers_fwd = make_quantizers(recipe, True, dev, delayed_state=self._delayed.get(True), **kw)
        self.quantizers_bwd = make_quantizers(recipe, False, dev, delayed_state=self._delayed.get(False), **kw)
        self.recipe = recipe
        self._weight_q = None

    def forward(self, inp: torch.Tensor, is_first_microbatch: Optional[bool] = None):
        if not FP8GlobalStateManager.is_fp8_enabled():
            out = torch.nn.functional.linear(inp, self.weight,
                                             self.bias if self.apply_bias_in_gemm else None)
            return self._finish(out)
        recipe = FP8GlobalStateManager.get_fp8_recipe()
        self._init_recipe_state(recipe)
        self._load_pending()
        grad_on = torch.is_grad_enabled()
        if recipe.delayed() and self.training and grad_on:
            FP8GlobalStateManager.add_fp8_tensors_to_global_buffer(self._delayed[True])
            FP8GlobalStateManager.add_fp8_tensors_to_global_buffer(self._delayed[False])
        weight_q = None
        if is_first_microbatch is False and self._weight_q is not None:
            weight_q = self._weight_q
        elif is_first_microbatch is True:
            q_w = self.quantizers_fwd[1]
            q_w.set_usage(rowwise=True, columnwise=recipe.backward_override is None)
            q_w.optimize_for_gemm = True
            weight_q = self._weight_q = q_w(self.weight.detach())
        inp = inp.to(self.weight.dtype) if inp.dtype != self.weight.dtype else inp
        # a bias that is added outside the GEMM (`return_bias`, row-parallel after the reduction) gets its
        # gradient from that add through autograd: `_Linear` must not see it, or bias.grad is counted twice
        # (reference linear.py: `bias if apply_bias else None` into `_Linear`, `gemm_bias_unfused_add`)
        out = _Linear.apply(inp, self.weight, self.bias if self.apply_bias_in_gemm else None, self, weight_q,
                            grad_on, *self._norm_params())
        return self._finish(out)

    def _norm_params(self):
        return ()

    def _symmetric_wanted(self) -> bool:
        if self.symmetric_ag is False or self._symm_failed:
            return False
        if self.symmetric_ag is None:   # auto: a CUDA tensor-parallel group of one NVSwitch domain over NCCL
            try:
                import torch.distributed as dist
                if dist.get_backend(self.tp_group) != "nccl" or self.tp_size > 8:
                    return False
            except Exception:
                return False
        return True

    def _reduce_scatter_for(self, which: str, rows_total: int, cols: int, dtype, device, applies: bool,
                            per_tensor_fp8: bool = True):
        """The fused GEMM + reduce-scatter for this output, or None (then GEMM followed by NCCL)."""
        if not (applies and self.sequence_parallel and self.tp_size > 1) or self.symmetric_rs is False:
            return None
        # Only the per-tensor FP8 GEMM with a 16-bit output sends 128-byte row pieces (staged through shared memory);
        # the other kinds store 32-byte fragments, which NVLink carries at less than half rate: measured slower than
        # NCCL at tp = 8, so by default they fuse only at tp = 2
        if self.symmetric_rs is None and self.tp_size > 2 and not (
                per_tensor_fp8 and dtype in (torch.bfloat16, torch.float16) and cols % 64 == 0
                and (rows_total // self.tp_size) % 128 == 0):
            return None
        if self.symmetric_rs is None and (os.environ.get("NVTE_B200_SYMM_RS", "1") == "0" or self._symm_failed
                                          or not self._symmetric_wanted()):
            return None
        if (dtype not in (torch.bfloat16, torch.float16, torch.float32) or rows_total % self.tp_size != 0
                or (rows_total // self.tp_size * cols * torch.empty(0, dtype=dtype).element_size()) % 32 != 0):
            return None
        key = (which, rows_total, cols, dtype)
        rs = self._rs.get(key)
        if rs is None:
            try:
                rs = SymmetricReduceScatter.shared(self.tp_group, rows_total, cols, dtype, device)
            except Exception as e:   # no symmetric memory on this system: NCCL from now on
                if self.symmetric_rs is True:
                    raise
                import warnings
                warnings.warn(f"symmetric-memory reduce-scatter unavailable ({str(e)[:120]}); using NCCL")
                self._symm_failed = True
                return None
            self._rs = {k: v for k, v in self._rs.items() if k[0] != which}   # one live buffer set per use
            self._rs[key] = rs
        return rs

    def _mx_gather_for(self, x: torch.Tensor, q):
        """The in-GEMM all-gather for an MXFP8 input shard, or None (then NCCL gathers data and scales)."""
        from .quantizers import MXFP8Quantizer
        if not (isinstance(q, MXFP8Quantizer) and self._symmetric_wanted() and self.symmetric_ag_flags
                and x.dtype in (torch.bfloat16, torch.float16) and self.out_features > 128):
            return None
        if os.environ.get("NVTE_B200_XCH_MXFP8", "1") == "0":
            return None
        rows, cols = x.shape
        key = (rows, cols)
        if self._mxg is None or self._mxg[0] != key:
            try:
                self._mxg = (key, SymmetricGatherMXFP8.shared(self.tp_group, rows, cols, x.device))
            except Exception as e:   # no symmetric memory on this system: NCCL from now on
                if self.symmetric_ag is True:
                    raise
                import warnings
                warnings.warn(f"symmetric-memory MXFP8 gather unavailable ({str(e)[:120]}); using NCCL")
                self._symm_failed = True
                self._mxg = (key, None)
        return self._mxg[1]

    def _exchange_for(self, x: torch.Tensor, q):
        """The flag-driven symmetric exchange for this input, or None when it does not apply (then the
        cast + barrier path or NCCL take over)."""
        if not (self._symmetric_wanted() and self.symmetric_ag_flags and isinstance(q, Float8CurrentScalingQuantizer)
                and q.with_amax_reduction):
            return None
        rows, cols = x.shape
        if (x.dtype not in (torch.bfloat16, torch.float16) or rows % 256 != 0 or cols % 16 != 0
                or self.out_features <= 128 or self.symmetric_ag_chunks != 1):
            return None
        if self._xch is None or (self._xch.rows_local, self._xch.cols) != (rows, cols):
            try:
                self._xch = SymmetricExchange.shared(self.tp_group, rows, cols, x.device)
            except Exception as e:   # no symmetric memory on this system: NCCL from now on
                if self.symmetric_ag is True:
                    raise
                import warnings
                warnings.warn(f"symmetric-memory exchange unavailable ({str(e)[:120]}); using NCCL")
                self._symm_failed = True
                return None
        return self._xch

    def _symmetric_cast_and_gather(self, x: torch.Tensor, q):
        if self._symm is None or (self._symm.rows_local, self._symm.cols) != tuple(x.shape):
            self._symm = SymmetricFP8Gather(self.tp_group, x.shape[0], x.shape[1], x.device,
                                            chunks=self.symmetric_ag_chunks)
        if isinstance(q, Float8CurrentScalingQuantizer):
            self._symm.allreduce_amax(x, q.amax)
            ops.compute_scale_from_amax(q.amax, q.scale, q.dtype, q.force_pow_2_scales, q.amax_epsilon)
            return self._symm.cast_and_gather(x, q.scale, q.dtype)
        return self._symm.cast_and_gather(x, q.scale, q.dtype, amax=q.amax)

    def _finish(self, out):
        ln_out = None
        if isinstance(out, tuple):   # LayerNormLinear(return_layernorm_output=True)
            out, ln_out = out
        if self.use_bias and not self.apply_bias_in_gemm:
            res = (out, self.bias) if self.return_bias else (out + self.bias,)
        else:
            res = (out, None) if self.return_bias else (out,)
        if ln_out is not None:
            res = res + (ln_out,)
        return res if len(res) > 1 else res[0]

    # -- checkpointing of the delayed-scaling state (reference module/base.py:1237-1355) --
    # The reference's format: a uint8 tensor holding a pickled dict (checkpoint.py); state written by the reference loads.
    def get_extra_state(self):
        from . import checkpoint
        return checkpoint.pack(self.recipe, [self._delayed] if self._delayed else [])

    def set_extra_state(self, state):
        from . import checkpoint
        st = checkpoint.unpack(state)
        self._pending_extra_state = st
        rec = st.get("recipe") if st else None
        if rec is not None and hasattr(rec, "delayed") and rec.delayed():
            # as the reference: the scaling tensors are created from the stored recipe and filled right away
            self._init_recipe_state(rec)
            self._load_pending()

    def _load_pending(self):
        from . import checkpoint
        st = getattr(self, "_pending_extra_state", None)
        if st and self._delayed:
            checkpoint.load_into(st, [self._delayed], type(self).__name__)   # warns and ignores on a shape mismatch
            self._pending_extra_state = None


class LayerNormLinear(Linear):
    """Drop-in for `te.LayerNormLinear` (reference module/layernorm_linear.py:1156-1700): LayerNorm
    or RMSNorm of the input followed by the Linear above, with the normalization kernel writing
    the quantized GEMM operand directly (per-tensor FP8 and MXFP8; current scaling takes the amax
    from the norm kernel; NVFP4 normalizes in bf16 and quantizes afterwards).

    Extra parameters, as the reference: `eps`, `normalization` ("LayerNorm" | "RMSNorm"),
    `zero_centered_gamma`, `return_layernorm_output`. Parameters `layer_norm_weight`,
    `layer_norm_bias` keep the reference's names. Row-parallel mode is not offered (the reference
    supports only column-parallel here, layernorm_linear.py:1345).
    """

    def __init__(self, in_features: int, out_features: int, eps: float = 1e-5, *, bias: bool = True,
                 normalization: str = "LayerNorm", zero_centered_gamma: bool = False,
                 return_layernorm_output: bool = False, params_dtype: Optional[torch.dtype] = None,
                 parallel_mode: Optional[str] = None, device="cuda", **kwargs):
        assert normalization in ("LayerNorm", "RMSNorm"), "Unsupported normalization type!"
        assert parallel_mode in (None, "column"), "LayerNormLinear supports column parallelism only"
        super().__init__(in_features, out_features, bias, params_dtype=params_dtype, parallel_mode=parallel_mode,
                         device=device, **kwargs)
        self.eps, self.normalization = eps, normalization
        self.zero_centered_gamma = zero_centered_gamma
        self.return_layernorm_output = return_layernorm_output
        dtype = self.weight.dtype
        init = torch.zeros if zero_centered_gamma else torch.ones
        self.layer_norm_weight = torch.nn.Parameter(init(in_features, dtype=dtype, device=device))
        if normalization == "LayerNorm":
            self.layer_norm_bias = torch.nn.Parameter(torch.zeros(in_features, dtype=dtype, device=device))
        else:
            self.register_parameter("layer_norm_bias", None)

    def _norm_params(self):
        return (self.layer_norm_weight, self.layer_norm_bias)

    def forward(self, inp: torch.Tensor, is_first_microbatch: Optional[bool] = None):
        if FP8GlobalStateManager.is_fp8_enabled():
            return super().forward(inp, is_first_microbatch)
        # outside autocast: the same normalization kernel, then a plain high-precision linear
        from .normalization import _Norm
        inp = inp.to(self.weight.dtype) if inp.dtype != self.weight.dtype else inp
        ln_out = _Norm.apply(inp, self.layer_norm_weight, self.layer_norm_bias, self.eps, self.zero_centered_gamma)
        out = torch.nn.functional.linear(ln_out, self.weight, self.bias if self.apply_bias_in_gemm else None)
        return self._finish((out, ln_out) if self.return_layernorm_output else out)
