"""`autocast(recipe=...)` and the per-module recipe state behind it.

Mirrors the reference's pytorch/quantization.py: `autocast` / `fp8_autocast` (:962-1055),
`FP8GlobalStateManager` (:405-960: nesting depth, first-module flag, delayed-scaling amax
reduction + fused scale update at region exit), `RecipeState.make_quantizers` (:1169-1725).
Roles of the quantizers a Linear needs: forward [input, weight, output], backward
[grad_output, grad_input] (reference module/base.py, `num_gemms=1` -> 3 fwd + 2 bwd slots).
"""
from contextlib import contextmanager
from typing import List, Optional

import torch

from . import ops
from .distributed import get_distributed_world_size
from .quantizers import (Float8CurrentScalingQuantizer, Float8Quantizer, MXFP8Quantizer,
                         NVFP4Quantizer)
from .recipe import DelayedScaling, Recipe, fp8_dtype_of


def get_default_fp8_recipe() -> Recipe:
    """MXFP8 on Blackwell (reference quantization.py:262-273 picks it for sm_100+)."""
    from .recipe import MXFP8BlockScaling
    return MXFP8BlockScaling()


class DelayedScalingState:
    """scale / amax-history tensors of one module and direction
    (reference quantization.py `DelayedScalingRecipeState`, :1213-1290)."""

    def __init__(self, recipe: DelayedScaling, forward: bool, num_quantizers: int, device):
        self.recipe, self.forward, self.n = recipe, forward, num_quantizers
        self.dtype = fp8_dtype_of(recipe.fp8_format, forward)
        self.scale = torch.ones(num_quantizers, dtype=torch.float32, device=device)
        self.amax_history = torch.zeros(recipe.amax_history_len, num_quantizers,
                                        dtype=torch.float32, device=device)

    def make_quantizers(self) -> list:
        # each quantizer sees a 1-element view of its scale and of history row 0
        return [Float8Quantizer(self.scale[i:i + 1], self.amax_history[0][i:i + 1], self.dtype)
                for i in range(self.n)]


class FP8GlobalStateManager:
    """Process-wide autocast state (reference quantization.py:405-960)."""
    FP8_ENABLED = False
    FP8_CALIBRATION = False
    FP8_RECIPE: Optional[Recipe] = None
    FP8_GROUP = None
    AUTOCAST_DEPTH = 0
    IS_FIRST_FP8_MODULE = False
    # delayed-scaling states touched in the current region, per direction
    _touched = {True: [], False: []}

    @classmethod
    def is_fp8_enabled(cls) -> bool:
        return cls.FP8_ENABLED

    @classmethod
    def is_fp8_calibration(cls) -> bool:
        return cls.FP8_CALIBRATION

    @classmethod
    def get_fp8_recipe(cls) -> Recipe:
        return cls.FP8_RECIPE if cls.FP8_RECIPE is not None else get_default_fp8_recipe()

    @classmethod
    def get_fp8_group(cls):
        return cls.FP8_GROUP

    @classmethod
    def is_first_fp8_module(cls) -> bool:
        """True for the first module that runs in an autocast region; its backward runs last
        and triggers the backward amax reduction (reference quantization.py:574-583)."""
        tmp = cls.IS_FIRST_FP8_MODULE
        cls.IS_FIRST_FP8_MODULE = False
        return tmp

    @classmethod
    def add_fp8_tensors_to_global_buffer(cls, state: DelayedScalingState) -> None:
        lst = cls._touched[state.forward]
        if all(s is not state for s in lst):
            lst.append(state)

    @classmethod
    def reduce_and_update_fp8_tensors(cls, forward: bool = True) -> None:
        """All-reduce(MAX) the newest amaxes over the amax reduction group, then one fused launch
        that rolls every history and recomputes every scale
        (reference quantization.py:646-720, common/recipe/delayed_scaling.cu:300-420)."""
        states: List[DelayedScalingState] = cls._touched[forward]
        cls._touched[forward] = []
        if not states:
            return
        recipe = states[0].recipe
        buf = torch.cat([s.amax_history[0] for s in states]).contiguous()
        group = cls.FP8_GROUP
        if recipe.reduce_amax and get_distributed_world_size(group) > 1:
            torch.distributed.all_reduce(buf, op=torch.distributed.ReduceOp.MAX, group=group)
        ops.fused_amax_and_scale_update_after_reduction(
            buf, [s.amax_history for s in states], [s.scale for s in states],
            recipe.amax_compute_algo, states[0].dtype, float(recipe.margin))

    @classmethod
    def autocast_enter(cls, enabled, calibrating, recipe, group):
        saved = (cls.FP8_ENABLED, cls.FP8_CALIBRATION, cls.FP8_RECIPE, cls.FP8_GROUP)
        cls.FP8_ENABLED, cls.FP8_CALIBRATION = enabled, calibrating
        cls.FP8_RECIPE = recipe if recipe is not None else get_default_fp8_recipe()
        cls.FP8_GROUP = group
        if cls.AUTOCAST_DEPTH == 0:
            cls.IS_FIRST_FP8_MODULE = True
        cls.AUTOCAST_DEPTH += 1
        return saved

    @classmethod
    def autocast_exit(cls, saved, enabled):
        cls.AUTOCAST_DEPTH -= 1
        # forward scales are refreshed once, when the outermost region closes
        # (reference quantization.py `fp8_autocast_exit`: skipped under no_grad)
        if cls.AUTOCAST_DEPTH == 0:
            if enabled and torch.is_grad_enabled():
                cls.reduce_and_update_fp8_tensors(forward=True)
            else:
                cls._touched[True] = []
        cls.FP8_ENABLED, cls.FP8_CALIBRATION, cls.FP8_RECIPE, cls.FP8_GROUP = saved


@contextmanager
def autocast(enabled: bool = True, calibrating: bool = False, recipe: Optional[Recipe] = None,
             amax_reduction_group=None):
    """Context manager selecting the quantization recipe for the modules run inside it
    (reference quantization.py:962-1055)."""
    saved = FP8GlobalStateManager.autocast_enter(enabled, calibrating, recipe, amax_reduction_group)
    try:
        yield
    finally:
        FP8GlobalStateManager.autocast_exit(saved, enabled)


@contextmanager
def fp8_autocast(enabled: bool = True, calibrating: bool = False, fp8_recipe: Optional[Recipe] = None,
                 fp8_group=None):
    """Deprecated spelling kept by the reference (quantization.py:1058-1080)."""
    with autocast(enabled=enabled, calibrating=calibrating, recipe=fp8_recipe,
                  amax_reduction_group=fp8_group):
        yield


def make_quantizers(recipe: Recipe, forward: bool, device, *, tp_group=None,
                    sequence_parallel: bool = False, parallel_mode: Optional[str] = None,
                    delayed_state: Optional[DelayedScalingState] = None) -> list:
    """Quantizers of one Linear for one direction: forward [input, weight, output],
    backward [grad_output, grad_input] (reference quantization.py:1169-1725 and the
    per-role customisation in module/linear.py:1960-2120)."""
    n = 3 if forward else 2
    if recipe.delayed():
        assert delayed_state is not None
        return delayed_state.make_quantizers()
    # amax must agree across the TP group when the quantized tensor is gathered afterwards:
    # the column-parallel input (fwd) and the row-parallel grad_output (bwd) under SP.
    gathered = sequence_parallel and ((forward and parallel_mode == "column") or
                                      (not forward and parallel_mode == "row"))
    if recipe.float8_current_scaling():
        dt = fp8_dtype_of(recipe.fp8_format, forward)
        roles = ([recipe.fp8_quant_fwd_inp, recipe.fp8_quant_fwd_weight, recipe.fp8_quant_fwd_inp]
                 if forward else [recipe.fp8_quant_bwd_grad, recipe.fp8_quant_bwd_grad])
        qs = [Float8CurrentScalingQuantizer(dt, device, force_pow_2_scales=r.power_2_scale,
                                            amax_epsilon=r.amax_epsilon) for r in roles]
        if gathered:
            qs[0].with_amax_reduction, qs[0].amax_reduction_group = True, tp_group
        return qs
    if recipe.mxfp8():
        dt = fp8_dtype_of(recipe.fp8_format, forward)
        return [MXFP8Quantizer(dt) for _ in range(n)]
    if recipe.nvfp4():
        roles = ([recipe.fp4_quant_fwd_inp, recipe.fp4_quant_fwd_weight, recipe.fp4_quant_fwd_inp]
                 if forward else [recipe.fp4_quant_bwd_grad, recipe.fp4_quant_bwd_grad])
        qs = [NVFP4Quantizer(with_rht=r.random_hadamard_transform,
                             with_post_rht_amax=r.random_hadamard_transform,
                             with_2d_quantization=r.fp4_2d_quantization,
                             stochastic_rounding=r.stochastic_rounding) for r in roles]
        if gathered:
            qs[0].with_amax_reduction, qs[0].amax_reduction_group = True, tp_group
        return qs
    raise NotImplementedError(f"recipe {type(recipe).__name__} is not part of the Blackwell Linear path")
