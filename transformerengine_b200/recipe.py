"""Quantization recipes for the low-precision Linear path.

Same class names, fields and defaults as the reference's user-facing recipes
(reference transformer_engine/common/recipe/__init__.py:20-617), restricted to the
recipes that exist on Blackwell for `te.Linear`: DelayedScaling, Float8CurrentScaling,
MXFP8BlockScaling, NVFP4BlockScaling. `Float8BlockScaling` is a Hopper recipe
(FP32 1x128 / 128x128 scales) and is not part of this path.
"""
import os
from dataclasses import dataclass, field
from enum import Enum
from typing import Callable, Literal, NamedTuple, Optional, Union

import torch


class _FormatHelper(NamedTuple):
    max_fwd: float
    max_bwd: float


class Format(Enum):
    """E4M3 everywhere, E5M2 everywhere, or HYBRID = E4M3 forward / E5M2 backward
    (reference recipe/__init__.py:29-51)."""
    E2M1 = _FormatHelper(max_fwd=6.0, max_bwd=6.0)
    E4M3 = _FormatHelper(max_fwd=448.0, max_bwd=448.0)
    E5M2 = _FormatHelper(max_fwd=57344.0, max_bwd=57344.0)
    HYBRID = _FormatHelper(max_fwd=448.0, max_bwd=57344.0)


def fp8_dtype_of(fmt: Format, fprop_tensor: bool) -> torch.dtype:
    """Element type of forward (input, weight) or backward (grad) tensors
    (reference pytorch/quantization.py:280-300)."""
    if fmt == Format.E4M3 or (fmt == Format.HYBRID and fprop_tensor):
        return torch.float8_e4m3fn
    return torch.float8_e5m2


@dataclass(frozen=True)
class MMParams:
    """GEMM options (reference recipe/__init__.py:54-64). The tcgen05 GEMM always accumulates
    in FP32 TMEM, so `use_split_accumulator` is accepted and has no effect."""
    use_split_accumulator: bool = True


@dataclass(frozen=True)
class QParams:
    """Per-tensor-role quantization options (reference recipe/__init__.py:78-90)."""
    power_2_scale: bool = False
    amax_epsilon: float = 0.0
    random_hadamard_transform: bool = False
    stochastic_rounding: bool = False
    fp4_2d_quantization: bool = False


class Recipe:
    """Base class; the predicates mirror reference recipe/__init__.py:107-170."""

    def delayed(self):
        return isinstance(self, DelayedScaling)

    def float8_current_scaling(self):
        return isinstance(self, Float8CurrentScaling)

    def mxfp8(self):
        return isinstance(self, MXFP8BlockScaling)

    def nvfp4(self):
        return isinstance(self, NVFP4BlockScaling)

    def float8_per_tensor_scaling(self):
        return self.delayed() or self.float8_current_scaling()


_OVERRIDES = (None, "high_precision", "dequantized")


@dataclass
class DelayedScaling(Recipe):
    """Per-tensor FP8 with scales derived from an amax history
    (reference recipe/__init__.py:173-283)."""
    margin: int = 0
    fp8_format: Format = Format.HYBRID
    amax_history_len: int = 1024
    amax_compute_algo: Union[Literal["max", "most_recent"], Callable] = "max"
    scaling_factor_compute_algo: Optional[Callable] = None
    reduce_amax: bool = True
    fp8_dpa: bool = False
    fp8_mha: bool = False
    backward_override: Optional[str] = None

    def __post_init__(self):
        assert self.fp8_format != Format.E5M2, "Pure E5M2 training is not supported."
        assert self.backward_override in _OVERRIDES
        assert callable(self.amax_compute_algo) or self.amax_compute_algo in ("max", "most_recent")
        if callable(self.amax_compute_algo) or self.scaling_factor_compute_algo is not None:
            raise NotImplementedError(
                "Custom amax / scaling-factor callables run on the host in the reference; only the "
                "fused 'max' and 'most_recent' device updates are implemented here")


@dataclass
class Float8CurrentScaling(Recipe):
    """Per-tensor FP8, scale from the current tensor's amax
    (reference recipe/__init__.py:286-334)."""
    use_power_2_scales: bool = os.getenv("NVTE_FP8_CURRENT_SCALING_POWER_2_SCALES", "0") == "1"
    fp8_format: Format = Format.HYBRID
    fp8_quant_fwd_inp: QParams = field(default_factory=QParams)
    fp8_quant_fwd_weight: QParams = field(default_factory=QParams)
    fp8_quant_bwd_grad: QParams = field(default_factory=QParams)
    fp8_gemm_fprop: MMParams = MMParams(use_split_accumulator=False)
    fp8_gemm_dgrad: MMParams = MMParams(use_split_accumulator=True)
    fp8_gemm_wgrad: MMParams = MMParams(use_split_accumulator=True)
    fp8_dpa: bool = False
    fp8_mha: bool = False
    backward_override: Optional[str] = None

    def __post_init__(self):
        assert self.fp8_format != Format.E5M2, "Pure E5M2 training is not supported."
        assert self.backward_override in _OVERRIDES
        if self.use_power_2_scales:
            q = QParams(power_2_scale=True)
            self.fp8_quant_fwd_inp = self.fp8_quant_fwd_weight = self.fp8_quant_bwd_grad = q


@dataclass
class MXFP8BlockScaling(Recipe):
    """MXFP8: 1x32 blocks with E8M0 scales (reference recipe/__init__.py:337-385)."""
    margin: int = 0
    fp8_format: Format = Format.E4M3
    fp8_dpa: bool = False
    fp8_mha: bool = False
    backward_override: Optional[str] = None

    def __post_init__(self):
        assert self.fp8_format != Format.E5M2, "Pure E5M2 training is not supported."
        assert self.backward_override in _OVERRIDES


@dataclass
class NVFP4BlockScaling(Recipe):
    """NVFP4: 1x16 blocks with E4M3 scales and an FP32 global scale; random Hadamard transform
    on the column-wise (wgrad) operands of activations and gradients; 16x16 (2-D) scale tiles
    for weights; stochastic rounding for gradients (reference recipe/__init__.py:479-617)."""
    disable_rht: bool = os.getenv("NVTE_NVFP4_DISABLE_RHT", "0") == "1"
    disable_stochastic_rounding: bool = os.getenv("NVTE_NVFP4_DISABLE_STOCHASTIC_ROUNDING", "0") == "1"
    disable_2d_quantization: bool = os.getenv("NVTE_NVFP4_DISABLE_2D_QUANTIZATION", "0") == "1"
    fp4_format: Format = Format.E2M1
    fp8_format: Format = Format.E4M3
    fp8_dpa: bool = False
    fp8_mha: bool = False
    backward_override: Optional[str] = None

    def __post_init__(self):
        assert self.fp4_format == Format.E2M1, "Only E2M1 is supported for NVFP4 scaling"
        assert self.fp8_format == Format.E4M3, "Only E4M3 is supported for NVFP4 scaling"
        assert self.backward_override in _OVERRIDES
        # RHT only on the column-wise usage, so that it serves the wgrad GEMM.
        self.fp4_quant_fwd_inp = QParams(random_hadamard_transform=not self.disable_rht)
        self.fp4_quant_fwd_weight = QParams(fp4_2d_quantization=not self.disable_2d_quantization)
        self.fp4_quant_bwd_grad = QParams(random_hadamard_transform=not self.disable_rht,
                                          stochastic_rounding=not self.disable_stochastic_rounding)
