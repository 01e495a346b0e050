"""Quantizer objects: "builder" classes that turn a high-precision tensor into a quantized
storage, one class per recipe. Same names, constructor arguments and methods as the reference
(reference pytorch/quantized_tensor.py:205-400, pytorch/tensor/float8_tensor.py,
mxfp8_tensor.py, nvfp4_tensor.py; C++ twins pytorch/csrc/quantizer.cpp:263-2626). Each
`quantize` is a short sequence of `nvte_*` calls through `ops`.
"""
from typing import Optional

import torch

from . import ops


class Quantizer:
    """Base class (reference quantized_tensor.py:205-400)."""

    def __init__(self, *, rowwise: bool = True, columnwise: bool = True):
        self.rowwise_usage = rowwise
        self.columnwise_usage = columnwise
        self.internal = False
        # write block scales directly in the tcgen05 layout (no separate swizzle launch)
        self.optimize_for_gemm = False

    def set_usage(self, *, rowwise: Optional[bool] = None, columnwise: Optional[bool] = None):
        if rowwise is not None:
            self.rowwise_usage = rowwise
        if columnwise is not None:
            self.columnwise_usage = columnwise

    def __call__(self, tensor: torch.Tensor):
        return self.quantize(tensor)

    def quantize(self, tensor: torch.Tensor, *, out=None, noop_flag=None):
        raise NotImplementedError

    def quantize_with_dbias(self, tensor: torch.Tensor):
        """(quantized tensor, tensor.sum(0)): the reference's `tex.bgrad_quantize`. Recipes whose
        cast cannot fuse the reduction (amax pre-pass) do it separately."""
        flat = tensor.reshape(-1, tensor.shape[-1])
        return self.quantize(tensor), flat.sum(dim=0).to(tensor.dtype)

    def update_quantized(self, src: torch.Tensor, dst, *, noop_flag: Optional[torch.Tensor] = None):
        """Re-quantize `src` into the existing storage `dst` (weight caching path)."""
        return self.quantize(src, out=dst, noop_flag=noop_flag)

    def copy(self):
        import copy
        return copy.copy(self)


class Float8Quantizer(Quantizer):
    """Per-tensor FP8 with an externally owned scale and amax slot (delayed scaling).
    Reference float8_tensor.py `Float8Quantizer(scale, amax, fp8_dtype, rowwise, columnwise)`."""

    def __init__(self, scale: torch.Tensor, amax: torch.Tensor, fp8_dtype: torch.dtype, *,
                 rowwise: bool = True, columnwise: bool = True):
        super().__init__(rowwise=rowwise, columnwise=columnwise)
        self.scale, self.amax, self.dtype = scale, amax, fp8_dtype

    def quantize(self, tensor, *, out=None, noop_flag=None):
        # sm_100 GEMMs read FP8 in either major-ness: no transpose buffer
        # (reference csrc/quantizer.cpp:298-308)
        return ops.fp8_quantize(tensor, self.scale, self.dtype, amax=self.amax, noop=noop_flag,
                                out=out)

    def quantize_with_dbias(self, tensor):
        return ops.fp8_quantize(tensor, self.scale, self.dtype, amax=self.amax, with_dbias=True)


class Float8CurrentScalingQuantizer(Quantizer):
    """Per-tensor FP8, scale computed from the tensor itself: amax -> (all-reduce MAX over the
    amax reduction group) -> scale -> cast (reference csrc/quantizer.cpp:886-938)."""

    def __init__(self, fp8_dtype: torch.dtype, device=None, *, rowwise: bool = True,
                 columnwise: bool = True, with_amax_reduction: bool = False,
                 amax_reduction_group=None, force_pow_2_scales: bool = False,
                 amax_epsilon: float = 0.0):
        super().__init__(rowwise=rowwise, columnwise=columnwise)
        device = torch.device("cuda") if device is None else device
        self.dtype = fp8_dtype
        self.scale = torch.empty(1, dtype=torch.float32, device=device)
        self.amax = torch.empty(1, dtype=torch.float32, device=device)
        self.with_amax_reduction = with_amax_reduction
        self.amax_reduction_group = amax_reduction_group
        self.force_pow_2_scales = force_pow_2_scales
        self.amax_epsilon = amax_epsilon

    def _amax_to_scale(self, tensor, noop_flag=None):
        ops.compute_amax(tensor, self.amax, noop=noop_flag)
        if self.with_amax_reduction:
            # stays on the compute stream: no host wait (NCCL all-reduce of one float)
            torch.distributed.all_reduce(self.amax, op=torch.distributed.ReduceOp.MAX,
                                         group=self.amax_reduction_group)
        ops.compute_scale_from_amax(self.amax, self.scale, self.dtype, self.force_pow_2_scales,
                                    self.amax_epsilon)

    def quantize(self, tensor, *, out=None, noop_flag=None):
        self._amax_to_scale(tensor, noop_flag)
        return ops.fp8_quantize(tensor, self.scale, self.dtype, amax=None, noop=noop_flag, out=out)

    def quantize_with_dbias(self, tensor):
        # the bias gradient rides on the cast pass (nvte_quantize_dbias); the amax pre-pass stays
        self._amax_to_scale(tensor)
        return ops.fp8_quantize(tensor, self.scale, self.dtype, amax=None, with_dbias=True)


class MXFP8Quantizer(Quantizer):
    """MXFP8 1x32 / 32x1 blocks, E8M0 scales (reference mxfp8_tensor.py `MXFP8Quantizer`)."""

    def __init__(self, fp8_dtype: torch.dtype, *, rowwise: bool = True, columnwise: bool = True):
        super().__init__(rowwise=rowwise, columnwise=columnwise)
        self.dtype = fp8_dtype

    def quantize(self, tensor, *, out=None, noop_flag=None):
        return ops.mxfp8_quantize(tensor, self.dtype, rowwise=self.rowwise_usage,
                                  columnwise=self.columnwise_usage,
                                  swizzled_scales=self.optimize_for_gemm, noop=noop_flag, out=out)

    def quantize_with_dbias(self, tensor):
        return ops.mxfp8_quantize(tensor, self.dtype, rowwise=self.rowwise_usage,
                                  columnwise=self.columnwise_usage,
                                  swizzled_scales=self.optimize_for_gemm, with_dbias=True)


class NVFP4Quantizer(Quantizer):
    """NVFP4 1x16 blocks, E4M3 block scales + FP32 global amax (reference nvfp4_tensor.py
    `NVFP4Quantizer`, csrc/quantizer.cpp:1778-2626). RHT applies to the column-wise operand only."""

    def __init__(self, fp4_dtype=None, *, rowwise: bool = True, columnwise: bool = True,
                 with_amax_reduction: bool = False, amax_reduction_group=None,
                 with_rht: bool = False, with_post_rht_amax: bool = False,
                 with_2d_quantization: bool = False, stochastic_rounding: bool = False):
        super().__init__(rowwise=rowwise, columnwise=columnwise)
        self.dtype = fp4_dtype
        self.with_amax_reduction = with_amax_reduction
        self.amax_reduction_group = amax_reduction_group
        self.with_rht = with_rht
        self.with_post_rht_amax = with_post_rht_amax
        self.with_2d_quantization = with_2d_quantization
        self.stochastic_rounding = stochastic_rounding

    def quantize(self, tensor, *, out=None, noop_flag=None):
        return ops.nvfp4_quantize(
            tensor, rowwise=self.rowwise_usage, columnwise=self.columnwise_usage,
            with_rht=self.with_rht, with_2d_quantization=self.with_2d_quantization,
            swizzled_scales=self.optimize_for_gemm,
            stochastic_rounding=self.stochastic_rounding,
            amax_reduction_group=self.amax_reduction_group if self.with_amax_reduction else None,
            out=out)
