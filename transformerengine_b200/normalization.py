"""LayerNorm / RMSNorm with the Linear path's quantization fused in.

Mirrors the reference's `tex.layernorm_fwd / rmsnorm_fwd / layernorm_bwd / rmsnorm_bwd`
(pytorch/csrc/extensions/normalization.cpp:60-330: pick the output the quantizer can take straight
from the norm kernel, otherwise normalize in high precision and quantize afterwards) and its
`te.LayerNorm` / `te.RMSNorm` modules (pytorch/module/layernorm.py, rmsnorm.py). Every device
operation is an `nvte_*` call through `ops`.
"""
from typing import Iterable, Optional, Union

import torch

from . import cabi as C
from . import ops
from .quantizers import (Float8CurrentScalingQuantizer, Float8Quantizer, MXFP8Quantizer, Quantizer)


def _norm_fwd(layernorm: bool, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], eps: float,
              quantizer: Optional[Quantizer], zero_centered_gamma: bool, want_high_precision: bool = False):
    """Returns (out, high-precision out or None, mu or None, rsigma). `out` is a quantized storage
    when `quantizer` is given, else the normalized tensor in x's dtype."""
    shape = tuple(x.shape)
    x2 = x.reshape(-1, shape[-1]).contiguous()
    N, H = x2.shape
    dev = x2.device

    def run(z):
        return ops.norm_fwd(layernorm, x2, weight, bias, eps, z, zero_centered_gamma)

    def high_precision(amax=None):
        y = torch.empty_like(x2)
        z = C.plain(y)
        if amax is not None:
            amax.zero_()
            z.set(C.kAmax, amax)
        mu, rs = run(z)
        return y, mu, rs

    e4m3 = quantizer is not None and getattr(quantizer, "dtype", None) == torch.float8_e4m3fn
    if quantizer is None:
        y, mu, rs = high_precision()
        return y.reshape(shape), None, mu, rs
    if want_high_precision or not e4m3:
        y, mu, rs = high_precision()
        return quantizer(y.reshape(shape)), y.reshape(shape), mu, rs
    if isinstance(quantizer, MXFP8Quantizer) and (not quantizer.columnwise_usage or N % 32 == 0):
        # norm -> MXFP8 in one kernel (the reference needs a cuDNN graph for this output)
        rs_shape, cs_shape = ops.mxfp8_scale_shapes(N, H)
        col = quantizer.columnwise_usage
        out = ops.MXFP8TensorStorage(
            rowwise_data=torch.empty(N, H, dtype=torch.uint8, device=dev),
            rowwise_scale_inv=torch.zeros(rs_shape, dtype=torch.uint8, device=dev),
            columnwise_data=torch.empty(N, H, dtype=torch.uint8, device=dev) if col else None,
            columnwise_scale_inv=torch.zeros(cs_shape, dtype=torch.uint8, device=dev) if col else None,
            fp8_dtype=quantizer.dtype, shape=(N, H), with_gemm_swizzled_scales=quantizer.optimize_for_gemm)
        mu, rs = run(out.handle())
        out.shape = shape
        out._handle = None
        return out, None, mu, rs
    if isinstance(quantizer, Float8Quantizer):
        # delayed scaling: z = fp8(y * scale), amax(y) into the quantizer's history slot
        out = ops.Float8TensorStorage(data=torch.empty(N, H, dtype=torch.uint8, device=dev),
                                      scale_inv=torch.empty(1, dtype=torch.float32, device=dev),
                                      fp8_dtype=quantizer.dtype, shape=shape)
        z = C.TensorHandle(C.DELAYED_TENSOR_SCALING)
        z.set(C.kRowwiseData, out.data, C.kFloat8E4M3).set(C.kRowwiseScaleInv, out.scale_inv)
        z.set(C.kScale, quantizer.scale).set(C.kAmax, quantizer.amax)
        mu, rs = run(z)
        out.data = out.data.reshape(shape)
        return out, None, mu, rs
    if isinstance(quantizer, Float8CurrentScalingQuantizer):
        # current scaling: the norm kernel also produces amax(y); reduce, derive the scale, cast
        y, mu, rs = high_precision(amax=quantizer.amax)
        if quantizer.with_amax_reduction:
            torch.distributed.all_reduce(quantizer.amax, op=torch.distributed.ReduceOp.MAX,
                                         group=quantizer.amax_reduction_group)
        ops.compute_scale_from_amax(quantizer.amax, quantizer.scale, quantizer.dtype,
                                    quantizer.force_pow_2_scales, quantizer.amax_epsilon)
        out = ops.fp8_quantize(y.reshape(shape), quantizer.scale, quantizer.dtype, amax=None)
        return out, None, mu, rs
    y, mu, rs = high_precision()
    return quantizer(y.reshape(shape)), y.reshape(shape), mu, rs


def layernorm_fwd(input: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, eps: float, out=None,
                  quantizer: Optional[Quantizer] = None, otype=None, sm_margin: int = 0,
                  zero_centered_gamma: bool = False):
    """Reference `tex.layernorm_fwd` (normalization.cpp:60-190): (out, mu, rsigma)."""
    assert out is None, "preallocated outputs are not supported"
    o, _, mu, rs = _norm_fwd(True, input, weight, bias, eps, quantizer, zero_centered_gamma)
    return o, mu, rs


def rmsnorm_fwd(input: torch.Tensor, weight: torch.Tensor, eps: float, out=None,
                quantizer: Optional[Quantizer] = None, otype=None, sm_margin: int = 0,
                zero_centered_gamma: bool = False):
    """Reference `tex.rmsnorm_fwd` (normalization.cpp:290-420): (out, None, rsigma)."""
    assert out is None, "preallocated outputs are not supported"
    o, _, _, rs = _norm_fwd(False, input, weight, None, eps, quantizer, zero_centered_gamma)
    return o, None, rs


def layernorm_bwd(dz, x, mu, rsigma, gamma, sm_margin: int = 0, zero_centered_gamma: bool = False):
    """Reference `tex.layernorm_bwd` (normalization.cpp:192-288): (dx, dgamma, dbeta)."""
    H = x.shape[-1]
    dx, dg, db = ops.norm_bwd(dz.reshape(-1, H), x.reshape(-1, H).contiguous(), mu, rsigma, gamma,
                              zero_centered_gamma)
    return dx.reshape(x.shape), dg, db


def rmsnorm_bwd(dz, x, rsigma, gamma, sm_margin: int = 0, zero_centered_gamma: bool = False):
    """Reference `tex.rmsnorm_bwd`: (dx, dgamma)."""
    H = x.shape[-1]
    dx, dg, _ = ops.norm_bwd(dz.reshape(-1, H), x.reshape(-1, H).contiguous(), None, rsigma, gamma,
                             zero_centered_gamma)
    return dx.reshape(x.shape), dg


def rmsnorm_bwd_add(dz, x, add, rsigma, gamma, sm_margin: int = 0, zero_centered_gamma: bool = False):
    """Reference `tex.rmsnorm_bwd_add`: dx also accumulates `add` (a residual branch's gradient)."""
    H = x.shape[-1]
    dx, dg, _ = ops.norm_bwd(dz.reshape(-1, H), x.reshape(-1, H).contiguous(), None, rsigma, gamma,
                             zero_centered_gamma, add=add.reshape(-1, H))
    return dx.reshape(x.shape), dg


class _Norm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, zero_centered_gamma):
        layernorm = bias is not None
        y, _, mu, rs = _norm_fwd(layernorm, x, weight, bias, eps, None, zero_centered_gamma)
        ctx.save_for_backward(x, weight, mu if layernorm else rs, rs)
        ctx.layernorm, ctx.zc = layernorm, zero_centered_gamma
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, mu, rs = ctx.saved_tensors
        H = x.shape[-1]
        dx, dg, db = ops.norm_bwd(dy.reshape(-1, H), x.reshape(-1, H).contiguous(), mu if ctx.layernorm else None,
                                  rs, weight, ctx.zc)
        return dx.reshape(x.shape), dg, db, None, None


class _NormModule(torch.nn.Module):
    _layernorm = True

    def __init__(self, normalized_shape: Union[Iterable[int], int, None] = None, eps: float = 1e-5,
                 sequence_parallel: Optional[bool] = None, params_dtype: Optional[torch.dtype] = None,
                 zero_centered_gamma: bool = False, hidden_size: Optional[int] = None, device="cuda", **kwargs):
        super().__init__()
        if normalized_shape is None:
            if hidden_size is None:
                raise RuntimeError("Neither `normalized_shape` nor `hidden_size` (deprecated) args are provided")
            normalized_shape = hidden_size
        if not isinstance(normalized_shape, int):
            normalized_shape = tuple(normalized_shape)
            assert len(normalized_shape) == 1, "only the last dimension is normalized"
            normalized_shape = normalized_shape[0]
        dtype = kwargs.pop("dtype", None) or params_dtype or torch.get_default_dtype()
        self.eps, self.zero_centered_gamma = eps, zero_centered_gamma
        self.sequence_parallel = sequence_parallel
        init = torch.zeros if zero_centered_gamma else torch.ones
        self.weight = torch.nn.Parameter(init(normalized_shape, dtype=dtype, device=device))
        if self._layernorm:
            self.bias = torch.nn.Parameter(torch.zeros(normalized_shape, dtype=dtype, device=device))
        else:
            self.register_parameter("bias", None)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        w = self.weight if self.weight.dtype in (x.dtype, torch.float32) else self.weight.to(x.dtype)
        b = None
        if self._layernorm:
            b = self.bias if self.bias.dtype == w.dtype else self.bias.to(w.dtype)
        return _Norm.apply(x, w, b, self.eps, self.zero_centered_gamma)


class LayerNorm(_NormModule):
    """Reference `te.LayerNorm` (pytorch/module/layernorm.py): y = (x - E[x]) / sqrt(Var[x] + eps) *
    gamma + beta over the last dimension, gamma stored zero-centered when `zero_centered_gamma`."""
    _layernorm = True


class RMSNorm(_NormModule):
    """Reference `te.RMSNorm` (pytorch/module/rmsnorm.py): y = x / sqrt(mean(x^2) + eps) * gamma."""
    _layernorm = False
