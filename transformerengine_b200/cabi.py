"""ctypes binding of the C-ABI library (`libtransformer_engine.so`).

This is the only place Python touches native code: torch supplies device
memory and streams, every operation on the path is an `nvte_*` entry point of
`include/transformerengine_b200.h`. There is no fallback: if the library is
missing or the device is not sm_100, calls raise.
"""
import ctypes
import os
import threading
from typing import Optional, Sequence

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtransformer_engine.so")

# NVTEDType (include/transformerengine_b200.h)
kByte, kInt16, kInt32, kInt64, kFloat32, kFloat16, kBFloat16 = range(7)
kFloat8E4M3, kFloat8E5M2, kFloat8E8M0, kFloat4E2M1 = 7, 8, 9, 10
# NVTETensorParam
(kRowwiseData, kColumnwiseData, kScale, kAmax, kRowwiseScaleInv, kColumnwiseScaleInv,
 kColumnwiseAmax, kWithGEMMSwizzledScales, kRowScaledNVFP4, kNVFP4E4M3Max) = range(10)
# NVTEScalingMode
DELAYED_TENSOR_SCALING, MXFP8_1D_SCALING, NVFP4_1D_SCALING = 0, 1, 4
# NVTEQuantizationConfigAttribute
(kQForcePow2Scales, kQAmaxEpsilon, kQNoopTensor, kQBlockScaleFormat, kQRNGState,
 kQNVFP42D, kQStochasticRounding, kQUseFastMath) = range(8)
kQB200PeerRowwiseData = 0x1000   # extension range (include/transformerengine_b200.h)
kQB200SymmetricExchange = 0x1001
# NVTEMatmulConfigAttribute
(kMBiasTensor, kMDBiasTensor, kMWithGELU, kMWithDGELU, kMEpilogueAux, kMUseSplitAccumulator,
 kMSMCount) = range(7)
kMOutputRowBlocks = 0x1000        # extension range
kMArrivalFlags = 0x1001
kMScatterOutput = 0x1002


class SymmetricExchangeStruct(ctypes.Structure):
    """NVTEB200SymmetricExchange (include/transformerengine_b200.h)."""
    _fields_ = [("world", ctypes.c_uint32), ("rank", ctypes.c_uint32), ("rows_per_flag", ctypes.c_uint32),
                ("flags_per_rank", ctypes.c_uint32), ("gather", ctypes.c_uint64 * 8), ("flags", ctypes.c_uint64 * 8),
                ("amax_slots", ctypes.c_uint64 * 8), ("step", ctypes.c_uint64), ("done", ctypes.c_uint64),
                ("error", ctypes.c_uint64)]


class ArrivalFlagsStruct(ctypes.Structure):
    """NVTEB200ArrivalFlags (include/transformerengine_b200.h)."""
    _fields_ = [("flags", ctypes.c_uint64), ("step", ctypes.c_uint64), ("error", ctypes.c_uint64),
                ("world", ctypes.c_uint32), ("rank", ctypes.c_uint32), ("rows_per_flag", ctypes.c_uint32),
                ("flags_per_rank", ctypes.c_uint32), ("push_unit_bytes", ctypes.c_uint32),
                ("reserved", ctypes.c_uint32), ("push_dst", ctypes.c_uint64 * 8),
                ("push_flags", ctypes.c_uint64 * 8), ("push_mc_dst", ctypes.c_uint64),
                ("push_mc_flags", ctypes.c_uint64), ("tile_counter", ctypes.c_uint64),
                ("push_src2", ctypes.c_uint64), ("push_mc_dst2", ctypes.c_uint64),
                ("push_unit2_bytes", ctypes.c_uint32), ("push_bytes2_per_flag", ctypes.c_uint32),
                ("advance_step", ctypes.c_uint32), ("reserved2", ctypes.c_uint32),
                ("ready", ctypes.c_uint64), ("ready_mc", ctypes.c_uint64)]

class ScatterOutputStruct(ctypes.Structure):
    """NVTEB200ScatterOutput (include/transformerengine_b200.h)."""
    _fields_ = [("dst", ctypes.c_uint64 * 8), ("world", ctypes.c_uint32), ("rank", ctypes.c_uint32),
                ("rows_per_rank", ctypes.c_uint32), ("reserved", ctypes.c_uint32)]


_TORCH_TO_NVTE = {
    torch.uint8: kByte, torch.int16: kInt16, torch.int32: kInt32, torch.int64: kInt64,
    torch.float32: kFloat32, torch.float16: kFloat16, torch.bfloat16: kBFloat16,
    torch.float8_e4m3fn: kFloat8E4M3, torch.float8_e5m2: kFloat8E5M2,
}


class NVTEShape(ctypes.Structure):
    _fields_ = [("data", ctypes.c_size_t * 15), ("ndim", ctypes.c_size_t)]


class NVTEBasicTensor(ctypes.Structure):
    _fields_ = [("data_ptr", ctypes.c_void_p), ("dtype", ctypes.c_int), ("shape", NVTEShape)]


_lib = None


def lib() -> ctypes.CDLL:
    """Load the C-ABI library (built by `transformerengine_b200.build_ext`)."""
    global _lib
    if _lib is None:
        path = os.environ.get("NVTE_B200_LIB", LIB_PATH)   # dev aid: a variant build of the same library
        if not os.path.exists(path):
            raise RuntimeError(
                f"{path} is missing: build it with `python -m transformerengine_b200.build_ext` "
                "(there is no non-CUDA fallback)")
        l = ctypes.CDLL(path)
        vp, sz, ci = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int
        l.nvte_create_tensor.restype = vp
        l.nvte_create_tensor.argtypes = [ci]
        l.nvte_destroy_tensor.argtypes = [vp]
        l.nvte_set_tensor_param_v2.argtypes = [vp, ci, vp, sz]
        l.nvte_get_tensor_param_v2.argtypes = [vp, ci, vp, sz, ctypes.POINTER(sz)]
        l.nvte_tensor_shape.restype = NVTEShape
        l.nvte_tensor_shape.argtypes = [vp]
        l.nvte_create_quantization_config.restype = vp
        l.nvte_destroy_quantization_config.argtypes = [vp]
        l.nvte_set_quantization_config_attribute.argtypes = [vp, ci, vp, sz]
        l.nvte_create_matmul_config.restype = vp
        l.nvte_destroy_matmul_config.argtypes = [vp]
        l.nvte_set_matmul_config_attribute.argtypes = [vp, ci, vp, sz]
        l.nvte_b200_last_error.restype = ctypes.c_char_p
        l.nvte_b200_set_tensor_param_flat.restype = ci
        l.nvte_b200_set_tensor_param_flat.argtypes = [vp, ci, vp, ci, ci, sz, sz, sz, sz]
        l.nvte_b200_create_plain_tensor.restype = vp
        l.nvte_b200_create_plain_tensor.argtypes = [vp, ci, ci, sz, sz, sz, sz]
        l.nvte_is_non_tn_fp8_gemm_supported.restype = ci
        f32 = ctypes.c_float
        sigs = {
            "nvte_quantize_v2_status": [vp, vp, vp, vp],
            "nvte_dequantize_status": [vp, vp, vp],
            "nvte_quantize_dbias_status": [vp, vp, vp, vp, vp],
            "nvte_b200_quantize_dbias_static_scratch": [vp, vp, vp, vp],
            "nvte_b200_reduce_partials": [vp, ctypes.c_uint32, sz, ci, vp, vp],
            "nvte_multi_tensor_quantize_status": [ctypes.POINTER(vp), ctypes.POINTER(vp), vp, sz, vp],
            "nvte_layernorm_fwd_status": [vp, vp, vp, f32, vp, vp, vp, vp, ci, ci, vp],
            "nvte_rmsnorm_fwd_status": [vp, vp, f32, vp, vp, vp, ci, ci, vp],
            "nvte_layernorm_bwd_status": [vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, vp],
            "nvte_rmsnorm_bwd_status": [vp, vp, vp, vp, vp, vp, vp, ci, ci, vp],
            "nvte_rmsnorm_bwd_add_status": [vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, vp],
            "nvte_compute_amax_with_config_status": [vp, vp, vp, vp],
            "nvte_compute_scale_from_amax_status": [vp, vp, vp],
            "nvte_delayed_scaling_recipe_amax_and_scale_update_status":
                [vp, vp, vp, vp, ctypes.c_char_p, ci, f32, vp],
            "nvte_b200_delayed_scaling_update_after_reduction_status":
                [vp, ctypes.POINTER(vp), ctypes.POINTER(vp), sz, ctypes.c_char_p, ci, f32, vp],
            "nvte_nvfp4_compute_per_tensor_scale_status":
                [vp, ctypes.c_bool, vp, ctypes.c_bool, f32, vp, vp],
            "nvte_hadamard_transform_status": [vp, vp, ci, ci, vp],
            "nvte_hadamard_transform_amax_status": [vp, vp, ci, ci, vp],
            "nvte_quantize_with_hadamard_transform_status": [vp, vp, vp, vp, vp],
            "nvte_swizzle_scaling_factors_status": [vp, vp, vp],
            "nvte_unswizzle_scaling_factors_status": [vp, vp, vp],
            "nvte_cublas_gemm_v2_status":
                [ci, ci, ctypes.POINTER(f32), vp, vp, ctypes.POINTER(f32), vp, vp, vp, vp, vp],
            "nvte_multi_tensor_gemm_status":
                [ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(vp),
                 ctypes.POINTER(vp), ci, ci, ci, ci, ctypes.POINTER(vp), ci, ci, ci, vp],
            "nvte_multi_tensor_swizzle_scaling_factors_status":
                [ctypes.POINTER(vp), ctypes.POINTER(vp), sz, vp],
        }
        for name, args in sigs.items():
            if hasattr(l, name):
                fn = getattr(l, name)
                fn.restype = ci
                fn.argtypes = args
        _lib = l
    return _lib


def check(status: int) -> None:
    if status != 0:
        raise RuntimeError(lib().nvte_b200_last_error().decode("utf-8", "replace"))


def stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    if stream is not None:
        return stream.cuda_stream
    # torch.cuda.current_stream() costs ~15 us of Python per call (device-index plumbing); the raw
    # getter is a single C call and every operator needs the stream
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())


def make_shape(dims: Sequence[int]) -> NVTEShape:
    s = NVTEShape()
    for i, d in enumerate(dims):
        s.data[i] = int(d)
    s.ndim = len(dims)
    return s


def basic_tensor(t: torch.Tensor, dtype: Optional[int] = None,
                 shape: Optional[Sequence[int]] = None) -> NVTEBasicTensor:
    b = NVTEBasicTensor()
    b.data_ptr = t.data_ptr()
    b.dtype = _TORCH_TO_NVTE[t.dtype] if dtype is None else dtype
    b.shape = make_shape(tuple(t.shape) if shape is None else shape)
    return b


class TensorHandle:
    """Owning wrapper of one NVTETensor handle (a non-owning view of torch buffers)."""

    def __init__(self, scaling_mode: int = DELAYED_TENSOR_SCALING):
        self._l = lib()
        self.ptr = self._l.nvte_create_tensor(scaling_mode)
        self._keep = []  # keep the torch tensors alive while the view exists

    def set(self, param: int, t: torch.Tensor, dtype: Optional[int] = None,
            shape: Optional[Sequence[int]] = None) -> "TensorHandle":
        assert t.is_contiguous(), "NVTETensor views need contiguous buffers"
        dims = tuple(t.shape) if shape is None else tuple(shape)
        if len(dims) <= 4:   # scalars only: no ctypes structure to build
            d = dims + (0,) * (4 - len(dims))
            if self._l.nvte_b200_set_tensor_param_flat(
                    self.ptr, param, t.data_ptr(), _TORCH_TO_NVTE[t.dtype] if dtype is None else dtype,
                    len(dims), d[0], d[1], d[2], d[3]) != 0:
                raise RuntimeError(self._l.nvte_b200_last_error().decode("utf-8", "replace"))
        else:
            b = basic_tensor(t, dtype, shape)
            self._l.nvte_set_tensor_param_v2(self.ptr, param, ctypes.byref(b), ctypes.sizeof(b))
        self._keep.append(t)
        return self

    def set_flag(self, param: int, value: bool) -> "TensorHandle":
        v = ctypes.c_uint8(1 if value else 0)
        self._l.nvte_set_tensor_param_v2(self.ptr, param, ctypes.byref(v), 1)
        return self

    def shape(self):
        s = self._l.nvte_tensor_shape(self.ptr)
        return tuple(s.data[i] for i in range(s.ndim))

    def reset_keep(self):
        """Forget the tensors kept alive for earlier `set` calls (for handles that are re-pointed
        every call; the caller holds the current tensors)."""
        self._keep = []

    def close(self):
        if self.ptr:
            self._l.nvte_destroy_tensor(self.ptr)
            self.ptr = None
            self._keep = []

    def __del__(self):
        try:
            self.close()
        except Exception:  # interpreter teardown
            pass


def plain(t: torch.Tensor) -> TensorHandle:
    """Handle of an unquantized (or raw) tensor: rowwise data only."""
    return TensorHandle(DELAYED_TENSOR_SCALING).set(kRowwiseData, t)


_scratch = threading.local()


def scratch_plain(role: str, t: torch.Tensor) -> TensorHandle:
    """`plain(t)` without creating and destroying a handle per call: one reusable handle per (thread, role), pointed
    at `t`. Valid until the next call with the same role on the same thread -- handles are host-side descriptors that a
    launch reads before it returns, so re-pointing one behind an enqueued kernel is safe. (Per thread: the autograd
    engine runs backward on its own threads.)"""
    hs = _scratch.__dict__.setdefault("handles", {})
    h = hs.get(role)
    if h is None:
        h = hs[role] = TensorHandle(DELAYED_TENSOR_SCALING)
    h._keep = []
    return h.set(kRowwiseData, t)


_DEFAULT_QCFG = None


def default_quant_config() -> "QuantConfig":
    """A shared all-defaults quantization config (never mutated): saves building one per call."""
    global _DEFAULT_QCFG
    if _DEFAULT_QCFG is None:
        _DEFAULT_QCFG = QuantConfig()
    return _DEFAULT_QCFG


class QuantConfig:
    def __init__(self, force_pow_2_scales=False, amax_epsilon=0.0, noop: Optional[TensorHandle] = None,
                 nvfp4_2d=False, stochastic_rounding=False, rng_state: Optional[TensorHandle] = None,
                 peer_rowwise=None, symmetric_exchange=None):
        l = lib()
        self._l = l
        self.ptr = l.nvte_create_quantization_config()
        self._keep = (noop, rng_state)

        def set8(attr, val):
            if not val:
                return   # a fresh config already holds the defaults
            v = ctypes.c_uint8(1)
            l.nvte_set_quantization_config_attribute(self.ptr, attr, ctypes.byref(v), 1)

        set8(kQForcePow2Scales, force_pow_2_scales)
        if amax_epsilon != 0.0:
            eps = ctypes.c_float(amax_epsilon)
            l.nvte_set_quantization_config_attribute(self.ptr, kQAmaxEpsilon, ctypes.byref(eps), 4)
        if noop is not None:
            p = ctypes.c_void_p(noop.ptr)
            l.nvte_set_quantization_config_attribute(self.ptr, kQNoopTensor, ctypes.byref(p), 8)
        if rng_state is not None:
            p = ctypes.c_void_p(rng_state.ptr)
            l.nvte_set_quantization_config_attribute(self.ptr, kQRNGState, ctypes.byref(p), 8)
        set8(kQNVFP42D, nvfp4_2d)
        set8(kQStochasticRounding, stochastic_rounding)
        if peer_rowwise:
            # extension: the cast also writes its rowwise bytes to these (peer-mapped) addresses
            arr = (ctypes.c_uint64 * 9)(len(peer_rowwise), *[int(p) for p in peer_rowwise])
            l.nvte_set_quantization_config_attribute(self.ptr, kQB200PeerRowwiseData, arr, 72)
        if symmetric_exchange is not None:
            # extension: amax push / scale wait / cast with arrival flags over peer-mapped memory
            l.nvte_set_quantization_config_attribute(self.ptr, kQB200SymmetricExchange,
                                                     ctypes.byref(symmetric_exchange),
                                                     ctypes.sizeof(symmetric_exchange))

    def __del__(self):
        try:
            if self.ptr:
                self._l.nvte_destroy_quantization_config(self.ptr)
                self.ptr = None
        except Exception:
            pass
