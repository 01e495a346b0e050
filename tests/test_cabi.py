"""CPU tests of the C-ABI boundary: the library loads without a GPU, exports every
symbol include/transformerengine_b200.h declares, and the host-only parts of the
API (handles, shapes, config objects, error convention) behave like the reference
(transformer_engine/common/transformer_engine.cpp:396-1145)."""
import ctypes
import os
import re

import pytest
import torch

from transformerengine_b200 import cabi as C

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "transformerengine_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nvte_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = C.lib()
    names = declared_symbols()
    assert len(names) > 60
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_handle_table_and_params_roundtrip():
    lib = C.lib()
    hs = [C.TensorHandle(C.MXFP8_1D_SCALING) for _ in range(100)]
    assert len({h.ptr for h in hs}) == 100 and all(h.ptr for h in hs)
    t = torch.zeros(6, 4, 32, dtype=torch.uint8)
    h = hs[0]
    h.set(C.kRowwiseData, t, C.kFloat8E4M3)
    assert h.shape() == (6, 4, 32)
    lib.nvte_tensor_numel.restype = ctypes.c_size_t
    lib.nvte_tensor_numel.argtypes = [ctypes.c_void_p]
    assert lib.nvte_tensor_numel(h.ptr) == 768
    lib.nvte_tensor_type.argtypes = [ctypes.c_void_p]
    assert lib.nvte_tensor_type(h.ptr) == C.kFloat8E4M3
    lib.nvte_tensor_scaling_mode.argtypes = [ctypes.c_void_p]
    assert lib.nvte_tensor_scaling_mode(h.ptr) == C.MXFP8_1D_SCALING
    # columnwise-only per-tensor FP8 / NVFP4 report the transposed logical shape
    n = C.TensorHandle(C.NVFP4_1D_SCALING).set(C.kColumnwiseData, torch.zeros(64, 8, dtype=torch.uint8),
                                               C.kFloat4E2M1, (64, 16))
    assert n.shape() == (16, 64)
    m = C.TensorHandle(C.MXFP8_1D_SCALING).set(C.kColumnwiseData, torch.zeros(64, 32, dtype=torch.uint8),
                                               C.kFloat8E4M3)
    assert m.shape() == (64, 32)
    # flags
    h.set_flag(C.kWithGEMMSwizzledScales, True)
    v, w = ctypes.c_uint8(0), ctypes.c_size_t(0)
    lib.nvte_get_tensor_param_v2(h.ptr, C.kWithGEMMSwizzledScales, ctypes.byref(v), 1, ctypes.byref(w))
    assert v.value == 1 and w.value == 1
    # freed handles are recycled and come back cleared
    old = h.ptr
    h.close()
    h2 = C.TensorHandle()
    assert h2.ptr == old and h2.shape() == (0,)
    # invalid handle: queries return defaults (reference behaviour)
    assert lib.nvte_tensor_type(None) == C.kFloat32


def test_quantization_config_attributes():
    lib = C.lib()
    cfg = C.QuantConfig(force_pow_2_scales=True, amax_epsilon=0.25)
    lib.nvte_get_quantization_config_attribute.argtypes = [
        ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
    w = ctypes.c_size_t(0)
    eps = ctypes.c_float(0)
    lib.nvte_get_quantization_config_attribute(cfg.ptr, C.kQAmaxEpsilon, ctypes.byref(eps), 4, ctypes.byref(w))
    assert eps.value == 0.25 and w.value == 4
    b = ctypes.c_uint8(0)
    lib.nvte_get_quantization_config_attribute(cfg.ptr, C.kQForcePow2Scales, ctypes.byref(b), 1, ctypes.byref(w))
    assert b.value == 1
    # extension attributes live in the 0x1000 range (the reference's NumAttributes stays 10 / 7, so an
    # attribute the reference adds later cannot be read as raw peer pointers) and report their exact size
    hdr = open(HEADER).read()
    assert "kNVTEQuantizationConfigNumAttributes = 10" in hdr and "kNVTEMatmulConfigNumAttributes = 7" in hdr
    assert C.kQB200PeerRowwiseData >= 0x1000 and C.kMOutputRowBlocks >= 0x1000
    lib.nvte_get_quantization_config_attribute(cfg.ptr, C.kQB200PeerRowwiseData, None, 0, ctypes.byref(w))
    assert w.value == 72
    peers = (ctypes.c_uint64 * 9)(2, 0x1000, 0x2000)
    lib.nvte_set_quantization_config_attribute.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                                           ctypes.c_size_t]
    lib.nvte_set_quantization_config_attribute(cfg.ptr, C.kQB200PeerRowwiseData, peers, 72)
    back = (ctypes.c_uint64 * 9)()
    lib.nvte_get_quantization_config_attribute(cfg.ptr, C.kQB200PeerRowwiseData, back, 72, ctypes.byref(w))
    assert list(back)[:3] == [2, 0x1000, 0x2000]


def test_reference_signature_of_after_reduction_is_exported():
    """recipe.h:71-74 is inside `extern "C"`: the unmangled symbol must exist (the reference's
    csrc/extensions/recipe.cpp binds it), next to the plain-array form for FFI callers."""
    lib = C.lib()
    assert hasattr(lib, "nvte_delayed_scaling_recipe_amax_and_scale_update_after_reduction")
    assert hasattr(lib, "nvte_b200_delayed_scaling_update_after_reduction")


def test_errors_surface_as_python_exceptions_without_gpu():
    """No CPU fallback: a compute call on a GPU-less box fails loudly."""
    x = torch.zeros(32, 32, dtype=torch.bfloat16)
    out = C.TensorHandle(C.MXFP8_1D_SCALING)
    out.set(C.kRowwiseData, torch.zeros(32, 32, dtype=torch.uint8), C.kFloat8E4M3)
    out.set(C.kRowwiseScaleInv, torch.zeros(128, 4, dtype=torch.uint8), C.kFloat8E8M0)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="sm_100"):
        C.check(C.lib().nvte_quantize_v2_status(C.plain(x).ptr, out.ptr, None, None))
    # invalid handle
    with pytest.raises(RuntimeError, match="Invalid tensor"):
        C.check(C.lib().nvte_quantize_v2_status(None, out.ptr, None, None))


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "transformerengine_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("test oracle", ""), os.path.join(dirpath, f)


def test_grouped_tensor_handles_and_cpp_helper_symbols():
    """Grouped-tensor views (reference transformer_engine.h:536-644) are host-side handles; the
    C++ helpers of libtransformer_engine.version:5-19 are exported with the reference's mangling."""
    import subprocess
    lib = C.lib()
    vp, sz = ctypes.c_void_p, ctypes.c_size_t
    lib.nvte_create_grouped_tensor.restype = vp
    lib.nvte_create_grouped_tensor.argtypes = [ctypes.c_int, sz, C.NVTEShape]
    lib.nvte_destroy_grouped_tensor.argtypes = [vp]
    lib.nvte_set_grouped_tensor_param.argtypes = [vp, ctypes.c_int, vp, sz]
    lib.nvte_get_grouped_tensor_param.argtypes = [vp, ctypes.c_int, vp, sz, ctypes.POINTER(sz)]
    lib.nvte_grouped_tensor_num_tensors.restype = sz
    lib.nvte_grouped_tensor_num_tensors.argtypes = [vp]
    lib.nvte_grouped_tensor_type.argtypes = [vp]
    lib.nvte_grouped_tensor_scaling_mode.argtypes = [vp]
    lib.nvte_get_grouped_tensor_logical_shape.restype = C.NVTEShape
    lib.nvte_get_grouped_tensor_logical_shape.argtypes = [vp]
    g = lib.nvte_create_grouped_tensor(C.MXFP8_1D_SCALING, 4, C.make_shape((512, 256)))
    assert g and lib.nvte_grouped_tensor_num_tensors(g) == 4
    assert lib.nvte_grouped_tensor_scaling_mode(g) == C.MXFP8_1D_SCALING
    s = lib.nvte_get_grouped_tensor_logical_shape(g)
    assert (s.ndim, s.data[0], s.data[1]) == (2, 512, 256)
    t = torch.zeros(512, 256, dtype=torch.uint8)
    b = C.basic_tensor(t, C.kFloat8E4M3)
    lib.nvte_set_grouped_tensor_param(g, 0, ctypes.byref(b), ctypes.sizeof(b))
    assert lib.nvte_grouped_tensor_type(g) == C.kFloat8E4M3
    back, w = C.NVTEBasicTensor(), sz(0)
    lib.nvte_get_grouped_tensor_param(g, 0, ctypes.byref(back), ctypes.sizeof(back), ctypes.byref(w))
    assert back.data_ptr == t.data_ptr() and w.value == ctypes.sizeof(back)
    flag = ctypes.c_uint8(1)
    lib.nvte_set_grouped_tensor_param(g, 10, ctypes.byref(flag), 1)
    out = ctypes.c_uint8(0)
    lib.nvte_get_grouped_tensor_param(g, 10, ctypes.byref(out), 1, ctypes.byref(w))
    assert out.value == 1 and w.value == 1
    lib.nvte_destroy_grouped_tensor(g)
    assert lib.nvte_get_num_compute_streams() == 1
    syms = subprocess.run(["nm", "-D", C.LIB_PATH], capture_output=True, text=True).stdout
    for mangled in ("_ZN18transformer_engine4cuda7sm_archEi", "_ZN18transformer_engine4cuda8sm_countEi",
                    "_ZN18transformer_engine4cuda14current_deviceEv",
                    "_ZN18transformer_engine4cuda18supports_multicastEi",
                    "_ZN18transformer_engine4cuda21stream_priority_rangeEPiS1_i",
                    "_ZN18transformer_engine4cuda14cublas_versionEv",
                    "_ZN18transformer_engine10typeToSizeENS_5DTypeE"):
        assert mangled in syms, mangled


def test_extension_structs_have_the_same_layout_in_c_and_in_the_ctypes_binding(tmp_path):
    """The extension structs travel by value with an exact-size check: the ctypes mirrors in cabi.py must have the
    header's size and field offsets (compiled here with gcc as C, which also proves the header is plain C)."""
    import ctypes
    import subprocess
    from transformerengine_b200 import cabi as C
    structs = {"NVTEB200SymmetricExchange": C.SymmetricExchangeStruct, "NVTEB200ArrivalFlags": C.ArrivalFlagsStruct,
               "NVTEB200ScatterOutput": C.ScatterOutputStruct}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{ROOT}/include/transformerengine_b200.h"',
             'int main(void) {']
    for name, cls in structs.items():
        lines.append(f'  printf("{name} size %zu\\n", sizeof({name}));')
        for field, _ in cls._fields_:
            lines.append(f'  printf("{name} {field} %zu\\n", offsetof({name}, {field}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-I/usr/local/cuda/include", "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    got = {tuple(l.split()[:2]): int(l.split()[2]) for l in out.splitlines()}
    for name, cls in structs.items():
        assert got[(name, "size")] == ctypes.sizeof(cls), name
        for field, _ in cls._fields_:
            assert got[(name, field)] == getattr(cls, field).offset, (name, field)


def test_group_entry_size_is_exported_and_fits_what_the_python_side_allocates():
    """The member table of a grouped GEMM launch: ops sizes the workspace from the library's own figure (a stale
    constant once made 64-expert lists fall back to one launch per member without any test noticing)."""
    import ctypes
    from transformerengine_b200 import cabi as C
    from transformerengine_b200 import ops
    fn = C.lib().nvte_b200_group_entry_bytes
    fn.restype = ctypes.c_size_t
    assert fn() >= 512 and fn() % 128 == 0
    assert ops._group_entry_bytes() == fn()
