"""Generate tests/golden/*: golden vectors for the oracle, made by the REFERENCE itself.

Run in the build container (needs /root/reference; never on the GPU box):
    python -m oracle.gen_golden
Two sources, both the reference's own code on seeded inputs:
  * its C++ host references compiled in place (oracle/_ref, `make -C oracle ref`)
    -> SHA-256 of every output buffer (tests/golden/host_ref_hashes.json);
  * its pure-torch NVFP4 quantizer with the random Hadamard transform
    (transformer_engine/pytorch/custom_recipes/quantization_ref_nvfp4.py), imported
    from /root/reference with the package's heavy imports stubbed out
    -> small .npz fixtures (tests/golden/nvfp4_rht_*.npz).
TEST INFRASTRUCTURE ONLY.
"""
import hashlib
import importlib.util
import json
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.environ.get("TE_REFERENCE", "/root/reference")
GOLD = os.path.join(ROOT, "tests", "golden")

from tests.util import bits16, make_input  # noqa: E402

HOST_CASES = [  # (rows, cols, kind, seed)
    (128, 128, "randn", 0), (256, 272, "uniform", 1), (304, 304, "outliers", 2),
    (993, 512, "uniform", 3), (1024, 1024, "randn", 4), (64, 96, "zeros", 5),
]
RHT_CASES = [(128, 128, "randn", 0), (256, 384, "randn", 1), (512, 256, "outliers", 2),
             (64, 128, "uniform", 3)]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load_python_reference():
    """Import quantization_ref_nvfp4.py from the reference with stubbed parents."""
    base = os.path.join(REF, "transformer_engine", "pytorch", "custom_recipes")
    for name in ("transformer_engine", "transformer_engine.pytorch",
                 "transformer_engine.pytorch.custom_recipes"):
        m = types.ModuleType(name)
        m.__path__ = []
        sys.modules[name] = m

    def load(modname, path):
        spec = importlib.util.spec_from_file_location(modname, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
        return mod

    utils = load("transformer_engine.pytorch.custom_recipes.utils", os.path.join(base, "utils.py"))
    qstub = load("transformer_engine.pytorch.custom_recipes.quantization",
                 os.path.join(base, "quantization.py"))
    qt = types.ModuleType("transformer_engine.pytorch.quantized_tensor")

    class Quantizer:  # minimal stand-in for the reference base class (attributes only)
        def __init__(self, rowwise=True, columnwise=True):
            self.rowwise_usage, self.columnwise_usage = rowwise, columnwise
            self.internal = False

    class QuantizedTensorStorage:
        pass

    qt.Quantizer, qt.QuantizedTensorStorage = Quantizer, QuantizedTensorStorage
    sys.modules[qt.__name__] = qt
    cr = sys.modules["transformer_engine.pytorch.custom_recipes"]
    cr.quantization, cr.utils = qstub, utils
    ref = load("transformer_engine.pytorch.custom_recipes.quantization_ref_nvfp4",
               os.path.join(base, "quantization_ref_nvfp4.py"))
    return ref, utils


def gen_host_hashes():
    from oracle import ref_host as R
    out = {}
    for rows, cols, kind, seed in HOST_CASES:
        x = bits16(make_input(rows, cols, kind, seed))
        xf = make_input(rows, cols, kind, seed, dtype=torch.float32).numpy()
        key = f"{rows}x{cols}_{kind}_{seed}"
        e = {}
        for e5 in (False, True):
            r = R.mxfp8_compute_ref(x, e5m2=e5)
            e[f"mxfp8_{'e5m2' if e5 else 'e4m3'}"] = {k: sha(v) for k, v in r.items()}
        e["mxfp8_e4m3_fp32in"] = {k: sha(v) for k, v in R.mxfp8_compute_ref(xf).items()}
        amax, scale, sinv = R.amax_scale(x)
        e["amax_scale"] = [float(amax), float(scale), float(sinv)]
        oc, ot, am = R.fp8_cast_transpose(x, scale)
        e["fp8_e4m3"] = {"data": sha(oc), "transpose": sha(ot), "amax": float(am)}
        oc, ot, am = R.fp8_cast_transpose(xf, 3.0, e5m2=True)
        e["fp8_e5m2_fp32in"] = {"data": sha(oc), "transpose": sha(ot), "amax": float(am)}
        if cols % 16 == 0 and rows % 16 == 0:
            for two_d in (False, True):
                o, s, o_t, s_t = R.nvfp4_compute_ref(x, amax, two_d)
                e[f"nvfp4_{'2d' if two_d else '1d'}"] = {
                    "data": sha(o), "scale": sha(s), "data_t": sha(o_t), "scale_t": sha(s_t)}
        out[key] = e
    rng = np.random.default_rng(11)
    for M, K in [(128, 4), (256, 8), (384, 36)]:
        sc = rng.integers(0, 255, (M, K), dtype=np.uint8)
        out[f"swizzle_{M}x{K}"] = {"row": sha(R.swizzle(sc, M, K, True)),
                                   "col": sha(R.swizzle(sc.T.copy(), M, K, False))}
    with open(os.path.join(GOLD, "host_ref_hashes.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    return out


def gen_rht_fixtures():
    ref, utils = load_python_reference()
    for rows, cols, kind, seed in RHT_CASES:
        x = make_input(rows, cols, kind, seed)
        q = ref.NVFP4QuantizerRef(dtype=utils.Fp4Formats.E2M1, rowwise=True, columnwise=True,
                                  pow_2_scales=False, eps=0.0, quant_tile_shape=(1, 16),
                                  with_rht=True, with_random_sign_mask=True)
        qx, sx, qx_t, sx_t, amax_row, amax_col = q._quantize(x)
        rht_t = q._apply_rht(x.t().contiguous())
        np.savez_compressed(
            os.path.join(GOLD, f"nvfp4_rht_{rows}x{cols}_{kind}_{seed}.npz"),
            data=qx.numpy(), scale=sx.view(torch.uint8).numpy(),
            data_t=qx_t.numpy(), scale_t=sx_t.view(torch.uint8).numpy(),
            amax_row=amax_row.numpy(), amax_col=amax_col.numpy(),
            rht_t_sha=np.array(sha(bits16(rht_t))))
    # the fixed sign vector and 16x16 matrix themselves (reference nvfp4_tensor.py:48-110)
    H = ref.NVFP4QuantizerRef._build_hadamard_matrix(16, torch.device("cpu"), torch.float32, True)
    np.save(os.path.join(GOLD, "rht_matrix_16x16.npy"), (H * 0.25).numpy())


NORM_CASES = [  # (rows, cols, layernorm, zero_centered_gamma, seed)
    (32, 256, True, False, 0), (32, 256, False, False, 1), (64, 512, True, True, 2),
    (48, 128, False, True, 3),
]


def gen_norm_fixtures():
    """LayerNorm / RMSNorm: the reference's host references (tests/cpp/operator/
    test_normalization.h, compiled in place into oracle/_ref) on seeded bf16 inputs."""
    from oracle import ref_host as R
    for rows, cols, ln, zc, seed in NORM_CASES:
        x = bits16(make_input(rows, cols, "randn", seed))
        dz = bits16(make_input(rows, cols, "uniform", seed + 100))
        add = bits16(make_input(rows, cols, "randn", seed + 200))
        g = torch.Generator().manual_seed(seed + 300)
        gamma = bits16((torch.rand(cols, generator=g) * 2 - (1.0 if zc else 0.0)).to(torch.bfloat16))
        beta = bits16((torch.rand(cols, generator=g) - 0.5).to(torch.bfloat16))
        eps = 1e-5
        mu, rs = R.norm_stats(x, eps, ln)
        y, amax = R.norm_output(x, gamma, beta, mu, rs, ln, zc)
        dx, dg, db = R.norm_backward(dz, x, mu, rs, gamma, ln, add=None if ln else add, zero_centered=zc)
        np.savez_compressed(
            os.path.join(GOLD, f"norm_{'ln' if ln else 'rms'}_{rows}x{cols}_{seed}.npz"),
            x=x, dz=dz, add=add, gamma=gamma, beta=beta, eps=np.float32(eps), layernorm=np.bool_(ln),
            zero_centered=np.bool_(zc), mu=mu, rsigma=rs, y=y, amax=np.float32(amax), dx=dx, dgamma=dg, dbeta=db)


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    gen_host_hashes()
    gen_rht_fixtures()
    gen_norm_fixtures()
    print("golden vectors written to", GOLD)
