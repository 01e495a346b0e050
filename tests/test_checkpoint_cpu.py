"""Delayed-scaling checkpoint state (`get_extra_state` / `set_extra_state`) in the reference's format
(reference module/base.py:1237-1355): a uint8 tensor holding a pickled dict. CPU only: the (de)serialisation and the
mapping of the reference's recipe classes; the module round trip on a GPU is in tests/test_linear_gpu.py."""
import pickle
import sys
import types

import pytest
import torch

from transformerengine_b200 import checkpoint, recipe


class _State:   # stands in for quantization.DelayedScalingState (CPU tensors)
    def __init__(self, n, hist, seed):
        g = torch.Generator().manual_seed(seed)
        self.scale = torch.rand(n, generator=g) + 0.5
        self.amax_history = torch.rand(hist, n, generator=g)


def _states(num, hist=4, seed=0):
    return [{True: _State(3, hist, seed + 2 * i), False: _State(2, hist, seed + 2 * i + 1)} for i in range(num)]


def test_no_fp8_state_is_an_empty_byte_tensor():
    t = checkpoint.pack(None, [])
    assert t.dtype == torch.uint8 and t.numel() == 0
    assert checkpoint.unpack(t) is None and checkpoint.unpack(None) is None


@pytest.mark.parametrize("num", [1, 3])
def test_round_trip_keeps_every_scale_and_history(num):
    rec = recipe.DelayedScaling(amax_history_len=4, fp8_format=recipe.Format.HYBRID)
    src = _states(num)
    blob = checkpoint.pack(rec, src)
    assert blob.dtype == torch.uint8 and blob.dim() == 1 and blob.numel() > 0
    st = checkpoint.unpack(blob)
    assert isinstance(st["recipe"], recipe.DelayedScaling) and st["recipe"].amax_history_len == 4
    assert st["scale_fwd"].shape == (3 * num,) and st["amax_history_bwd"].shape == (4, 2 * num)
    dst = _states(num, seed=100)
    assert checkpoint.load_into(st, dst, "test")
    for a, b in zip(src, dst):
        for d in (True, False):
            assert torch.equal(a[d].scale, b[d].scale) and torch.equal(a[d].amax_history, b[d].amax_history)


def test_shape_mismatch_warns_and_leaves_the_state_alone():
    st = checkpoint.unpack(checkpoint.pack(recipe.DelayedScaling(amax_history_len=4), _states(1)))
    dst = _states(1, hist=8, seed=7)
    before = dst[0][True].scale.clone()
    with pytest.warns(UserWarning):
        assert not checkpoint.load_into(st, dst, "test")
    assert torch.equal(dst[0][True].scale, before)


def test_state_pickled_by_the_reference_loads_without_transformer_engine_installed():
    """A blob whose recipe object belongs to `transformer_engine.common.recipe` (what the reference writes): the
    unpickler maps the class to ours of the same name."""
    mod = types.ModuleType("transformer_engine.common.recipe")

    class DelayedScaling:   # the reference's dataclass, reduced to what pickling needs
        def __init__(self):
            self.margin, self.amax_history_len, self.fp8_format = 0, 4, recipe.Format.HYBRID
            self.amax_compute_algo, self.scaling_factor_compute_algo, self.reduce_amax = "max", None, True
            self.fp8_dpa = self.fp8_mha = False
            self.backward_override = None

    DelayedScaling.__module__, DelayedScaling.__qualname__ = "transformer_engine.common.recipe", "DelayedScaling"
    mod.DelayedScaling = DelayedScaling
    pkgs = {"transformer_engine": types.ModuleType("transformer_engine"),
            "transformer_engine.common": types.ModuleType("transformer_engine.common"),
            "transformer_engine.common.recipe": mod}
    saved = {k: sys.modules.get(k) for k in pkgs}
    sys.modules.update(pkgs)
    try:
        src = _states(1)
        state = {"recipe": DelayedScaling(), "scale_fwd": src[0][True].scale, "amax_history_fwd": src[0][True].amax_history,
                 "scale_bwd": src[0][False].scale, "amax_history_bwd": src[0][False].amax_history,
                 "extra_fp8_variables": {"fp8_checkpoint": True, "num_gemms": 1}}
        blob = torch.frombuffer(bytearray(pickle.dumps(state)), dtype=torch.uint8)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    st = checkpoint.unpack(blob)
    assert type(st["recipe"]) is recipe.DelayedScaling and st["recipe"].amax_history_len == 4
    dst = _states(1, seed=50)
    assert checkpoint.load_into(st, dst, "test") and torch.equal(dst[0][True].scale, src[0][True].scale)


def test_dict_written_by_an_earlier_version_still_loads():
    src = _states(1)
    old = {"key": (4, recipe.Format.HYBRID), "scale_fwd": src[0][True].scale, "amax_history_fwd": src[0][True].amax_history,
           "scale_bwd": src[0][False].scale, "amax_history_bwd": src[0][False].amax_history}
    dst = _states(1, seed=9)
    assert checkpoint.load_into(checkpoint.unpack(old), dst, "test")
    assert torch.equal(dst[0][False].amax_history, src[0][False].amax_history)
