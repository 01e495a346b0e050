"""Delayed-scaling checkpoint state in the reference's format (module/base.py:1237-1355): `get_extra_state` returns a
uint8 tensor holding a pickled dict

    {"recipe": <recipe object>, "scale_fwd": [n_fwd], "amax_history_fwd": [len, n_fwd],
     "scale_bwd": [n_bwd], "amax_history_bwd": [len, n_bwd], "extra_fp8_variables": {...}}

with CPU tensors (n_fwd = 3 and n_bwd = 2 per GEMM: input / weight / output and grad_output / grad_input; a grouped
module concatenates its members' blocks, reference module/grouped_linear.py `_offsets`), and an empty uint8 tensor when
there is no FP8 state. `unpack` reads that tensor -- also one written by the reference itself: its recipe classes
(`transformer_engine.common.recipe.*`) are mapped to the classes of `transformerengine_b200.recipe` of the same name,
so `transformer_engine` does not have to be importable -- and the plain dict an earlier version of this package wrote.
"""
import io
import pickle
import warnings
from typing import List, Optional

import torch

from . import recipe as _recipe

_REF_RECIPE_MODULES = ("transformer_engine.common.recipe", "transformer_engine.common.recipe.__init__")


class _Unpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module in _REF_RECIPE_MODULES or module.startswith("transformer_engine.common.recipe"):
            if hasattr(_recipe, name):
                return getattr(_recipe, name)
        return super().find_class(module, name)


def pack(recipe, states: List[dict]) -> torch.Tensor:
    """states: one {True: DelayedScalingState, False: DelayedScalingState} per GEMM (empty list: no FP8 state)."""
    if not states:
        return torch.empty(0, dtype=torch.uint8)
    cpu = lambda t: t.detach().to("cpu", copy=True)
    state = {
        "recipe": recipe,
        "scale_fwd": torch.cat([cpu(s[True].scale) for s in states]),
        "amax_history_fwd": torch.cat([cpu(s[True].amax_history) for s in states], dim=1),
        "scale_bwd": torch.cat([cpu(s[False].scale) for s in states]),
        "amax_history_bwd": torch.cat([cpu(s[False].amax_history) for s in states], dim=1),
        "extra_fp8_variables": {"num_gemms": len(states), "fp8_checkpoint": True},
    }
    return torch.frombuffer(bytearray(pickle.dumps(state)), dtype=torch.uint8)


def unpack(state) -> Optional[dict]:
    """The dict stored by `pack` (or by the reference, or by an older version of this package); None when empty."""
    if state is None:
        return None
    if isinstance(state, torch.Tensor):
        if state.numel() == 0:
            return None
        return _Unpickler(io.BytesIO(state.detach().cpu().numpy().tobytes())).load()
    if isinstance(state, io.BytesIO):   # the reference's deprecated format
        state.seek(0)
        return torch.load(state, map_location="cpu")
    if isinstance(state, dict):
        return state or None
    raise RuntimeError("Unsupported checkpoint format.")


def load_into(state: dict, states: List[dict], who: str) -> bool:
    """Copy a stored dict into the live scaling states; False (with a warning) when the shapes do not fit."""
    if not states:
        return False
    n = len(states)
    hist = states[0][True].amax_history.shape[0]
    ok = (state["scale_fwd"].numel() == 3 * n and state["scale_bwd"].numel() == 2 * n and
          tuple(state["amax_history_fwd"].shape) == (hist, 3 * n) and tuple(state["amax_history_bwd"].shape) == (hist, 2 * n))
    if not ok:
        warnings.warn(f"{who}: the checkpoint's delayed-scaling state (amax history "
                      f"{tuple(state['amax_history_fwd'].shape)}) does not fit the module's ({hist}, {3 * n}); ignored")
        return False
    for i, s in enumerate(states):
        s[True].scale.copy_(state["scale_fwd"][3 * i:3 * i + 3])
        s[True].amax_history.copy_(state["amax_history_fwd"][:, 3 * i:3 * i + 3])
        s[False].scale.copy_(state["scale_bwd"][2 * i:2 * i + 2])
        s[False].amax_history.copy_(state["amax_history_bwd"][:, 2 * i:2 * i + 2])
    return True
