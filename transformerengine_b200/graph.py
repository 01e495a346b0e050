"""`make_graphed_callables`: CUDA-graph a module's forward and backward under `autocast`.

Mirrors the reference's `te.make_graphed_callables` (pytorch/graph.py:60-1100) for the part this
path needs: a `te.Linear` step at small shapes is bound by host launch latency (a dozen `nvte_*`
calls of a few microseconds of GPU work each), so the whole forward and the whole backward --
casts, GEMMs, the fused delayed-scaling amax/scale update -- are captured once and replayed.
Every kernel on the path is capture-safe: nothing reads device memory on the host, amax / scale /
history live in static tensors, skip flags are device tensors.
"""
from typing import Optional

import torch

from .quantization import autocast
from .recipe import Recipe


class _AutocastModule(torch.nn.Module):
    def __init__(self, module, enabled, recipe, group):
        super().__init__()
        self.module, self._enabled, self._recipe, self._group = module, enabled, recipe, group

    def forward(self, *args):
        with autocast(enabled=self._enabled, recipe=self._recipe, amax_reduction_group=self._group):
            return self.module(*args)


def make_graphed_callables(modules, sample_args, num_warmup_iters: int = 3, allow_unused_input: bool = False,
                           enabled: bool = False, recipe: Optional[Recipe] = None,
                           amax_reduction_group=None, pool=None, **deprecated):
    """Graphed versions of `modules` (one module or a tuple), run inside `autocast(enabled, recipe)`.

    Same call shape as the reference (`fp8_enabled` / `fp8_recipe` / `fp8_group` are accepted as the
    deprecated spellings). The returned callables take the sample arguments' shapes and dtypes.
    """
    enabled = deprecated.pop("fp8_enabled", enabled)
    recipe = deprecated.pop("fp8_recipe", recipe)
    amax_reduction_group = deprecated.pop("fp8_group", amax_reduction_group)
    assert not deprecated, f"unsupported arguments: {sorted(deprecated)}"
    single = not isinstance(modules, (tuple, list))
    mods = (modules,) if single else tuple(modules)
    args = (sample_args,) if single else tuple(sample_args)
    wrapped = tuple(_AutocastModule(m, enabled, recipe, amax_reduction_group) for m in mods)
    # Warm-up and capture run the backward on uninitialised grad_outputs: their amaxes must not
    # stay in the delayed-scaling histories (reference graph.py save_fp8_tensors / restore_fp8_tensors).
    from .linear import Linear
    linears = [m for mod in mods for m in mod.modules() if isinstance(m, Linear)]
    saved = [{d: (st.scale.clone(), st.amax_history.clone()) for d, st in lin._delayed.items() if d in (True, False)}
             for lin in linears]
    graphed = torch.cuda.make_graphed_callables(wrapped, args, num_warmup_iters=num_warmup_iters,
                                                allow_unused_input=allow_unused_input, pool=pool)
    for lin, snap in zip(linears, saved):
        for d in (True, False):
            st = lin._delayed.get(d)
            if st is None:
                continue
            if d in snap:      # in place: the captured kernels hold these tensors' addresses
                st.scale.copy_(snap[d][0])
                st.amax_history.copy_(snap[d][1])
            else:
                st.scale.fill_(1.0)
                st.amax_history.zero_()
    return graphed[0] if single else graphed
