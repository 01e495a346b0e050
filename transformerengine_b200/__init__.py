"""B200-native low-precision Linear path (drop-in for Transformer Engine's nvte_* C API).

`import transformerengine_b200 as te` gives the reference's user-facing names for this path:
`te.Linear`, `te.GroupedLinear`, `te.LayerNormLinear`, `te.LayerNorm`, `te.RMSNorm`, `te.autocast` / `te.fp8_autocast`, `te.recipe.*` (reference
transformer_engine/pytorch/__init__.py:21-51). Everything below them is an `nvte_*` call
into `libtransformer_engine.so` (hand-written sm_100a CUDA); there is no fallback.
"""
from . import cabi  # noqa: F401
from . import recipe  # noqa: F401
from .graph import make_graphed_callables  # noqa: F401
from .grouped_linear import GroupedLinear  # noqa: F401
from .linear import LayerNormLinear, Linear  # noqa: F401
from .normalization import LayerNorm, RMSNorm  # noqa: F401
from .quantization import FP8GlobalStateManager, autocast, fp8_autocast  # noqa: F401
from .quantizers import (Float8CurrentScalingQuantizer, Float8Quantizer, MXFP8Quantizer,  # noqa: F401
                         NVFP4Quantizer)

__version__ = "0.2.0"
