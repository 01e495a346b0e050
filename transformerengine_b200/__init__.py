"""B200-native low-precision Linear path (drop-in for Transformer Engine's nvte_* C API)."""
from . import cabi  # noqa: F401

__version__ = "0.1.0"
