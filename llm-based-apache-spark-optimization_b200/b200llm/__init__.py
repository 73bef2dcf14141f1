"""b200llm -- B200-native Llama decode engine (Python host over the libb200llm.so C ABI).

Importing this package loads the CUDA library; there is no CPU or PyTorch fallback.
"""
from . import _lib, geometry, weights  # noqa: F401
from ._lib import B200LLMError  # noqa: F401
from .engine import Engine, GenTiming, SamplingParams  # noqa: F401
from .geometry import Geometry, resolve  # noqa: F401
