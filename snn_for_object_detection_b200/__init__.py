"""B200-native SODa spiking forward hot path (event binning + Conv/Norm/LIF/Pool over T).

Importing this package loads ``libsoda_b200.so`` (C ABI, ``include/soda_b200.h``); there is no CPU
fallback — a missing library is an ImportError, a non-CUDA tensor a RuntimeError.
"""

from . import _lib  # noqa: F401  (fails loudly when the CUDA library is not built)
from . import ops, events, sharding  # noqa: F401
from .generator import BackboneGen, BlockGen, Head, HeadGen, ListGen, ListState, ModelGen, NeckGen  # noqa: F401
from .modules import *  # noqa: F401,F403
from .soda import SODa  # noqa: F401
from .tiny_yolo import TinyYolo  # noqa: F401

__version__ = "0.1.0"
