"""kima_b200 — B200-native batched likelihood engine for kima's RVmodel hot path."""
from .data import RVData
from .flags import ModelFlags, layout_from_flags
from .engine import LikelihoodEngine
from ._lib import KimaCudaError, load_library

__all__ = ["RVData", "ModelFlags", "layout_from_flags", "LikelihoodEngine", "KimaCudaError",
           "load_library"]
