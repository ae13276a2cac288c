"""kima_b200 — B200-native batched likelihood engine for kima's RVmodel hot path."""
from .data import RVData
from .flags import ModelFlags, layout_from_flags
from .engine import LikelihoodEngine
from ._lib import KimaCudaError, load_library
from .posterior import PosteriorPredictive, params_from_samples, sample_columns

__all__ = ["RVData", "ModelFlags", "layout_from_flags", "LikelihoodEngine", "KimaCudaError",
           "load_library", "PosteriorPredictive", "params_from_samples", "sample_columns"]
