"""expann_b200 — B200-native (sm_100a) implementation of expANN's query hot path behind a C ABI.
See DESIGN.md. The CUDA library (expann_b200/libexpann_gpu.so) is mandatory: nothing here computes on the CPU."""
from ._lib import METRIC_IP, METRIC_L2, ExpannError, LIB_PATH, build_library, load  # noqa: F401
from .engine import (BruteForceEngine, GpuEngine, GraphEngine, QuantScanEngine, pair_distances,  # noqa: F401
                     QUANT_SIMPLE_U8, QUANT_RANGED_Q8)

__all__ = ["BruteForceEngine", "GpuEngine", "GraphEngine", "QuantScanEngine", "pair_distances", "QUANT_SIMPLE_U8", "QUANT_RANGED_Q8", "METRIC_L2", "METRIC_IP", "ExpannError", "build_library", "load", "LIB_PATH"]
