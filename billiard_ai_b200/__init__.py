"""billiard_ai_b200: B200-native batched strike evaluation for billiard-ai (one hot path).

The product is libbilliard_cuda.so (hand-written sm_100a CUDA behind the C ABI of include/billiard_cuda.h).
This package is the Python host-side binding used by tests and bench.py; it has no CPU path and raises
when the CUDA library is missing.
"""
from .engine import (  # noqa: F401
    BcuError,
    Engine,
    LIB_PATH,
    F_COMPLETED,
    F_VALID,
    F_OVERFLOW,
    F_NOEVENT,
    load_library,
)
