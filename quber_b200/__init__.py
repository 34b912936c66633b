"""quber_b200 — B200-native (sm_100a) implementation of QuBER's maskrefiner inference hot path.

Host code is Python; the arithmetic lives in hand-written CUDA behind the C ABI of
include/quber_b200.h (libquber_b200.so).  There is no CPU or PyTorch fallback.
"""
__version__ = "0.1.0"
