"""quber_b200 — B200-native (sm_100a) implementation of QuBER's maskrefiner inference hot path.

Host code is Python; the arithmetic lives in hand-written CUDA behind the C ABI of
include/quber_b200.h (libquber_b200.so).  There is no CPU or PyTorch fallback.
"""
__version__ = "0.1.0"


def register(override=False):
    """Register the meta-architecture, backbones, head and pixel decoder with detectron2's registries (iff detectron2 is
    importable) and install the `MultiScaleDeformableAttention` module name — see quber_b200/registry.py."""
    from .registry import register as _r
    return _r(override)
