"""ggpm_b200 — B200-native (sm_100a) implementation of ggpm's hierarchical message-passing encoder.

Drop-in classes (same names / constructor / forward contract as ggpm/encoder.py and ggpm/rnn.py):
    from ggpm_b200 import HierMPNEncoder, MotifEncoder, MPNEncoder, GRU, LSTM
    from ggpm_b200 import IncHierMPNEncoder, IncEncoder, IncMPNEncoder      # the decoder's incremental encoders
    from ggpm_b200 import PropertyOptimizer, PropertyRegressor              # ggpm/property_optimizer.py heads
The CUDA library (libggpm_b200.so, C ABI in include/ggpm_b200.h) is loaded on first use; without it, or
without a CUDA device, the modules raise — there is no CPU fallback.
"""
from .encoder import (HierMPNEncoder, IncEncoder, IncHierMPNEncoder, IncMPNEncoder, MotifEncoder,  # noqa: F401
                      MPNEncoder)
from .rnn import GRU, LSTM                                         # noqa: F401
from .property_optimizer import PropertyOptimizer, PropertyRegressor   # noqa: F401

__all__ = ["HierMPNEncoder", "MotifEncoder", "MPNEncoder", "IncHierMPNEncoder", "IncEncoder", "IncMPNEncoder", "GRU", "LSTM",
           "PropertyOptimizer", "PropertyRegressor"]
