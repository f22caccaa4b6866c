"""houv_b200 -- B200-native (sm_100a) HOUV registration hot path.

Host side: Python/PyTorch mirror of the reference's operator surface.
Device side: hand-written CUDA behind the C ABI of include/houv_b200.h
(houv_b200/libhouv_b200.so).  There is no CPU fallback.
"""
__version__ = "0.1"
