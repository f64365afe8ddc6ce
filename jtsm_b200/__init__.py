"""jtsm_b200: B200 (sm_100a) implementation of JTSM's per-proposal hot path.

Python host code mirrors the reference's nn.Module / op surface; all arithmetic runs in
hand-written CUDA kernels behind the C ABI of include/jtsm_b200.h (libjtsm_b200.so).
"""
__version__ = "0.1.0"
