"""skghoi_b200 -- B200-native (sm_100a) implementation of the SKGHOI interaction head.

The package holds only what the hot path needs: ``csrc/`` (hand-written CUDA kernels and the
C-ABI of include/skghoi_b200.h), ``_lib`` (ctypes binding), and the host-side mirror of the
reference's operator / module interface (``ops``, ``kge``, ``head``).  There is no CPU fallback:
every operator raises when the CUDA library is missing or a tensor is not on a CUDA device.
"""
__version__ = "0.1.0"
