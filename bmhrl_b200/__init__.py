"""bmhrl_b200 -- B200-native (sm_100a) implementation of the BMHRL data-parallel hot path.

Host side: Python/PyTorch modules that keep the reference's module API (Berghojo/bmhrl, model/*.py and
loss/*.py) and state_dict layout.  Compute: hand-written CUDA behind the C ABI declared in
include/bmhrl_b200.h (libbmhrl_b200.so, loaded with ctypes; no torch types cross the boundary).
There is no CPU fallback: calling any op without the built library or without a CUDA device raises.
"""
__version__ = "0.1.0"
