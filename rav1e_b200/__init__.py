"""rav1e_b200 -- B200 (sm_100a) implementation of rav1e's data-parallel inner
loop (SAD/SATD, 8-tap sub-pel MC, forward transforms, motion search) behind a
C ABI (include/rav1e_cuda.h).  `rav1e_b200.api` mirrors the reference interface
for this path; `rav1e_b200.capi` is the raw ctypes binding.  GPU only: importing
works anywhere, creating a Context needs the built library and a CUDA device.
"""
from . import capi  # noqa: F401
from .api import *  # noqa: F401,F403
