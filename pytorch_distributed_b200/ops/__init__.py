"""Python face of the sm_100a kernel library (``lib/libpdb200.so``, C ABI in include/pdb200.h).

Every function here launches hand-written CUDA on the current stream through ctypes; there is no
silent fallback — if the library is missing on a GPU box the first call raises (the driver checks
that the native code is what actually runs).
"""
from __future__ import annotations

from ._lib import available, launch_count, lib, require  # noqa: F401
from .elementwise import *  # noqa: F401,F403
from . import elementwise  # noqa: E402,F401
