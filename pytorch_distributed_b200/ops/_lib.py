"""Loader for ``libpdb200.so`` and small ctypes helpers."""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional

import torch

_lib: Optional[ctypes.CDLL] = None
_err: Optional[str] = None


def _load() -> Optional[ctypes.CDLL]:
    global _lib, _err
    if _lib is not None or _err is not None:
        return _lib
    from pathlib import Path
    from ..build import LIB
    # PDB200_LIB: load another build of the kernel library (same C ABI) — for A/B timing of kernel variants
    path = Path(os.environ["PDB200_LIB"]) if os.environ.get("PDB200_LIB") else LIB / "libpdb200.so"
    if not path.exists():
        _err = f"{path} not built (run: python -m pytorch_distributed_b200.build)"
        return None
    if not torch.cuda.is_available():
        _err = "no CUDA device"
        return None
    try:
        torch.cuda.init()
        _lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        from . import _sigs
        _sigs.declare(_lib)
    except OSError as e:  # pragma: no cover
        _err = f"cannot load {path}: {e}"
        _lib = None
    return _lib


def available() -> bool:
    if os.environ.get("PDB200_DISABLE_NATIVE") == "1":
        return False
    return _load() is not None


def require() -> ctypes.CDLL:
    L = _load()
    if L is None:
        raise RuntimeError(f"pytorch_distributed_b200 native kernels unavailable: {_err}")
    return L


def lib() -> ctypes.CDLL:
    return require()


def stream_ptr(device=None) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def check(code: int, what: str) -> None:
    if code != 0:
        L = require()
        L.pdb_last_error.restype = ctypes.c_char_p
        msg = L.pdb_last_error()
        raise RuntimeError(f"{what} failed (code {code}): {msg.decode() if msg else ''}")


_launches = 0
_tls = threading.local()


def count_launch(n: int = 1) -> None:
    global _launches
    _launches += n
    _tls.n = getattr(_tls, "n", 0) + n


def thread_launch_count() -> int:
    """Launches issued by the calling thread (CUDA-graph capture runs per thread in the DataParallel mode)."""
    return getattr(_tls, "n", 0)


def launch_count() -> int:
    """Number of kernels of this library launched so far by this process (bench evidence)."""
    return _launches
