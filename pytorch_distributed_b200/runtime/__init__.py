"""ctypes bindings of the host runtime (``include/pdb200_runtime.h``)."""
from __future__ import annotations

import ctypes
from ctypes import POINTER, c_char_p, c_int, c_int32, c_int64, c_void_p
from typing import List, Optional, Sequence, Tuple

_lib: Optional[ctypes.CDLL] = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        from ..build import LIB, build_runtime
        path = LIB / "libpdb200_rt.so"
        if not path.exists():
            build_runtime()
        L = ctypes.CDLL(str(path))
        L.pdb_plan_buckets.restype = c_int
        L.pdb_plan_buckets.argtypes = [c_int, POINTER(c_int64), POINTER(c_int64), c_int, c_int64, c_int64,
                                       c_int64, c_int, POINTER(c_int64), POINTER(c_int64), POINTER(c_int32)]
        L.pdb_reducer_create.restype = c_void_p
        L.pdb_reducer_create.argtypes = [c_int, POINTER(c_int32), c_int]
        L.pdb_reducer_destroy.argtypes = [c_void_p]
        L.pdb_reducer_destroy.restype = None
        L.pdb_reducer_reset.argtypes = [c_void_p]
        L.pdb_reducer_reset.restype = None
        L.pdb_reducer_mark_ready.restype = c_int
        L.pdb_reducer_mark_ready.argtypes = [c_void_p, c_int, POINTER(c_int32), c_int]
        L.pdb_reducer_flush.restype = c_int
        L.pdb_reducer_flush.argtypes = [c_void_p, POINTER(c_int32), c_int]
        L.pdb_reducer_done.restype = c_int
        L.pdb_reducer_done.argtypes = [c_void_p]
        L.pdb_reducer_pending.restype = c_int
        L.pdb_reducer_pending.argtypes = [c_void_p, c_int]
        L.pdb_bind_numa_for_pci.restype = c_int
        L.pdb_bind_numa_for_pci.argtypes = [c_char_p]
        L.pdb_bind_numa_node.restype = c_int
        L.pdb_bind_numa_node.argtypes = [c_int]
        L.pdb_affinity_count.restype = c_int
        L.pdb_runtime_version.restype = c_int
        _lib = L
    return _lib


def plan_buckets(offsets: Sequence[int], numels: Sequence[int], elem_bytes: int = 4,
                 first_cap_bytes: int = 1 << 20, cap_bytes: int = 25 << 20, align: int = 64
                 ) -> Tuple[List[Tuple[int, int]], List[int]]:
    """Returns ``(ranges in launch order, bucket index per parameter)``."""
    n = len(offsets)
    off = (c_int64 * n)(*offsets)
    num = (c_int64 * n)(*numels)
    bs, be, pb = (c_int64 * n)(), (c_int64 * n)(), (c_int32 * n)()
    nb = lib().pdb_plan_buckets(n, off, num, elem_bytes, first_cap_bytes, cap_bytes, align, n, bs, be, pb)
    if nb < 0:
        raise ValueError("pdb_plan_buckets: invalid arguments (offsets must be ascending and disjoint)")
    return [(bs[i], be[i]) for i in range(nb)], list(pb)


class Reducer:
    """Releases gradient buckets in a rank-invariant order as parameters become ready."""

    def __init__(self, param_bucket: Sequence[int], n_buckets: int):
        n = len(param_bucket)
        self._n_buckets = n_buckets
        self._h = lib().pdb_reducer_create(n, (c_int32 * n)(*param_bucket), n_buckets)
        if not self._h:
            raise ValueError("pdb_reducer_create failed")
        self._out = (c_int32 * max(1, n_buckets))()

    def reset(self) -> None:
        lib().pdb_reducer_reset(self._h)

    def mark_ready(self, param: int) -> List[int]:
        k = lib().pdb_reducer_mark_ready(self._h, param, self._out, self._n_buckets)
        if k < 0:
            raise RuntimeError(f"parameter {param} marked ready twice in one backward pass "
                               "(or index out of range)")
        return [self._out[i] for i in range(k)]

    def flush(self) -> List[int]:
        k = lib().pdb_reducer_flush(self._h, self._out, self._n_buckets)
        return [self._out[i] for i in range(k)]

    def done(self) -> bool:
        return bool(lib().pdb_reducer_done(self._h))

    def pending(self, bucket: int) -> int:
        return lib().pdb_reducer_pending(self._h, bucket)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and _lib is not None:
            _lib.pdb_reducer_destroy(h)


def bind_numa_for_pci(pci_bus_id: str) -> int:
    return lib().pdb_bind_numa_for_pci(pci_bus_id.encode())


def bind_numa_node(node: int) -> int:
    return lib().pdb_bind_numa_node(node)


def affinity_count() -> int:
    return lib().pdb_affinity_count()
