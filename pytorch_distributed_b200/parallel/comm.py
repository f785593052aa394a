"""Process-group plumbing and the two gradient-communication backends.

``init_process_group`` keeps the reference's call (ref:restnet_ddp.py:94:
``backend='nccl', init_method='tcp://ip:port', world_size, rank``); torch's TCPStore does the
bootstrap (torch:distributed/rendezvous.py:162-239).  On top of the group sit two
interchangeable backends for the arena collectives DDP needs:

* ``TorchComm``  — ``torch.distributed`` calls on arena slices (gloo on CPU; NCCL as the
  *labelled baseline* on GPU, ``--backend nccl``).
* ``NativeComm`` — (parallel/symm.py) symmetric-memory arenas + our sm_100a peer-memory
  kernels (K1–K3, K5); the product path on a multi-GPU box.
"""
from __future__ import annotations

import datetime
import time
from typing import List, Optional

import torch
import torch.distributed as dist


def init_process_group(backend: str = "nccl", init_method: Optional[str] = None, world_size: int = -1,
                       rank: int = -1, timeout: Optional[datetime.timedelta] = None, **kw):
    if backend == "nccl" and not torch.cuda.is_available():
        backend = "gloo"          # CPU plumbing path (BASELINE.json config 1)
    if timeout is None:
        timeout = datetime.timedelta(minutes=10)   # ProcessGroupNCCL.hpp:156-157 default
    return dist.init_process_group(backend=backend, init_method=init_method, world_size=world_size,
                                   rank=rank, timeout=timeout, **kw)


def get_rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def get_world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def is_initialized() -> bool:
    return dist.is_available() and dist.is_initialized()


def destroy_process_group() -> None:
    global _active_native
    _active_native = None
    if is_initialized():
        dist.destroy_process_group()


def reduce(tensor: torch.Tensor, dst: int = 0, op=None, group=None):
    """``dist.reduce(x, 0)`` of the validation loop (ref:restnet_ddp.py:63-64)."""
    if not is_initialized() or get_world_size() == 1:
        return None
    nat = _active_native
    if nat is not None and tensor.is_cuda and tensor.dtype == torch.float32 and op in (None, dist.ReduceOp.SUM) \
            and group is None:
        nat.allreduce_small_(tensor)      # K5: one 16-byte peer-memory reduction (every rank gets the sum)
        return None
    return dist.reduce(tensor, dst, op=op if op is not None else dist.ReduceOp.SUM, group=group)


class TorchComm:
    """Arena collectives through torch.distributed (gloo / NCCL)."""

    name = "torch"
    fuses_unscale = False

    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if is_initialized() else 1
        self.rank = dist.get_rank(group) if is_initialized() else 0
        self._works: List = []
        self.profile = False                      # --profile: wall time per bucket (launch -> completion of its wait)
        self._bucket_marks: List = []

    def register_arena(self, arena) -> None:      # nothing to map
        return None

    def broadcast_(self, flat: torch.Tensor, src: int = 0) -> None:
        if self.world > 1:
            dist.broadcast(flat, src, group=self.group)

    def allreduce_range(self, flat: torch.Tensor, start: int, end: int, *, average: bool = True,
                        wire_dtype: Optional[torch.dtype] = None, inv_scale=None, found_inf=None) -> None:
        if self.world == 1:
            return
        seg = flat[start:end]
        t0 = time.perf_counter() if self.profile else None
        if wire_dtype is not None and wire_dtype != seg.dtype:
            # fp16/bf16_compress_hook semantics (torch:distributed/algorithms/ddp_comm_hooks/
            # default_hooks.py:57-134): cast, pre-divide, reduce, copy back.
            wire = seg.to(wire_dtype)
            if average:
                wire.div_(self.world)
            work = dist.all_reduce(wire, group=self.group, async_op=True)
            self._works.append((work, seg, wire, t0))
        else:
            if average:
                seg.div_(self.world)
            work = dist.all_reduce(seg, group=self.group, async_op=True)
            self._works.append((work, None, seg, t0))

    def wait_all(self) -> None:
        for work, seg, wire, t0 in self._works:
            work.wait()
            if seg is not None:
                seg.copy_(wire)
            if t0 is not None:
                self._bucket_marks.append((wire.numel() * 4, time.perf_counter() - t0))
        self._works.clear()

    def bucket_times(self, link_gbs: float = 770.0):
        """Same record layout as NativeComm.bucket_times, but HOST-timed (launch to completed wait), which is all a
        torch.distributed work handle offers; bytes are counted as fp32 gradients."""
        out = []
        for nbytes, sec in self._bucket_marks:
            us = 1e6 * sec
            bus = 2.0 * (self.world - 1) / self.world * nbytes / max(us, 1e-3) / 1e3
            out.append({"bytes": nbytes, "us": round(us, 1), "bus_gbs": round(bus, 3),
                        "roofline_frac": round(bus / link_gbs, 4), "timed": "host"})
        self._bucket_marks.clear()
        return out


_active_native = None        # the NativeComm of the default group, once one exists (used by reduce())


def make_comm(backend: str = "auto", group=None):
    """auto: the native peer-memory backend on multi-GPU CUDA, torch.distributed on CPU / single rank.

    On a multi-GPU CUDA job a failure to bring up the native backend is an ERROR, not a quiet switch to NCCL:
    ask for the library path explicitly (``backend='nccl'``) or set ``PDB200_ALLOW_NCCL_FALLBACK=1``."""
    import os
    if backend in ("nccl", "gloo", "torch"):
        return TorchComm(group)
    want_native = backend == "native"
    if torch.cuda.is_available() and is_initialized() and dist.get_backend(group) == "nccl" \
            and dist.get_world_size(group) > torch.cuda.device_count() and not want_native:
        # More ranks than local GPUs: the job spans hosts (the reference's "Apex * 4" run).  Peer memory and NVLS end at
        # the NVSwitch domain, so the gradient path of a multi-node job is NCCL; this is a decision, not a fallback.
        import warnings
        warnings.warn("pdb200: multi-node job — gradient all-reduce goes through NCCL (the native peer-memory backend "
                      "covers one NVSwitch domain)")
        return TorchComm(group)
    if torch.cuda.is_available() and is_initialized() and dist.get_backend(group) == "nccl" \
            and dist.get_world_size(group) > 1:
        try:
            from .symm import NativeComm
            global _active_native
            if _active_native is not None and group is None:
                return _active_native
            c = NativeComm(group)
            if group is None:
                _active_native = c
            return c
        except Exception as e:
            if want_native or os.environ.get("PDB200_ALLOW_NCCL_FALLBACK", "0") != "1":
                raise RuntimeError(
                    f"the native peer-memory communication backend could not be initialised ({type(e).__name__}: {e}); "
                    "pass backend='nccl' (--backend nccl) or set PDB200_ALLOW_NCCL_FALLBACK=1 to run on "
                    "torch.distributed instead") from e
            import warnings
            warnings.warn(f"pdb200: native comm backend unavailable ({type(e).__name__}: {e}); using torch.distributed")
    elif want_native and is_initialized() and dist.get_world_size(group) > 1:
        raise RuntimeError("backend='native' needs CUDA + an NCCL process group for the bootstrap")
    return TorchComm(group)
