"""NativeComm — arena collectives as our own sm_100a kernels over NVLink peer memory.

torch's symmetric-memory allocator is used for exactly two things: allocating buffers that can be mapped by
peers, and exchanging the handles (``rendezvous``; bootstrap over the NCCL group's store).  Everything that
moves or reduces data is ``csrc/comm/collectives.cu``: in-place two-shot all-reduce with fused
1/world · 1/loss-scale · non-finite check (K1), pull broadcast (K2/K3), one-shot small all-reduce for SyncBN
statistics and metric vectors (K5/K6).  Kernels run on a high-priority side stream so that a bucket's
all-reduce overlaps the rest of backward.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .. import ops
from ..ops._lib import check, count_launch, ptr, require


class _Mapped:
    """A symmetric tensor after rendezvous: per-rank base pointers, signal pads, multicast address."""

    def __init__(self, tensor: torch.Tensor, group):
        import torch.distributed._symmetric_memory as symm_mem
        self.tensor = tensor
        self.handle = symm_mem.rendezvous(tensor, group if group is not None else dist.group.WORLD)
        h = self.handle
        self.world, self.rank = h.world_size, h.rank
        self.bufs = (ctypes.c_void_p * self.world)(*[int(p) for p in h.buffer_ptrs])
        self.pads = (ctypes.c_void_p * self.world)(*[int(p) for p in h.signal_pad_ptrs])
        mc = 0
        try:
            mc = int(h.multicast_ptr) if h.has_multicast_support else 0
        except Exception:
            mc = 0
        self.multicast = mc
        self.pad_bytes = int(getattr(h, "signal_pad_size", 2048))


class NativeComm:
    name = "native"
    fuses_unscale = True

    def __init__(self, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        require()
        L = require()
        if not hasattr(L, "pdb_allreduce_inplace"):
            raise RuntimeError("libpdb200.so was built without csrc/comm")
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.stream = torch.cuda.Stream(self.device, priority=-1)
        self._mapped: Dict[int, _Mapped] = {}
        # symmetric scratch: [0:64) non-finite flags, then the staging area float[2][world][slot] of the small sums
        # (SyncBN statistics inside the conv / BN kernels, metric vectors): two halves used alternately, see
        # csrc/comm/peer_sync.cuh
        self.slot = 8192
        self._aux_t = symm_mem.empty(64 + 2 * self.world * self.slot, dtype=torch.float32, device=self.device)
        self._aux_t.zero_()
        self._aux = _Mapped(self._aux_t, group)
        self.flag = self._aux_t[:1]
        fl = [int(p) for p in self._aux.handle.buffer_ptrs]
        self._flag_ptrs = (ctypes.c_void_p * self.world)(*fl)
        self._staging_ptrs = (ctypes.c_void_p * self.world)(*[p + 64 * 4 for p in fl])
        from ..ops import SyncPeersC
        self._sync = SyncPeersC()
        for r in range(self.world):
            self._sync.staging[r] = fl[r] + 64 * 4
            self._sync.pads[r] = int(self._aux.handle.signal_pad_ptrs[r])
        self._sync.rank, self._sync.world, self._sync.slot_stride, self._sync.row = self.rank, self.world, self.slot, 0
        self.oneshot_bytes = int(os.environ.get("PDB200_ONESHOT_BYTES", str(256 << 10)))
        self.use_multicast = os.environ.get("PDB200_NVLS", "1") != "0"
        # PDB200_COMM_SMALL_CTA=1: 256-thread / 80-register CTAs that fit beside a persistent convolution CTA, twice as many
        self.blocks = int(os.environ.get("PDB200_COMM_BLOCKS", "64" if os.environ.get("PDB200_COMM_SMALL_CTA") == "1" else "32"))
        self._wire: Dict[torch.dtype, tuple] = {}       # wire dtype -> (mapped staging, half_elems)
        self._events: List[torch.cuda.Event] = []
        self._pass_started = False
        self._found_inf: Optional[torch.Tensor] = None
        # --profile: an event pair around every bucket's kernel (per-bucket bus GB/s in the JSONL log)
        self.profile = False
        self._bucket_marks: List[tuple] = []
        self._trace: Optional[torch.Tensor] = None
        self._trace_read = 0
        torch.cuda.synchronize(self.device)
        dist.barrier(group)

    # -- registration -----------------------------------------------------------------------------------
    def _map(self, t: torch.Tensor) -> _Mapped:
        key = t.untyped_storage().data_ptr()
        m = self._mapped.get(key)
        if m is None:
            m = _Mapped(t, self.group)
            self._mapped[key] = m
        return m

    def register_arena(self, arena) -> None:
        if not (getattr(arena, "grads_symmetric", False) and getattr(arena, "params_symmetric", False)):
            raise RuntimeError("NativeComm needs arenas in symmetric memory: create the process group before "
                               "the model's arenas (ParamArena allocates symmetric memory when it is initialised)")
        self._map(arena.grads)
        self._map(arena.params)

    def _wire_staging(self, dtype: torch.dtype, n: int):
        """Symmetric 16-bit staging (send half | receive half) for the compressed-wire all-reduce; sized once
        for the largest bucket.  Collective (rendezvous): every rank reaches it for the same bucket."""
        import torch.distributed._symmetric_memory as symm_mem
        cur = self._wire.get(dtype)
        if cur is None or cur[1] < n:
            half = (max(n, 8 << 20) + 63) // 64 * 64
            t = symm_mem.empty(2 * half, dtype=dtype, device=self.device)
            m = _Mapped(t, self.group)
            cur = (m, half)
            self._wire[dtype] = cur
        return cur

    def sync_peers(self, n_floats: int):
        """The handle a kernel needs to sum ``n_floats`` values over the ranks by itself (SyncBN statistics inside
        the convolution epilogue / the BN backward reduction), or None when the payload does not fit a staging slot."""
        return self._sync if self.world > 1 and n_floats <= self.slot else None

    def _blocks(self, m: _Mapped, n: int) -> int:
        cap = require().pdb_comm_max_blocks(self.world, m.pad_bytes)
        want = max(1, min(self.blocks, cap, (n // 4 + 4095) // 4096))
        return want

    # -- collectives ------------------------------------------------------------------------------------
    def broadcast_(self, flat: torch.Tensor, src: int = 0) -> None:
        if self.world == 1:
            return
        sym = flat.untyped_storage().data_ptr() in self._mapped or getattr(flat, "_pdb200_symmetric", False)
        if not sym:
            try:
                m = self._map(flat)
            except Exception:
                dist.broadcast(flat, src, group=self.group)
                return
        m = self._map(flat)
        start = (flat.data_ptr() - m.tensor.data_ptr()) // 4
        n = flat.numel() * flat.element_size() // 4
        if n % 4 or start % 4:
            dist.broadcast(flat, src, group=self.group)
            return
        check(require().pdb_broadcast_inplace(m.bufs, m.pads, self.rank, self.world, src, start, n,
                                              self._blocks(m, n), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "pdb_broadcast_inplace")
        count_launch()

    def broadcast_async(self, flat: torch.Tensor, src: int = 0):
        """The same broadcast on the comm stream, ordered after everything enqueued so far on the current stream; returns
        the event that marks its completion (or None when nothing was launched).  DDP uses it to refresh the BN buffers
        for the NEXT forward while this step's backward runs, instead of opening every forward with a cross-rank meeting."""
        if self.world == 1:
            return None
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.stream.wait_event(ev)
        with torch.cuda.stream(self.stream):
            self.broadcast_(flat, src)
            done = torch.cuda.Event()
            done.record(self.stream)
        return done

    def allreduce_range(self, flat: torch.Tensor, start: int, end: int, *, average: bool = True,
                        wire_dtype=None, inv_scale=None, found_inf=None) -> None:
        if self.world == 1:
            return
        m = self._map(flat)
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.stream.wait_event(ev)
        with torch.cuda.stream(self.stream):
            if not self._pass_started:
                self._pass_started = True
                self.flag.zero_()
            self._found_inf = found_inf
            n = end - start
            t0 = None
            if self.profile:
                t0 = torch.cuda.Event(enable_timing=True)
                t0.record(self.stream)
            if wire_dtype in (torch.float16, torch.bfloat16) and n % 8 == 0:
                wm, half = self._wire_staging(wire_dtype, n)
                check(require().pdb_allreduce_inplace_wire16(
                    m.bufs, m.pads, self._flag_ptrs if found_inf is not None else None, wm.bufs, half,
                    0 if wire_dtype == torch.bfloat16 else 1, self.rank, self.world, start, n,
                    (1.0 / self.world) if average else 1.0, ptr(inv_scale), self._blocks(m, n),
                    ctypes.c_void_p(self.stream.cuda_stream)), "pdb_allreduce_inplace_wire16")
                count_launch()
                done = torch.cuda.Event(enable_timing=t0 is not None)
                done.record(self.stream)
                self._events.append(done)
                if t0 is not None:
                    self._bucket_marks.append((4 * n, t0, done))
                return
            if 4 * n <= self.oneshot_bytes and n % 4 == 0:
                # small bucket: latency-bound, one NVLink round instead of reduce-scatter + all-gather
                check(require().pdb_allreduce_oneshot(
                    m.bufs, m.pads, self._flag_ptrs if found_inf is not None else None, self.rank, self.world, start, n,
                    (1.0 / self.world) if average else 1.0, ptr(inv_scale),
                    ctypes.c_void_p(self.stream.cuda_stream)), "pdb_allreduce_oneshot")
                count_launch()
                done = torch.cuda.Event(enable_timing=t0 is not None)
                done.record(self.stream)
                if t0 is not None:
                    self._bucket_marks.append((4 * n, t0, done))
                self._events.append(done)
                return
            if t0 is not None and self._trace is None and hasattr(require(), "pdb_comm_set_trace"):
                # in-kernel timestamps (entry / after the first meeting / exit) for every two-shot launch from now on
                self._trace = torch.zeros(3 * 4096, dtype=torch.int64, device=self.device)
                require().pdb_comm_set_trace(ptr(self._trace), 4096)
                self._trace_read = 0
            if t0 is None and self._trace is not None:
                # profiling was switched off: stop stamping, so that marks and stamps stay paired when it comes back
                require().pdb_comm_set_trace(None, 0)
                self._trace = None
            traced = t0 is not None and self._trace is not None
            check(require().pdb_allreduce_inplace(
                m.bufs, m.pads, self._flag_ptrs if found_inf is not None else None,
                ctypes.c_void_p(m.multicast if self.use_multicast else 0), self.rank, self.world, start, n,
                (1.0 / self.world) if average else 1.0, ptr(inv_scale), self._blocks(m, n),
                ctypes.c_void_p(self.stream.cuda_stream)), "pdb_allreduce_inplace")
            count_launch()
            done = torch.cuda.Event(enable_timing=t0 is not None)
            done.record(self.stream)
            if t0 is not None:
                self._bucket_marks.append((4 * n, t0, done, traced))
        self._events.append(done)

    def bucket_times(self, link_gbs: float = 770.0):
        """Device time of every bucket all-reduce since the last call (needs ``profile = True`` and a completed
        step): bytes, microseconds, bus bandwidth 2(w-1)/w * bytes / t, and its fraction of the NVLink roofline.
        In-step numbers include waiting for the slowest rank at the kernel's first cross-rank barrier."""
        out = []
        stamps = None
        for mark in self._bucket_marks:
            nbytes, t0, t1 = mark[:3]
            t1.synchronize()
            us = 1e3 * t0.elapsed_time(t1)
            bus = 2.0 * (self.world - 1) / self.world * nbytes / max(us, 1e-3) / 1e3        # GB/s
            rec = {"bytes": nbytes, "us": round(us, 1), "bus_gbs": round(bus, 1), "roofline_frac": round(bus / link_gbs, 3)}
            if len(mark) > 3 and mark[3] and self._trace is not None and self._trace_read < 4096:
                # the kernel's own clock: how long it waited for the slowest rank, and how long the collective itself took
                if stamps is None:
                    stamps = self._trace.view(-1, 3).cpu()
                a, b, c = (int(v) for v in stamps[self._trace_read])
                self._trace_read += 1
                if a and b >= a and c >= b:
                    xfer = (c - b) / 1e3
                    xbus = 2.0 * (self.world - 1) / self.world * nbytes / max(xfer, 1e-3) / 1e3
                    rec.update(wait_us=round((b - a) / 1e3, 1), xfer_us=round(xfer, 1), xfer_bus_gbs=round(xbus, 1),
                               xfer_roofline_frac=round(xbus / link_gbs, 3))
            out.append(rec)
        self._bucket_marks.clear()
        return out

    def wait_all(self) -> None:
        cur = torch.cuda.current_stream(self.device)
        if self._events:
            cur.wait_event(self._events[-1])          # stream order on the comm stream covers the earlier ones
            self._events.clear()
        if self._found_inf is not None:
            torch.maximum(self._found_inf, self.flag[0], out=self._found_inf)
            self._found_inf = None
        self._pass_started = False

    def allreduce_small_(self, data: torch.Tensor) -> None:
        """In-place sum over ranks of a small fp32 vector (SyncBN statistics, metric accumulators)."""
        if self.world == 1:
            return
        n = data.numel()
        if n > self.slot or data.dtype != torch.float32 or not data.is_contiguous():
            dist.all_reduce(data, group=self.group)
            return
        check(require().pdb_small_allreduce(self._staging_ptrs, self._aux.pads, self.rank, self.world, ptr(data), n,
                                            self.slot, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "pdb_small_allreduce")
        count_launch()

    def bn_stats_sync(self, x2d: torch.Tensor, sums: torch.Tensor) -> None:
        """Per-channel sum / sum-of-squares of x2d[rows, C] over ALL ranks, exchanged inside the statistics
        kernel (K6)."""
        rows, C = x2d.shape
        scratch = ops.elementwise.bn_stats_scratch(C, x2d.device)
        check(require().pdb_bn_stats_sync(ptr(x2d), ops.dtype_tag(x2d.dtype), rows, C, ptr(sums), ptr(scratch),
                                          scratch.numel(), self._staging_ptrs, self._aux.pads, self.rank, self.world,
                                          self.slot, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "pdb_bn_stats_sync")
        count_launch()
