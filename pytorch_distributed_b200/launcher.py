"""Process launcher: one process per GPU with error propagation and NUMA binding.

Call shape of ``hfai.multiprocessing.spawn(main, args=(), nprocs=ngpus, bind_numa=True)``
(ref:restnet_ddp.py:153-155).  Semantics follow torch:multiprocessing/spawn.py:99-340:
``spawn`` start method, ``fn(i, *args)`` in child i, the first failure terminates the siblings
(SIGTERM, grace period, SIGKILL) and is re-raised in the parent with the child's traceback.
Differences by design: a child that exits with ``suspend.SUSPEND_EXIT_CODE`` (cooperative
pre-emption, quirk Q10) tears the group down *without* being reported as a failure, and
``bind_numa`` pins each child to its GPU's NUMA node through the C++ runtime
(csrc/runtime/numa.cpp).
"""
from __future__ import annotations

import multiprocessing as mp
import os
import pickle
import signal
import sys
import tempfile
import time
import traceback
from typing import Callable, Optional, Sequence

from .suspend import SUSPEND_EXIT_CODE


class ProcessRaisedException(Exception):
    def __init__(self, msg: str, error_index: int, pid: int):
        super().__init__(msg)
        self.error_index, self.pid = error_index, pid


class ProcessExitedException(Exception):
    def __init__(self, msg: str, error_index: int, pid: int, exit_code: int, signal_name: Optional[str] = None):
        super().__init__(msg)
        self.error_index, self.pid, self.exit_code, self.signal_name = error_index, pid, exit_code, signal_name


def _bind_numa(local_rank: int) -> None:
    try:
        import torch
        from . import runtime
        if torch.cuda.is_available() and local_rank < torch.cuda.device_count():
            # PCI bus id without creating a CUDA context in this process yet (NVML).
            pci = None
            try:
                import pynvml
                pynvml.nvmlInit()
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                phys = int(vis.split(",")[local_rank]) if vis and vis.split(",")[local_rank].isdigit() else local_rank
                pci = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(phys)).busId
                if isinstance(pci, bytes):
                    pci = pci.decode()
            except Exception:
                pci = None
            if pci:
                runtime.bind_numa_for_pci(pci)
    except Exception as e:
        # affinity is an optimisation and must not fail a launch, but it must not fail silently either
        import warnings
        warnings.warn(f"pdb200: NUMA binding of local rank {local_rank} skipped: {type(e).__name__}: {e}")


def _child(fn: Callable, i: int, args: Sequence, error_file: str, bind_numa: bool) -> None:
    try:
        import ctypes
        try:  # die with the parent (torch:multiprocessing/spawn.py:79-96)
            ctypes.CDLL("libc.so.6").prctl(1, signal.SIGINT)
        except Exception:
            pass
        if bind_numa:
            _bind_numa(i)
        fn(i, *args)
    except KeyboardInterrupt:
        pass
    except SystemExit:
        raise
    except BaseException:
        with open(error_file, "wb") as f:
            pickle.dump(traceback.format_exc(), f)
        sys.exit(1)


class ProcessContext:
    def __init__(self, processes, error_files, grace_period: float = 30.0):
        self.processes, self.error_files, self.grace_period = processes, error_files, grace_period
        self.suspended = False

    def pids(self):
        return [p.pid for p in self.processes]

    def _terminate_others(self, skip: int) -> None:
        for j, p in enumerate(self.processes):
            if j != skip and p.is_alive():
                p.terminate()
        deadline = time.monotonic() + self.grace_period
        for j, p in enumerate(self.processes):
            if j == skip:
                continue
            p.join(max(0.0, deadline - time.monotonic()))
            if p.is_alive():
                p.kill()
                p.join()

    def join(self, timeout: Optional[float] = None) -> bool:
        """True when every child has exited cleanly; raises on the first failure."""
        sentinels = {p.sentinel: i for i, p in enumerate(self.processes) if p.exitcode is None}
        if not sentinels:
            return True
        mp.connection.wait(list(sentinels), timeout=timeout)
        bad = None
        for i, p in enumerate(self.processes):
            if p.exitcode is not None and p.exitcode != 0:
                bad = i
                break
        if bad is None:
            return all(p.exitcode is not None for p in self.processes)
        p = self.processes[bad]
        self._terminate_others(bad)
        if p.exitcode == SUSPEND_EXIT_CODE:
            self.suspended = True
            return True
        ef = self.error_files[bad]
        if os.path.exists(ef) and os.path.getsize(ef) > 0:
            with open(ef, "rb") as f:
                tb = pickle.load(f)
            raise ProcessRaisedException(f"\n\n-- Process {bad} terminated with the following error:\n{tb}",
                                         bad, p.pid)
        if p.exitcode < 0:
            try:
                name = signal.Signals(-p.exitcode).name
            except ValueError:
                name = f"<unknown signal {-p.exitcode}>"
            raise ProcessExitedException(f"process {bad} terminated with signal {name}", bad, p.pid,
                                         p.exitcode, name)
        raise ProcessExitedException(f"process {bad} terminated with exit code {p.exitcode}", bad, p.pid,
                                     p.exitcode)


def start_processes(fn: Callable, args: Sequence = (), nprocs: int = 1, join: bool = True,
                    daemon: bool = False, start_method: str = "spawn", bind_numa: bool = False,
                    grace_period: float = 30.0):
    ctx = mp.get_context(start_method)
    tmpdir = tempfile.mkdtemp(prefix="pdb200_spawn_")
    procs, files = [], []
    for i in range(nprocs):
        ef = os.path.join(tmpdir, f"rank{i}.err")
        p = ctx.Process(target=_child, args=(fn, i, tuple(args), ef, bind_numa), daemon=daemon)
        p.start()
        procs.append(p)
        files.append(ef)
    pc = ProcessContext(procs, files, grace_period)
    if not join:
        return pc
    try:
        while not pc.join():
            pass
    finally:
        for f in files:
            try:
                os.remove(f)
            except OSError:
                pass
        try:
            os.rmdir(tmpdir)
        except OSError:
            pass
    return pc


def spawn(fn: Callable, args: Sequence = (), nprocs: int = 1, join: bool = True, daemon: bool = False,
          bind_numa: bool = False, grace_period: float = 30.0):
    return start_processes(fn, args, nprocs, join, daemon, "spawn", bind_numa, grace_period)
