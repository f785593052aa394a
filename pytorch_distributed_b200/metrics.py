"""Prints with the reference's wording plus machine-readable JSONL and device timers.

Reference prints: step ticker (ref:resnet_single_gpu.py:23-24), validation line
(ref:restnet_ddp.py:70), epoch time (``:146``), new-best line (``:149``).
"""
from __future__ import annotations

import json
import time
from typing import Optional

import torch


def print_step(epoch: int, step: int) -> None:
    print("epoch: {}, step: {}".format(epoch, step), flush=True)


def print_validation(epoch: int, loss_val: float, acc1: float, acc5: float) -> None:
    print(f"Epoch: {epoch}, Loss: {loss_val}, Acc1: {acc1:.2f}%, Acc5: {acc5:.2f}%", flush=True)


def print_epoch_time(seconds: float) -> None:
    print("cost time per epoch: {:.4f} s".format(seconds), flush=True)


def print_new_best(acc: float) -> None:
    print(f"New Best Acc: {100 * acc:.2f}%!", flush=True)


class JsonlLogger:
    def __init__(self, path: Optional[str]):
        self._f = open(path, "a") if path else None

    @property
    def enabled(self) -> bool:
        return self._f is not None

    def log(self, **rec) -> None:
        if self._f is not None:
            rec.setdefault("ts", time.time())
            self._f.write(json.dumps(rec) + "\n")
            self._f.flush()

    def close(self) -> None:
        if self._f is not None:
            self._f.close()
            self._f = None


class DeviceTimer:
    """CUDA-event timer on the current stream; wall clock on CPU.  Multi-GPU numbers must be
    the max over ranks (``max_over_ranks``), never a host clock (BASELINE.json)."""

    def __init__(self, device: torch.device):
        self.cuda = device.type == "cuda"
        if self.cuda:
            self._a = torch.cuda.Event(enable_timing=True)
            self._b = torch.cuda.Event(enable_timing=True)
        self._t0 = 0.0
        self._ms = 0.0

    def start(self) -> None:
        if self.cuda:
            self._a.record()
        else:
            self._t0 = time.perf_counter()

    def stop(self) -> float:
        if self.cuda:
            self._b.record()
            self._b.synchronize()
            self._ms = self._a.elapsed_time(self._b)
        else:
            self._ms = (time.perf_counter() - self._t0) * 1e3
        return self._ms


def max_over_ranks(value: float, device: torch.device) -> float:
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device if device.type == "cuda" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
