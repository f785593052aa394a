"""``latest.pt`` / ``best.pt`` — formats loadable by and from stock torch.

* ``latest.pt``: ``{'model', 'optimizer', 'scheduler', 'acc', 'epoch', 'step'}`` where
  ``model`` is the *bare* (unwrapped) state_dict and ``step`` is the next step to run
  (ref:restnet_ddp.py:37-45).  We add an optional ``'scaler'`` key for the amp mode
  (quirk Q6: the reference forgets it); the reference's loader indexes by name, so the
  extra key is harmless to it.
* ``best.pt``: bare model state_dict (ref:restnet_ddp.py:150).
"""
from __future__ import annotations

import os
from pathlib import Path
from typing import Any, Dict, Optional, Tuple

import torch


def unwrap(model):
    return model.module if hasattr(model, "module") and isinstance(getattr(model, "module"), torch.nn.Module) else model


def _atomic_save(obj: Any, path: Path) -> None:
    path.parent.mkdir(parents=True, exist_ok=True)
    tmp = path.with_suffix(path.suffix + f".tmp{os.getpid()}")
    torch.save(obj, tmp)
    os.replace(tmp, path)        # a suspend that is interrupted never leaves half a file


def _cpu_contiguous(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    # Arena-backed parameters are strided views; serialise plain dense copies so the file
    # is independent of our memory layout and of the device.
    return {k: (v.detach().to("cpu").contiguous() if torch.is_tensor(v) else v) for k, v in sd.items()}


def save_latest(save_dir: Path, model, optimizer, scheduler, acc: float, epoch: int, step: int,
                scaler=None) -> Path:
    state = {
        "model": _cpu_contiguous(unwrap(model).state_dict()),
        "optimizer": optimizer.state_dict(),
        "scheduler": scheduler.state_dict(),
        "acc": acc,
        "epoch": epoch,
        "step": step,
    }
    if scaler is not None:
        state["scaler"] = scaler.state_dict()
    path = Path(save_dir) / "latest.pt"
    _atomic_save(state, path)
    return path


def save_best(save_dir: Path, model) -> Path:
    path = Path(save_dir) / "best.pt"
    _atomic_save(_cpu_contiguous(unwrap(model).state_dict()), path)
    return path


def load_latest(save_dir: Path, model, optimizer, scheduler, scaler=None) -> Optional[Tuple[float, int, int]]:
    """Returns ``(best_acc, start_epoch, start_step)`` or None (ref:restnet_ddp.py:126-132)."""
    path = Path(save_dir) / "latest.pt"
    if not path.exists():
        return None
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    unwrap(model).load_state_dict(ckpt["model"])
    optimizer.load_state_dict(ckpt["optimizer"])
    scheduler.load_state_dict(ckpt["scheduler"])
    if scaler is not None and "scaler" in ckpt:
        scaler.load_state_dict(ckpt["scaler"])
    return ckpt["acc"], ckpt["epoch"], ckpt["step"]
