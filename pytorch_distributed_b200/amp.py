"""Mixed precision: an explicit dtype plan and a GradScaler without host round trips.

Reference usage (ref:resnet_ddp_apex.py:27-34,107): forward + loss under
``torch.cuda.amp.autocast()``, ``scaler.scale(loss).backward()``, ``scaler.step(optimizer)``,
``scaler.update()``, ``GradScaler(enabled=True)``.

* ``autocast``: the delegate interposes on the dispatcher and casts per op
  (torch:include/ATen/autocast_mode.h:820-932).  Our native ResNet graph has a *static* plan —
  16-bit activations and weight shadows, fp32 master weights, fp32 BN statistics and loss —
  so the context manager only records the requested compute dtype for the model to pick up;
  on the stock-op path (CPU, or ``--engine torch``) it defers to ``torch.autocast``.
* ``GradScaler``: same constants and state machine as torch:amp/grad_scaler.py:123-159,484-559
  (init 65536, ×2 every 2000 clean steps, ×0.5 on overflow).  The delegate synchronises the
  host every step to decide whether to skip (``found_inf.item()``, ``:363-373``); here
  ``found_inf`` stays on the device, the fused SGD kernel reads it, and ``update()`` is a
  device-side select.  The un-scale itself is folded into the gradient all-reduce epilogue
  (K1) or the SGD prologue (K9) instead of being a separate pass (K13).
"""
from __future__ import annotations

import contextlib
import threading
from typing import Optional

import torch

_state = threading.local()

_DTYPES = {"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16,
           torch.float32: torch.float32, torch.bfloat16: torch.bfloat16, torch.float16: torch.float16}


def compute_dtype() -> Optional[torch.dtype]:
    """dtype requested by the innermost active ``autocast`` (None outside)."""
    return getattr(_state, "dtype", None)


@contextlib.contextmanager
def autocast(dtype="fp16", enabled: bool = True, device_type: Optional[str] = None, stock: bool = False):
    """``with autocast():`` — default fp16 like ``torch.cuda.amp.autocast`` on CUDA."""
    if not enabled:
        yield
        return
    dt = _DTYPES[dtype]
    prev = getattr(_state, "dtype", None)
    _state.dtype = dt
    try:
        if stock:
            dev = device_type or ("cuda" if torch.cuda.is_available() else "cpu")
            if dev == "cpu" and dt == torch.float16:
                dt = torch.bfloat16            # CPU autocast kernels are bf16-first
            with torch.autocast(device_type=dev, dtype=dt):
                yield
        else:
            yield
    finally:
        _state.dtype = prev


def _native_ops():
    """The kernel library when it can be used in this process, else None."""
    from . import ops
    return ops if ops.available() else None


class GradScaler:
    def __init__(self, init_scale: float = 65536.0, growth_factor: float = 2.0,
                 backoff_factor: float = 0.5, growth_interval: int = 2000, enabled: bool = True,
                 device: Optional[torch.device] = None):
        self._enabled = enabled
        self._init_scale = float(init_scale)
        self._growth_factor = float(growth_factor)
        self._backoff_factor = float(backoff_factor)
        self._growth_interval = int(growth_interval)
        self._device = torch.device(device) if device is not None else None
        self._scale: Optional[torch.Tensor] = None
        self._growth_tracker: Optional[torch.Tensor] = None
        self._found_inf: Optional[torch.Tensor] = None
        self._inv_scale: Optional[torch.Tensor] = None
        self._unscaled = False
        self._grads_carry_inv_scale = False

    # -- lazy device state (torch creates it on first ``scale()`` too) -----------------
    def _lazy_init(self, device: torch.device) -> None:
        if self._scale is None:
            self._device = self._device or device
            self._scale = torch.full((), self._init_scale, dtype=torch.float32, device=self._device)
            # a state_dict loaded before the first scale() (every resume) parked its counter here
            self._growth_tracker = torch.full((), int(getattr(self, "_pending_tracker", 0)), dtype=torch.int32,
                                              device=self._device)
            self._pending_tracker = 0
            self._found_inf = torch.zeros((), dtype=torch.float32, device=self._device)
            self._inv_scale = torch.full((), 1.0 / self._init_scale, dtype=torch.float32,
                                         device=self._device)

    def is_enabled(self) -> bool:
        return self._enabled

    def get_scale(self) -> float:
        if not self._enabled:
            return 1.0
        return self._init_scale if self._scale is None else float(self._scale.item())

    def scale_tensor(self, device) -> torch.Tensor:
        self._lazy_init(torch.device(device))
        return self._scale

    def found_inf_tensor(self, device) -> torch.Tensor:
        self._lazy_init(torch.device(device))
        return self._found_inf

    def inv_scale_tensor(self, device) -> torch.Tensor:
        self._lazy_init(torch.device(device))
        return self._inv_scale

    def scale(self, loss: torch.Tensor) -> torch.Tensor:
        if not self._enabled:
            return loss
        self._lazy_init(loss.device)
        return loss * self._scale.to(loss.dtype)

    def mark_unscaled_by_comm(self) -> None:
        """The fused all-reduce epilogue (K1) already multiplied the gradients by 1/scale and
        accumulated ``found_inf``; ``step`` must not un-scale again."""
        self._unscaled = True

    def unscale_(self, optimizer) -> None:
        if not self._enabled or self._unscaled:
            return
        params = [p for g in optimizer.param_groups for p in g["params"] if p.grad is not None]
        if not params:
            return
        self._lazy_init(params[0].grad.device)
        arena = getattr(optimizer, "arena", None)
        grads = [arena.grads] if arena is not None else [p.grad for p in params]
        for g in grads:
            # non-finite check before the multiply so that inf*0 cannot hide as nan-only
            self._found_inf.add_((~torch.isfinite(g)).any().to(torch.float32)).clamp_(max=1.0)
            g.mul_(self._inv_scale.to(g.dtype))
        self._unscaled = True

    def step(self, optimizer, *args, **kwargs):
        if not self._enabled:
            return optimizer.step(*args, **kwargs)
        params = [p for g in optimizer.param_groups for p in g["params"]]
        self._lazy_init(params[0].device)
        if hasattr(optimizer, "_step_arena"):
            # Our SGD: un-scale and skip-on-inf happen inside the fused step, on the device.
            inv = None if self._unscaled else self._inv_scale
            if not self._unscaled:
                arena = getattr(optimizer, "arena", None)
                grads = [arena.grads] if arena is not None else [p.grad for p in params if p.grad is not None]
                if grads and grads[0].is_cuda:
                    from . import ops
                    if ops.available():
                        for g in grads:
                            if g.numel() % 4 == 0 and g.is_contiguous():
                                ops.check_finite(g, self._found_inf)        # K13, one read of the arena
                            else:
                                self._found_inf.add_((~torch.isfinite(g)).any().to(torch.float32)).clamp_(max=1.0)
                    else:
                        for g in grads:
                            self._found_inf.add_((~torch.isfinite(g)).any().to(torch.float32)).clamp_(max=1.0)
                else:
                    for g in grads:
                        self._found_inf.add_((~torch.isfinite(g)).any().to(torch.float32)).clamp_(max=1.0)
            self._unscaled = True
            return optimizer.step(*args, inv_scale=inv, found_inf=self._found_inf, **kwargs)
        # Foreign optimizer: torch semantics, host sync included.
        self.unscale_(optimizer)
        if float(self._found_inf.item()) == 0.0:
            return optimizer.step(*args, **kwargs)
        return None

    def update(self, new_scale: Optional[float] = None) -> None:
        if not self._enabled or self._scale is None:
            return
        if new_scale is None and self._scale.is_cuda and _native_ops() is not None:
            # K13: one device thread does the whole update (scale, tracker, 1/scale, flag reset)
            _native_ops().scaler_update(self._scale, self._growth_tracker, self._found_inf, self._inv_scale,
                                        self._growth_factor, self._backoff_factor, self._growth_interval)
            self._unscaled = False
            return
        if new_scale is not None:
            self._scale.fill_(float(new_scale))
        else:
            # torch:amp/grad_scaler.py:484-559 / ATen AmpKernels amp_update_scale, as tensor selects.
            inf = self._found_inf > 0
            grown_tracker = self._growth_tracker + 1
            grow = grown_tracker >= self._growth_interval
            cand = self._scale * self._growth_factor
            cand = torch.where(torch.isfinite(cand), cand, self._scale)
            new_scale_t = torch.where(inf, self._scale * self._backoff_factor,
                                      torch.where(grow, cand, self._scale))
            new_tracker = torch.where(inf | grow, torch.zeros_like(grown_tracker), grown_tracker)
            self._scale.copy_(new_scale_t)
            self._growth_tracker.copy_(new_tracker)
        self._inv_scale.copy_(1.0 / self._scale.double())
        self._found_inf.zero_()
        self._unscaled = False

    def state_dict(self):
        if not self._enabled:
            return {}
        return {"scale": self.get_scale(), "growth_factor": self._growth_factor,
                "backoff_factor": self._backoff_factor, "growth_interval": self._growth_interval,
                "_growth_tracker": 0 if self._growth_tracker is None else int(self._growth_tracker.item())}

    def load_state_dict(self, sd) -> None:
        if not self._enabled or not sd:
            return
        self._init_scale = float(sd["scale"])
        self._growth_factor = float(sd["growth_factor"])
        self._backoff_factor = float(sd["backoff_factor"])
        self._growth_interval = int(sd["growth_interval"])
        if self._scale is not None:
            self._scale.fill_(self._init_scale)
            self._inv_scale.fill_(1.0 / self._init_scale)
            self._growth_tracker.fill_(int(sd["_growth_tracker"]))
        else:
            self._pending_tracker = int(sd.get("_growth_tracker", 0))
