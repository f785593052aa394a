"""SGD with momentum and weight decay — one fused pass over the parameter arena.

Math (torch:optim/sgd.py:322-380, dampening 0, no nesterov as in ref:restnet_ddp.py:122):
``g = grad + wd·p``; first step ``buf = g``, later ``buf = m·buf + g``; ``p -= lr·buf``.
The reference's delegate runs this as ~4 foreach launches over 161 tensors
(torch:optim/sgd.py:383-477).  With arena-backed parameters (utils/arena.py) it is a
single kernel over three flat arrays that also, optionally, un-scales the gradient,
honours a device-side ``found_inf`` flag (no ``.item()`` host sync — compare
torch:amp/grad_scaler.py:363-373), writes the 16-bit weight shadow the tensor-core path
consumes, and zeroes the gradient arena for the next step.

``state_dict()`` is interchangeable with ``torch.optim.SGD``'s
(torch:optim/optimizer.py:684): per-parameter ``momentum_buffer`` entries are views into
the momentum arena.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch.optim import Optimizer

from ..utils.arena import ParamArena, _view


class SGD(Optimizer):
    def __init__(self, params, lr: float = 1e-3, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False, *, maximize: bool = False,
                 arena: Optional[ParamArena] = None):
        if lr < 0.0 or momentum < 0.0 or weight_decay < 0.0:
            raise ValueError("lr, momentum and weight_decay must be non-negative")
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                        nesterov=nesterov, maximize=maximize, foreach=None, differentiable=False,
                        fused=None)
        params = list(params)
        super().__init__(params, defaults)
        self.arena = arena
        if arena is None:
            self.arena = self._discover_arena()
        self._first_step_done = False
        self.fuse_zero_grad = False     # harness may enable: step() also clears the grad arena

    def _discover_arena(self) -> Optional[ParamArena]:
        """Parameters re-homed by ParamArena are recognisable by their storage."""
        ps = [p for g in self.param_groups for p in g["params"]]
        if not ps or len(self.param_groups) != 1:
            return None
        arena = getattr(ps[0], "_pdb200_arena", None)
        if arena is None:
            return None
        # The fused step walks the WHOLE flat arena, so it is only valid when this optimizer was given exactly the
        # arena's parameters and all of them train; SGD(model.fc.parameters()) or a frozen layer takes the
        # per-tensor path and touches nothing else (torch.optim.SGD semantics).
        own = arena.parameters()
        if len(ps) != len(own) or {id(p) for p in ps} != {id(p) for p in own} or not all(p.requires_grad for p in own):
            return None
        return arena

    # -- grads ---------------------------------------------------------------------
    def zero_grad(self, set_to_none: bool = True) -> None:
        if self.arena is not None:
            # ``.grad`` must keep aliasing the arena (SURVEY.md hard part 5): clear in place.
            self.arena.rebind_grads()
            self.arena.zero_grad()
            return
        super().zero_grad(set_to_none=set_to_none)

    # -- step ----------------------------------------------------------------------
    @torch.no_grad()
    def step(self, closure=None, *, inv_scale: Optional[torch.Tensor] = None,
             found_inf: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self.arena is not None:
            self._step_arena(inv_scale, found_inf)
        else:
            self._step_tensors(inv_scale, found_inf)
        return loss

    def _step_arena(self, inv_scale, found_inf) -> None:
        a, g = self.arena, self.param_groups[0]
        first = a.momentum is None and g["momentum"] != 0
        mom = a.ensure_momentum() if g["momentum"] != 0 else None
        if a.params.is_cuda:
            from .. import ops
            ops.sgd_step(a.params, a.grads, mom, a.shadow, lr=g["lr"], momentum=g["momentum"],
                         dampening=g["dampening"], weight_decay=g["weight_decay"],
                         nesterov=g["nesterov"], maximize=g["maximize"], first_step=first,
                         inv_scale=inv_scale, found_inf=found_inf, zero_grad=self.fuse_zero_grad)
        else:
            _flat_sgd(a.params, a.grads, mom, lr=g["lr"], momentum=g["momentum"],
                      dampening=g["dampening"], weight_decay=g["weight_decay"], nesterov=g["nesterov"],
                      maximize=g["maximize"], first_step=first, inv_scale=inv_scale, found_inf=found_inf)
            a.refresh_shadow()
            if self.fuse_zero_grad:
                a.grads.zero_()
        if mom is not None and first:
            for e, p in zip(a.entries, a.parameters()):
                self.state[p]["momentum_buffer"] = _view(mom, e)

    def _step_tensors(self, inv_scale, found_inf) -> None:
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                first = "momentum_buffer" not in st and g["momentum"] != 0
                if first:
                    st["momentum_buffer"] = torch.zeros_like(p)
                _flat_sgd(p, p.grad, st.get("momentum_buffer"), lr=g["lr"], momentum=g["momentum"],
                          dampening=g["dampening"], weight_decay=g["weight_decay"],
                          nesterov=g["nesterov"], maximize=g["maximize"], first_step=first,
                          inv_scale=inv_scale, found_inf=found_inf)

    # -- state_dict: momentum buffers must land back inside the arena ----------------
    def load_state_dict(self, state_dict) -> None:
        super().load_state_dict(state_dict)
        a = self.arena
        if a is None:
            return
        loaded = [self.state.get(p, {}).get("momentum_buffer") for p in a.parameters()]
        if any(b is not None for b in loaded):
            mom = a.ensure_momentum()
            for e, p, b in zip(a.entries, a.parameters(), loaded):
                v = _view(mom, e)
                if b is not None:
                    v.copy_(b)
                self.state[p]["momentum_buffer"] = v


def _flat_sgd(p, grad, buf, *, lr, momentum, dampening, weight_decay, nesterov, maximize, first_step,
              inv_scale=None, found_inf=None) -> None:
    """Reference implementation on plain tensors (CPU path and numerics oracle for K9)."""
    g = grad if not maximize else -grad
    if inv_scale is not None:
        g = g * inv_scale.to(g.dtype)
    if weight_decay != 0:
        g = g.add(p, alpha=weight_decay)
    if momentum != 0:
        new_buf = g.clone() if first_step else buf.mul(momentum).add_(g, alpha=1 - dampening)
        d = g.add(new_buf, alpha=momentum) if nesterov else new_buf
    else:
        new_buf, d = None, g
    new_p = p.add(d, alpha=-lr)
    if found_inf is not None:
        keep = found_inf.reshape(()).to(torch.bool)          # device-side skip, no host sync
        new_p = torch.where(keep, p, new_p)
        if new_buf is not None:
            new_buf = torch.where(keep, buf, new_buf)
    p.copy_(new_p)
    if new_buf is not None:
        buf.copy_(new_buf)
