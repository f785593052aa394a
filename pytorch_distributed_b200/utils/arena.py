"""Arenas: parameters, gradients, momentum and BN buffers as single flat allocations.

The reference's delegates keep 161 separate parameter tensors and pay for it three
times: DDP flattens gradients into buckets and copies them back
(torch:include/torch/csrc/distributed/c10d/reducer.hpp:276-286), DataParallel
coalesces 10 MiB chunks to broadcast/reduce (torch:nn/parallel/comm.py:50-168), and the
optimizer walks a tensor list (torch:optim/sgd.py:383-477).  Here every ``nn.Parameter``
is a *view* into one contiguous fp32 arena and ``.grad`` is a view into a second arena
of the same geometry, so

* a gradient bucket is an address range — no flatten, no copy-in/copy-out,
* the SGD step is one kernel over three flat arrays,
* a peer-memory broadcast / reduce is one kernel over one range,

while ``state_dict()`` / ``load_state_dict()`` / ``torch.save`` still see torchvision's
names and shapes (hard part 5).  Convolution weights are stored
physically as [Cout, kh, kw, Cin] (the K-major operand layout the tcgen05 implicit
GEMM wants) and exposed as a channels_last-strided [Cout, Cin, kh, kw] view.
"""
from __future__ import annotations

from typing import Dict, List, NamedTuple, Optional

import torch
from torch import nn

ALIGN_ELEMS = 64  # 256 B for fp32: keeps every tensor 16 B-aligned for TMA / 128-bit access


class ArenaEntry(NamedTuple):
    name: str
    offset: int      # in elements
    numel: int
    shape: torch.Size
    channels_last: bool


def _round_up(n: int, a: int) -> int:
    return (n + a - 1) // a * a


def want_symmetric(device: torch.device) -> bool:
    """Arenas are placed in NVLink-symmetric memory when they will take part in peer-memory collectives:
    CUDA device, NCCL process group (used only for the handle exchange), more than one rank."""
    import os
    if os.environ.get("PDB200_SYMM", "1") == "0" or torch.device(device).type != "cuda":
        return False
    try:
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized() and dist.get_backend() == "nccl" and dist.get_world_size() > 1
    except Exception:
        return False


def alloc_flat(n: int, dtype: torch.dtype, device: torch.device, symmetric: bool):
    """Zero-initialised flat buffer; returns (tensor, is_symmetric).  Symmetric allocations must happen in
    the same order on every rank (they do: every rank builds the same model)."""
    if symmetric:
        try:
            import torch.distributed._symmetric_memory as symm_mem
            t = symm_mem.empty(n, dtype=dtype, device=device)
            t.zero_()
            return t, True
        except Exception:
            pass
    return torch.zeros(n, dtype=dtype, device=device), False


def _view(flat: torch.Tensor, e: ArenaEntry) -> torch.Tensor:
    seg = flat[e.offset: e.offset + e.numel]
    if e.channels_last:
        o, i, h, w = e.shape
        return seg.view(o, h, w, i).permute(0, 3, 1, 2)
    return seg.view(e.shape)


class ParamArena:
    def __init__(self, module: nn.Module, device: Optional[torch.device] = None,
                 channels_last_conv: bool = True):
        params = [(n, p) for n, p in module.named_parameters()]
        if not params:
            raise ValueError("module has no parameters")
        device = torch.device(device) if device is not None else params[0][1].device
        self.device = device
        self.entries: List[ArenaEntry] = []
        off = 0
        for n, p in params:
            cl = channels_last_conv and p.dim() == 4
            self.entries.append(ArenaEntry(n, off, p.numel(), p.shape, cl))
            off = _round_up(off + p.numel(), ALIGN_ELEMS)
        self.numel = off
        symmetric = want_symmetric(device)
        self.params, self.params_symmetric = alloc_flat(off, torch.float32, device, symmetric)
        self.grads, self.grads_symmetric = alloc_flat(off, torch.float32, device, symmetric)
        self.momentum: Optional[torch.Tensor] = None
        self.shadow: Optional[torch.Tensor] = None     # 16-bit copy of the weights for the MMA path
        self.by_name: Dict[str, ArenaEntry] = {e.name: e for e in self.entries}
        self._parameters: List[nn.Parameter] = []
        with torch.no_grad():
            for e, (_, p) in zip(self.entries, params):
                v = _view(self.params, e)
                v.copy_(p.detach().to(device=device, dtype=torch.float32))
                p.data = v
                p.grad = _view(self.grads, e)
                p._pdb200_arena = self
                self._parameters.append(p)
        module._pdb200_arena = self

    # -- views -------------------------------------------------------------------
    def param_view(self, name: str) -> torch.Tensor:
        return _view(self.params, self.by_name[name])

    def grad_view(self, name: str) -> torch.Tensor:
        return _view(self.grads, self.by_name[name])

    def shadow_view(self, name: str) -> torch.Tensor:
        e = self.by_name[name]
        return self.shadow[e.offset: e.offset + e.numel]

    def parameters(self) -> List[nn.Parameter]:
        return self._parameters

    # -- state -------------------------------------------------------------------
    def ensure_momentum(self) -> torch.Tensor:
        if self.momentum is None:
            self.momentum = torch.zeros_like(self.params)
        return self.momentum

    def ensure_shadow(self, dtype: torch.dtype) -> torch.Tensor:
        if self.shadow is None or self.shadow.dtype != dtype:
            self.shadow = self.params.to(dtype)
        return self.shadow

    def refresh_shadow(self) -> None:
        if self.shadow is not None:
            self.shadow.copy_(self.params)

    def zero_grad(self) -> None:
        self.grads.zero_()

    def rebind_grads(self) -> None:
        """Re-attach ``.grad`` views (a foreign ``zero_grad(set_to_none=True)`` drops them)."""
        for e, p in zip(self.entries, self._parameters):
            if p.grad is None or p.grad.data_ptr() != self.grads.data_ptr() + 4 * e.offset:
                p.grad = _view(self.grads, e)

    def bucket_ranges(self, cap_bytes: int = 25 << 20, first_cap_bytes: int = 1 << 20):
        """Contiguous [start, end) element ranges walking parameters in *reverse* order —
        the order gradients become ready (torch:nn/parallel/distributed.py:1252-1258) — with
        torch's 1 MiB-first / 25 MiB-cap policy (reducer.hpp:30-31).  Returns ranges in
        launch order (last layers first)."""
        ranges, cur_end, cur_start, cap = [], None, None, first_cap_bytes
        for e in reversed(self.entries):
            end = _round_up(e.offset + e.numel, ALIGN_ELEMS)
            if cur_end is None:
                cur_end = end
            cur_start = e.offset
            if (cur_end - cur_start) * 4 >= cap:
                ranges.append((cur_start, cur_end))
                cur_end, cap = None, cap_bytes
        if cur_end is not None:
            ranges.append((cur_start, cur_end))
        return ranges


class BufferArena:
    """BN running statistics in one fp32 range (+ the int64 counters in another), so the
    per-iteration DDP buffer broadcast (torch:nn/parallel/distributed.py:2214-2257) is a single
    contiguous transfer instead of flatten→broadcast→unflatten of 159 tensors."""

    def __init__(self, module: nn.Module, device: Optional[torch.device] = None):
        bufs = list(module.named_buffers())
        f = [(n, b) for n, b in bufs if b.is_floating_point()]
        i = [(n, b) for n, b in bufs if not b.is_floating_point()]
        device = torch.device(device) if device is not None else (bufs[0][1].device if bufs else torch.device("cpu"))
        self.float_entries: List[ArenaEntry] = []
        off = 0
        for n, b in f:
            self.float_entries.append(ArenaEntry(n, off, b.numel(), b.shape, False))
            off = _round_up(off + b.numel(), 4)
        n_float = _round_up(max(off, 1), 4)
        self.int_entries: List[ArenaEntry] = []
        ioff = 0
        for n, b in i:
            self.int_entries.append(ArenaEntry(n, ioff, b.numel(), b.shape, False))
            ioff += b.numel()
        n_int = max(ioff, 1)
        # one allocation: fp32 statistics followed by the int64 counters (viewed through the same bytes),
        # so the per-iteration DDP buffer broadcast is a single transfer
        total = _round_up(n_float + 2 * n_int, 4)
        self.storage, self.symmetric = alloc_flat(total, torch.float32, device, want_symmetric(device))
        self.floats = self.storage[:n_float]
        # The counters are updated with real in-place ops (num_batches_tracked += 1).  A plain view would
        # share the storage's autograd version counter with running_mean/var, which BN saves for backward,
        # and every later layer's increment would look like an illegal in-place modification.  A DLPack
        # round trip aliases the same bytes with an independent version counter.
        from torch.utils import dlpack
        ints_alias = dlpack.from_dlpack(dlpack.to_dlpack(self.storage[n_float: n_float + 2 * n_int]))
        self.ints = ints_alias.view(torch.int64)
        owners = {}
        for mod_name, mod in module.named_modules():
            for bn, _ in mod.named_buffers(recurse=False):
                owners[(mod_name + "." if mod_name else "") + bn] = (mod, bn)
        with torch.no_grad():
            for e, (n, b) in zip(self.float_entries, f):
                v = self.floats[e.offset: e.offset + e.numel].view(e.shape)
                v.copy_(b)
                mod, bn = owners[n]
                mod._buffers[bn] = v
            for e, (n, b) in zip(self.int_entries, i):
                v = self.ints[e.offset: e.offset + e.numel].view(e.shape)
                v.copy_(b)
                mod, bn = owners[n]
                mod._buffers[bn] = v
        self.by_name = {e.name: e for e in self.float_entries + self.int_entries}
        module._pdb200_buffer_arena = self
