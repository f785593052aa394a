"""Scatter / gather along the batch dimension with autograd, for the single-process DP mode.

Spec: torch:nn/parallel/_functions.py:82-155 (``Gather``/``Scatter``: each is the other's
backward) and torch:nn/parallel/scatter_gather.py:49-150 (recursion over tuples / lists / dicts,
non-tensors replicated by reference).  On CUDA the copies are peer-memory *pull* kernels
(K4a/K4c, `peer_pull_kernel` in csrc/comm/collectives.cu): the destination GPU reads its chunk straight from the source
GPU's memory over NVLink, instead of the source pushing seven cudaMemcpyPeerAsync in a row.
"""
from __future__ import annotations

from typing import Any, List, Sequence

import torch


def _chunk_sizes(n: int, parts: int) -> List[int]:
    # torch.chunk semantics: ceil-sized chunks, possibly fewer than `parts`
    size = -(-n // parts)
    out = []
    while n > 0:
        out.append(min(size, n))
        n -= size
    return out


def _copy_to(t: torch.Tensor, device: torch.device) -> torch.Tensor:
    if t.device == device:
        return t
    if t.is_cuda and device.type == "cuda":
        from .. import ops
        if ops.available():
            return ops.peer_copy(t, device)
    return t.to(device, non_blocking=True)


class Scatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, devices, dim, x):
        ctx.dim, ctx.src_device = dim, x.device
        sizes = _chunk_sizes(x.size(dim), len(devices))
        chunks = x.split(sizes, dim)
        return tuple(_copy_to(c.contiguous() if c.device != d else c, d) for c, d in zip(chunks, devices))

    @staticmethod
    def backward(ctx, *grads):
        return None, None, Gather.apply(ctx.src_device, ctx.dim, *grads)


class Gather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, target, dim, *xs):
        ctx.dim, ctx.devices = dim, [x.device for x in xs]
        ctx.scalar = all(x.dim() == 0 for x in xs)
        if ctx.scalar:          # torch warns and unsqueezes (torch:nn/parallel/_functions.py:94-104)
            xs = tuple(x.view(1) for x in xs)
        ctx.sizes = [x.size(dim) for x in xs]
        return torch.cat([_copy_to(x, target) for x in xs], dim)

    @staticmethod
    def backward(ctx, g):
        parts = g.split(ctx.sizes, ctx.dim)
        outs = tuple(_copy_to(p.contiguous() if p.device != d else p, d) for p, d in zip(parts, ctx.devices))
        if ctx.scalar:
            outs = tuple(o[0] for o in outs)
        return (None, None) + outs


def scatter(obj: Any, devices: Sequence[torch.device], dim: int = 0) -> List[Any]:
    n = len(devices)

    def rec(o):
        if torch.is_tensor(o):
            return list(Scatter.apply(list(devices), dim, o))
        if isinstance(o, tuple) and o:
            return [tuple(z) for z in zip(*map(rec, o))]
        if isinstance(o, list) and o:
            return [list(z) for z in zip(*map(rec, o))]
        if isinstance(o, dict) and o:
            return [type(o)(z) for z in zip(*map(rec, o.items()))]
        return [o for _ in range(n)]

    return rec(obj)


def scatter_kwargs(inputs, kwargs, devices, dim: int = 0):
    si = scatter(inputs, devices, dim) if inputs else []
    sk = scatter(kwargs, devices, dim) if kwargs else []
    m = min([len(x) for x in (si, sk) if x] or [len(devices)])   # fewer chunks than devices is legal
    si = list(si[:m]) + [()] * (m - len(si)) if si else [() for _ in range(m)]
    sk = list(sk[:m]) + [{}] * (m - len(sk)) if sk else [{} for _ in range(m)]
    return si, sk


def gather(outputs: Sequence[Any], target: torch.device, dim: int = 0):
    def rec(outs):
        o = outs[0]
        if torch.is_tensor(o):
            return Gather.apply(target, dim, *outs)
        if o is None:
            return None
        if isinstance(o, dict):
            return type(o)((k, rec([d[k] for d in outs])) for k in o)
        return type(o)(map(rec, zip(*outs)))

    return rec(list(outputs))
