"""SyncBatchNorm: batch statistics over every replica of the data-parallel group.

Not used by the reference scripts but named by BASELINE.json.  Delegate
algorithm (torch:nn/modules/_functions.py:7-209): local per-channel statistics → all-gather of
``[mean, invstd, count]`` → combine → normalise; backward: local ``[sum_dy, sum_dy_xmu]`` →
all-reduce → dx.  Here the exchanged quantities are the additive ``[sum, sum_sq, count]`` so a
single all-reduce suffices in each direction.  This module is the stock-op implementation (CPU / ``--engine torch``);
on the native engine the exchange happens inside the kernels that already hold the sums — the last CTA of the
producing convolution (forward) and the folding block of the BN backward reduction — over peer memory (K6,
``csrc/comm/peer_sync.cuh::exchange_sum``): zero extra launches and zero NCCL calls for the 53 BN layers.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch import nn


class _SyncBNFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, eps, momentum, group):
        c = x.size(1)
        dims = [0] + list(range(2, x.dim()))
        xf = x.float()
        local = torch.cat([xf.sum(dims), (xf * xf).sum(dims),
                           torch.tensor([float(x.numel() // c)], device=x.device)])
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(local, group=group)
        n = local[-1]
        mean = local[:c] / n
        var = (local[c:2 * c] / n - mean * mean).clamp_(min=0)
        invstd = torch.rsqrt(var + eps)
        if running_mean is not None:
            with torch.no_grad():
                running_mean.mul_(1 - momentum).add_(mean, alpha=momentum)
                running_var.mul_(1 - momentum).add_(var * (n / (n - 1).clamp(min=1)), alpha=momentum)
        shape = [1, c] + [1] * (x.dim() - 2)
        xhat = (xf - mean.view(shape)) * invstd.view(shape)
        ctx.save_for_backward(xhat, weight, invstd, n)
        ctx.group, ctx.dims, ctx.shape = group, dims, shape
        return (xhat * weight.view(shape) + bias.view(shape)).to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        xhat, weight, invstd, n = ctx.saved_tensors
        dyf = dy.float()
        sum_dy = dyf.sum(ctx.dims)
        sum_dy_xhat = (dyf * xhat).sum(ctx.dims)
        both = torch.cat([sum_dy, sum_dy_xhat])
        if dist.is_initialized() and dist.get_world_size(ctx.group) > 1:
            dist.all_reduce(both, group=ctx.group)
        c = sum_dy.numel()
        g_dy, g_dyx = both[:c] / n, both[c:] / n
        dx = (dyf - g_dy.view(ctx.shape) - xhat * g_dyx.view(ctx.shape)) * (weight * invstd).view(ctx.shape)
        # weight/bias gradients stay local: DDP's gradient all-reduce averages them like any other
        return dx.to(dy.dtype), sum_dy_xhat, sum_dy, None, None, None, None, None


class SyncBatchNorm(nn.BatchNorm2d):
    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True,
                 process_group=None):
        super().__init__(num_features, eps, momentum, affine, track_running_stats)
        self.process_group = process_group

    def forward(self, x):
        if not self.training:
            return super().forward(x)
        if self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        return _SyncBNFunction.apply(x, self.weight, self.bias, self.running_mean, self.running_var,
                                     self.eps, self.momentum if self.momentum is not None else 0.1,
                                     self.process_group)


def convert_sync_batchnorm(module: nn.Module, process_group=None) -> nn.Module:
    """Replace every BatchNorm2d by SyncBatchNorm, keeping parameters, buffers and names
    (torch:nn/modules/batchnorm.py:845)."""
    out = module
    if isinstance(module, nn.BatchNorm2d) and not isinstance(module, SyncBatchNorm):
        out = SyncBatchNorm(module.num_features, module.eps, module.momentum, module.affine,
                            module.track_running_stats, process_group)
        if module.affine:
            with torch.no_grad():
                out.weight, out.bias = module.weight, module.bias
        out.running_mean, out.running_var = module.running_mean, module.running_var
        out.num_batches_tracked = module.num_batches_tracked
        out.training = module.training
    for name, child in module.named_children():
        new = convert_sync_batchnorm(child, process_group)
        if new is not child:
            out.add_module(name, new)
    return out
