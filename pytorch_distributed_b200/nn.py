"""Loss and metric modules that route to the fused sm_100a kernels on CUDA.

``CrossEntropyLoss`` = mean-reduced log-softmax + NLL over ``[B, classes]`` logits
(torch:nn/modules/loss.py:1194) as used by ref:restnet_ddp.py:121; on CUDA it is one fused
forward+backward kernel with fp32 math (K10) that can also apply the amp loss scale to the
gradient.  ``topk_correct`` is the validation metric of ref:restnet_ddp.py:58-60 as a fused
"rank of the label logit" counter (K14).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn


class _FusedXent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        from . import ops
        loss, dlogits = ops.xent_fwd_bwd(logits, labels)          # dlogits: fp32 (softmax - onehot)/B
        ctx.save_for_backward(dlogits)
        ctx.logits_dtype = logits.dtype
        return loss

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        # g carries the amp loss scale (65536 does not even fit fp16): multiply in fp32, round once afterwards
        return (dlogits * g.float()).to(ctx.logits_dtype), None


class CrossEntropyLoss(nn.Module):
    def forward(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        if logits.is_cuda and logits.dim() == 2 and logits.dtype in (torch.bfloat16, torch.float16, torch.float32):
            from . import ops
            if ops.available():
                return _FusedXent.apply(logits, labels)
        return F.cross_entropy(logits.float(), labels)


def topk_correct(logits: torch.Tensor, labels: torch.Tensor, ks=(1, 5)) -> torch.Tensor:
    """Returns a float tensor ``[len(ks)]`` with the number of rows whose label is in the top-k."""
    if logits.is_cuda and logits.dim() == 2:
        from . import ops
        if ops.available():
            return ops.topk_correct(logits, labels, ks)
    kmax = min(max(ks), logits.size(-1))      # fewer than five classes: top-5 is "any class" (torch.topk would raise)
    _, preds = logits.topk(kmax, -1, True, True)
    hit = torch.eq(preds, labels.unsqueeze(1))
    return torch.stack([hit[:, :k].sum() for k in ks]).to(torch.float32)
