"""Tensor-native image transforms with the reference's pipeline shape.

The reference composes torchvision PIL transforms (ref:resnet_single_gpu.py:87-103):
train = RandomResizedCrop(224) → RandomHorizontalFlip → ToTensor → Normalize,
val = Resize(256) → CenterCrop(224) → ToTensor → Normalize.  There is no PIL decode
in a synthetic pipeline, so these work directly on uint8/float CHW tensors and can be
applied per sample or per batch.
"""
from __future__ import annotations

import math
from typing import Callable, Optional, Sequence

import torch
import torch.nn.functional as F

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def _as_float_chw(img: torch.Tensor) -> torch.Tensor:
    if img.dtype == torch.uint8:
        return img.to(torch.float32)
    return img


def _resize(img: torch.Tensor, size) -> torch.Tensor:
    """Antialiased bilinear resize.  uint8 in -> uint8 out (rounded), like torchvision on uint8 tensors / PIL images,
    so that the ToTensor that follows still sees 8-bit pixels and rescales them to [0, 1]."""
    x = _as_float_chw(img).unsqueeze(0)
    x = F.interpolate(x, size=size, mode="bilinear", align_corners=False, antialias=True).squeeze(0)
    if img.dtype == torch.uint8:
        x = x.round_().clamp_(0, 255).to(torch.uint8)
    return x


class Compose:
    def __init__(self, transforms: Sequence[Callable]):
        self.transforms = list(transforms)

    def __call__(self, img, generator: Optional[torch.Generator] = None):
        for t in self.transforms:
            img = t(img, generator) if getattr(t, "random", False) else t(img)
        return img


class RandomResizedCrop:
    random = True

    def __init__(self, size: int, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0)):
        self.size, self.scale, self.ratio = size, scale, ratio

    def _params(self, h, w, g):
        area = h * w
        log_r = (math.log(self.ratio[0]), math.log(self.ratio[1]))
        for _ in range(10):
            u = torch.rand(2, generator=g)
            target = area * (self.scale[0] + (self.scale[1] - self.scale[0]) * u[0].item())
            ar = math.exp(log_r[0] + (log_r[1] - log_r[0]) * u[1].item())
            cw, ch = int(round(math.sqrt(target * ar))), int(round(math.sqrt(target / ar)))
            if 0 < cw <= w and 0 < ch <= h:
                i = int(torch.randint(0, h - ch + 1, (1,), generator=g))
                j = int(torch.randint(0, w - cw + 1, (1,), generator=g))
                return i, j, ch, cw
        ch, cw = h, w                      # fallback: whole image (torchvision does a centre crop)
        return 0, 0, ch, cw

    def __call__(self, img, generator=None):
        _, h, w = img.shape
        i, j, ch, cw = self._params(h, w, generator)
        return _resize(img[:, i:i + ch, j:j + cw], (self.size, self.size))


class RandomHorizontalFlip:
    random = True

    def __init__(self, p: float = 0.5):
        self.p = p

    def __call__(self, img, generator=None):
        if torch.rand(1, generator=generator).item() < self.p:
            return img.flip(-1)
        return img


class Resize:
    def __init__(self, size: int):
        self.size = size

    def __call__(self, img):
        _, h, w = img.shape
        if h <= w:
            nh, nw = self.size, max(1, int(self.size * w / h))
        else:
            nh, nw = max(1, int(self.size * h / w)), self.size
        return _resize(img, (nh, nw))


class CenterCrop:
    def __init__(self, size: int):
        self.size = size

    def __call__(self, img):
        _, h, w = img.shape
        i, j = max(0, (h - self.size) // 2), max(0, (w - self.size) // 2)
        return img[:, i:i + self.size, j:j + self.size]


class ToTensor:
    """uint8 [0,255] → float32 [0,1]; float input passes through."""

    def __call__(self, img):
        if img.dtype == torch.uint8:
            return img.to(torch.float32).div_(255.0)
        return img.to(torch.float32)


class Normalize:
    def __init__(self, mean=IMAGENET_MEAN, std=IMAGENET_STD):
        self.mean = torch.tensor(mean, dtype=torch.float32).view(-1, 1, 1)
        self.std = torch.tensor(std, dtype=torch.float32).view(-1, 1, 1)

    def __call__(self, img):
        return (img - self.mean) / self.std


def train_transform(size: int = 224) -> Compose:
    return Compose([RandomResizedCrop(size), RandomHorizontalFlip(), ToTensor(), Normalize()])


def val_transform(size: int = 224) -> Compose:
    return Compose([Resize(int(size * 256 / 224)), CenterCrop(size), ToTensor(), Normalize()])
