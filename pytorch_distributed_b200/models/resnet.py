"""ResNet-50 (v1.5) with torchvision-compatible parameter names.

The reference builds ``torchvision.models.resnet50()`` (ref:resnet_single_gpu.py:83).  The
architecture spec is tv:models/resnet.py:108-163 (Bottleneck, stride on the 3×3),
``:166-285`` (stem 7×7/2 → BN → ReLU → maxpool 3×3/2 → stages [3,4,6,3] → avgpool → fc) and
``:208-213`` (kaiming-normal fan_out conv init, BN weight 1 / bias 0).  This module declares
the same module tree, so ``state_dict()`` has torchvision's 320 keys in the same order and
checkpoints move freely between the two — but execution is different: on a CUDA device the
module hands the whole forward/backward to the native sm_100a step graph
(models/native_resnet.py); the ``nn.Module`` tree is only the parameter/bookkeeping façade.
The stock-op ``forward`` below is the CPU path and the numerics oracle.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.nn.functional as F
from torch import nn


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * self.expansion, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * self.expansion)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)


class ResNet(nn.Module):
    def __init__(self, layers: Sequence[int] = (3, 4, 6, 3), num_classes: int = 1000, width: int = 64,
                 zero_init_residual: bool = False):
        super().__init__()
        self.layers_cfg = tuple(layers)
        self.width = width
        self.num_classes = num_classes
        self.inplanes = width
        self.conv1 = nn.Conv2d(3, width, 7, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, stride=2, padding=1)
        self.layer1 = self._make_layer(width, layers[0], 1)
        self.layer2 = self._make_layer(width * 2, layers[1], 2)
        self.layer3 = self._make_layer(width * 4, layers[2], 2)
        self.layer4 = self._make_layer(width * 8, layers[3], 2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(width * 8 * Bottleneck.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1.0)
                nn.init.constant_(m.bias, 0.0)
        if zero_init_residual:
            for m in self.modules():
                if isinstance(m, Bottleneck):
                    nn.init.constant_(m.bn3.weight, 0.0)
        self._engine = None           # set by enable_native()

    def _make_layer(self, planes: int, blocks: int, stride: int) -> nn.Sequential:
        downsample = None
        if stride != 1 or self.inplanes != planes * Bottleneck.expansion:
            downsample = nn.Sequential(
                nn.Conv2d(self.inplanes, planes * Bottleneck.expansion, 1, stride=stride, bias=False),
                nn.BatchNorm2d(planes * Bottleneck.expansion))
        mods: List[nn.Module] = [Bottleneck(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * Bottleneck.expansion
        for _ in range(1, blocks):
            mods.append(Bottleneck(self.inplanes, planes))
        return nn.Sequential(*mods)

    # -- execution ---------------------------------------------------------------------
    def forward_stock(self, x: torch.Tensor) -> torch.Tensor:
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = torch.flatten(self.avgpool(x), 1)
        return self.fc(x)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        eng = self._engine
        if eng is not None and x.is_cuda:
            return eng(x)
        return self.forward_stock(x)

    def enable_native(self, batch_size: int, dtype: torch.dtype = torch.bfloat16, **kw):
        """Bind the sm_100a step graph.  Fails loudly if the CUDA library is missing: a silent
        fallback to stock ops on a GPU box would hide that the product path is not running."""
        from .native_resnet import NativeResNetEngine
        object.__setattr__(self, "_engine", NativeResNetEngine(self, batch_size=batch_size, dtype=dtype, **kw))
        return self

    def disable_native(self):
        object.__setattr__(self, "_engine", None)
        return self


def resnet50(num_classes: int = 1000, **kw) -> ResNet:
    return ResNet((3, 4, 6, 3), num_classes=num_classes, **kw)
