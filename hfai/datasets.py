"""``hfai.datasets.ImageNet(split, transform).loader(...)`` (ref:restnet_ddp.py:107-119).

Synthetic ImageNet-shaped data by default; with ``PDB200_DATA_ROOT=<dir>`` in the environment the same call
reads a real image-folder tree ``<dir>/{train,val}/<class>/<image>`` (the reference's ffrecord files are only
readable on its own cluster).  ``transform`` may be a torchvision pipeline, as in the reference scripts.
"""
import os

from pytorch_distributed_b200.data import ImageNet as _ImageNet, SyntheticImageNet  # noqa: F401


def ImageNet(split="train", transform=None, **kw):
    root = kw.pop("root", None) or os.environ.get("PDB200_DATA_ROOT")
    if root and "synthetic" not in kw:
        return _ImageNet(split, transform, synthetic=False, root=root, **kw)
    return _ImageNet(split, transform, **kw)
