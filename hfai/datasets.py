"""``hfai.datasets.ImageNet(split, transform).loader(...)`` (ref:restnet_ddp.py:107-119)."""
from pytorch_distributed_b200.data import ImageNet, SyntheticImageNet  # noqa: F401
