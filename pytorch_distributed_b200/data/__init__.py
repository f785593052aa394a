from .sampler import DistributedSampler
from .dataset import ImageNet, SyntheticImageNet, BatchLoader
from . import transforms

__all__ = ["DistributedSampler", "ImageNet", "SyntheticImageNet", "BatchLoader", "transforms"]
