from .sampler import DistributedSampler
from .dataset import ImageNet, SyntheticImageNet, BatchLoader
from .folder import ImageFolder, FolderLoader
from . import transforms

__all__ = ["DistributedSampler", "ImageNet", "SyntheticImageNet", "BatchLoader", "ImageFolder", "FolderLoader", "transforms"]
