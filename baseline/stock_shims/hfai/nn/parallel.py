from torch.nn.parallel import DistributedDataParallel  # noqa: F401
