from .distributed import DistributedDataParallel
from .data_parallel import DataParallel
from .sync_bn import SyncBatchNorm, convert_sync_batchnorm
from . import comm

__all__ = ["DistributedDataParallel", "DataParallel", "SyncBatchNorm", "convert_sync_batchnorm", "comm"]
