"""``hfai.nn.parallel.DistributedDataParallel`` (ref:restnet_ddp.py:16,99)."""
from pytorch_distributed_b200.parallel import DataParallel, DistributedDataParallel  # noqa: F401
