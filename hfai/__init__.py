"""``hfai`` compatibility shim backed by pytorch_distributed_b200.

The reference imports High-Flyer's closed ``hfai`` package for data, the suspend client, the
process launcher and thin wrappers over torch.distributed / DDP (ref:restnet_ddp.py:14-16,36,107,
155).  These modules expose the same names so reference-style scripts run unchanged on top of the
B200-native framework.
"""
from pytorch_distributed_b200 import suspend as client
from pytorch_distributed_b200.suspend import go_suspend, receive_suspend_command  # Q2: hfai.go_suspend()

from . import datasets, multiprocessing, nccl, nn  # noqa: E402

__all__ = ["client", "datasets", "multiprocessing", "nccl", "nn", "go_suspend", "receive_suspend_command"]
