"""Reference-arm shims: bind the closed `hfai` names the reference imports to STOCK PyTorch
(torch.distributed, torch DDP, torch.multiprocessing).  None of pytorch_distributed_b200 is used
here — this is the baseline the product is measured against."""
from . import client, datasets, multiprocessing, nccl, nn  # noqa: F401


def go_suspend():
    raise SystemExit(75)
