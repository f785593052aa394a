"""``hfai.nccl.distributed`` — the torch.distributed surface the reference touches
(init_process_group / get_rank / get_world_size / reduce, ref:restnet_ddp.py:36,63-67,94)."""
from torch.distributed import ReduceOp, all_reduce, barrier, broadcast  # noqa: F401

from pytorch_distributed_b200.parallel.comm import (destroy_process_group, get_rank, get_world_size,  # noqa: F401
                                                    init_process_group, is_initialized, reduce)
