from torch.distributed import *  # noqa: F401,F403
from torch.distributed import get_rank, get_world_size, init_process_group, reduce  # noqa: F401
