"""``hfai.multiprocessing.spawn(fn, args, nprocs, bind_numa=True)`` (ref:restnet_ddp.py:155)."""
from pytorch_distributed_b200.launcher import (ProcessExitedException, ProcessRaisedException,  # noqa: F401
                                               spawn, start_processes)
