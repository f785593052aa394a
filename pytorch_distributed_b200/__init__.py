"""pytorch_distributed_b200 — a B200-native data-parallel ResNet training framework.

Same capabilities as HFAiLab/pytorch_distributed (four training modes: single GPU,
single-process DataParallel, multi-process DistributedDataParallel, DDP + mixed
precision), rebuilt for Blackwell: hand-written sm_100a CUDA kernels behind a C ABI
(``include/pdb200.h``), peer-memory collectives over NVLink 5, arenas instead of
tensor lists.  See DESIGN.md for the architecture and SURVEY.md for the parity map.
"""
from . import config  # noqa: F401

__version__ = "0.1.0"

__all__ = ["config", "__version__"]
