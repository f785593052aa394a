"""Hyper-parameter block and flag system.

The reference has no flags: every script hard-codes the same literals at the top of
``main()`` (ref:resnet_single_gpu.py:70-78, ref:resnet_dp.py:67-77,
ref:restnet_ddp.py:76-84, ref:resnet_ddp_apex.py:80-88).  Here those literals are the
*defaults* of one dataclass, and an argparse layer can override them for benchmarking
(SURVEY.md §5.6).  Running a script with no flags reproduces the reference settings.
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from pathlib import Path
from typing import Optional, Sequence

MODES = ("single", "dp", "ddp", "ddp_amp")

_SAVE_DIRS = {
    "single": "output/resnet_single",   # ref:resnet_single_gpu.py:77
    "dp": "output/resnet_dp",           # ref:resnet_dp.py:74
    "ddp": "output/resnet_ddp",         # ref:restnet_ddp.py:83
    "ddp_amp": "output/resnet_ddp_amp", # ref:resnet_ddp_apex.py:87
}


@dataclasses.dataclass
class TrainConfig:
    mode: str = "single"
    # -- reference literals ------------------------------------------------------
    epochs: int = 100
    batch_size: int = 400          # per GPU; DP mode multiplies by #GPUs (ref:resnet_dp.py:69)
    num_workers: int = 4
    lr: float = 0.1
    momentum: float = 0.9
    weight_decay: float = 1e-4
    step_size: int = 30            # StepLR (ref:restnet_ddp.py:123)
    gamma: float = 0.1
    save_path: Optional[str] = None
    # -- bench / portability knobs (not in the reference) ------------------------
    synthetic: bool = True         # no ImageNet on this box; data.py makes ImageNet-shaped tensors
    data_root: Optional[str] = None   # <root>/{train,val}/<class>/<image>: real data through data/folder.py
    image_size: int = 224
    num_classes: int = 1000
    train_samples: int = 1_281_167
    val_samples: int = 50_000
    max_steps: Optional[int] = None      # stop an epoch early (bench/tests)
    max_val_steps: Optional[int] = None
    dtype: str = "auto"            # auto -> the reference's precision: fp32 (TF32 GEMMs) for single/dp/ddp, fp16 for ddp_amp
    backend: str = "auto"          # auto -> nccl/native on CUDA, gloo on CPU
    engine: str = "auto"           # auto -> native sm_100a step graph on CUDA, stock torch ops on CPU
    sync_bn: bool = False
    empty_cache: bool = False      # reference calls synchronize()+empty_cache() every step (Q7)
    gpus: Optional[int] = None     # number of local devices (default: all visible)
    device: str = "auto"           # auto -> cuda if available else cpu
    seed: int = 0
    log_jsonl: Optional[str] = None
    profile: bool = False          # NVTX ranges per phase + per-step device timing in the JSONL log
    arch: str = "resnet50"
    width: int = 64                # tests shrink the net; 64 == torchvision resnet50
    layers: Sequence[int] = (3, 4, 6, 3)

    def __post_init__(self):
        if self.mode not in MODES:
            raise ValueError(f"mode must be one of {MODES}, got {self.mode!r}")
        if self.save_path is None:
            self.save_path = _SAVE_DIRS[self.mode]
        if self.data_root:
            self.synthetic = False

    @property
    def save_dir(self) -> Path:
        return Path(self.save_path)

    def resolved_dtype(self) -> str:
        """`auto` is the reference's precision for the mode: fp16 autocast for the amp script
        (ref:resnet_ddp_apex.py:27 ``torch.cuda.amp.autocast()``), fp32 storage with TF32 tensor-core math for the
        other three (ref:resnet_single_gpu.py:26-30 et al.: no autocast, cuDNN's TF32 default).  On a CUDA device
        all three run on the native sm_100a step graph (``kind::f16`` / ``kind::tf32`` tcgen05 GEMMs); ``--dtype
        bf16`` selects bf16 activations with fp32 master weights for any mode."""
        if self.dtype != "auto":
            return self.dtype
        return "fp16" if self.mode == "ddp_amp" else "fp32"


def add_arguments(p: argparse.ArgumentParser, mode: str) -> argparse.ArgumentParser:
    d = TrainConfig(mode=mode)
    p.add_argument("--epochs", type=int, default=d.epochs)
    p.add_argument("--batch-size", type=int, default=d.batch_size)
    p.add_argument("--workers", dest="num_workers", type=int, default=d.num_workers)
    p.add_argument("--lr", type=float, default=d.lr)
    p.add_argument("--momentum", type=float, default=d.momentum)
    p.add_argument("--weight-decay", type=float, default=d.weight_decay)
    p.add_argument("--save-path", type=str, default=d.save_path)
    p.add_argument("--synthetic", action=argparse.BooleanOptionalAction, default=True)
    p.add_argument("--data-root", type=str, default=None,
                   help="image-folder ImageNet (<root>/{train,val}/<class>/<image>); implies --no-synthetic")
    p.add_argument("--image-size", type=int, default=d.image_size)
    p.add_argument("--num-classes", type=int, default=d.num_classes)
    p.add_argument("--train-samples", type=int, default=d.train_samples)
    p.add_argument("--val-samples", type=int, default=d.val_samples)
    p.add_argument("--steps", dest="max_steps", type=int, default=None)
    p.add_argument("--val-steps", dest="max_val_steps", type=int, default=None)
    p.add_argument("--dtype", choices=("auto", "fp32", "bf16", "fp16"), default="auto")
    p.add_argument("--backend", choices=("auto", "native", "nccl", "gloo"), default="auto")
    p.add_argument("--engine", choices=("auto", "native", "torch"), default="auto")
    p.add_argument("--sync-bn", action="store_true")
    p.add_argument("--empty-cache", action=argparse.BooleanOptionalAction, default=False)
    p.add_argument("--gpus", type=int, default=None)
    p.add_argument("--device", choices=("auto", "cuda", "cpu"), default="auto")
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--log-jsonl", type=str, default=None)
    p.add_argument("--profile", action="store_true", help="NVTX ranges per phase, per-step device time in --log-jsonl")
    p.add_argument("--width", type=int, default=64)
    return p


def from_args(mode: str, argv: Optional[Sequence[str]] = None) -> TrainConfig:
    p = argparse.ArgumentParser(description=f"ResNet-50 training, mode={mode}")
    add_arguments(p, mode)
    ns = p.parse_args(argv)
    return TrainConfig(mode=mode, **vars(ns))


def multi_node_addressing(local_rank: int, gpus: int, env=os.environ):
    """Global rank arithmetic of the reference's DDP scripts (ref:restnet_ddp.py:87-94).

    ``WORLD_SIZE`` is the number of *hosts* and ``RANK`` the *host* index (quirk Q12).
    Returns ``(init_method, world_size, global_rank, host_rank)``.
    """
    ip = env.get("MASTER_IP", "127.0.0.1")
    port = env.get("MASTER_PORT", "29500")
    hosts = int(env.get("WORLD_SIZE", "1"))
    host_rank = int(env.get("RANK", "0"))
    return f"tcp://{ip}:{port}", hosts * gpus, host_rank * gpus + local_rank, host_rank
