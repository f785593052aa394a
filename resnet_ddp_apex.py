"""ResNet-50 under DDP + mixed precision with dynamic loss scaling ("Apex" in the reference).

Counterpart of ref:resnet_ddp_apex.py (159 lines).  Launch exactly like the reference: python resnet_ddp_apex.py on every host with
MASTER_IP, MASTER_PORT, WORLD_SIZE (number of *hosts*) and RANK (*host* index) in the environment
(ref:restnet_ddp.py:87-94, quirk Q12; all default to a single local host).  One child per local
GPU is spawned with NUMA binding (ref:restnet_ddp.py:153-155).
"""
import hf_env
hf_env.set_env('202111')

import sys

import torch

import hfai
from pytorch_distributed_b200 import config, harness


def main(local_rank, cfg):
    return harness.main(cfg, local_rank)


if __name__ == '__main__':
    cfg = config.from_args("ddp_amp", sys.argv[1:])
    ngpus = cfg.gpus or (torch.cuda.device_count() if torch.cuda.is_available() and cfg.device != "cpu" else 2)
    cfg.gpus = ngpus
    hfai.multiprocessing.spawn(main, args=(cfg,), nprocs=ngpus, bind_numa=True)
