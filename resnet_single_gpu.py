"""ResNet-50 on one GPU — entry point with the reference's name and defaults.

Counterpart of ref:resnet_single_gpu.py (137 lines).  Running it with no flags uses the
reference's hyper-parameters (batch 400, lr 0.1, momentum 0.9, wd 1e-4, 100 epochs, checkpoints
under output/resnet_single); flags exist only for benchmarking (see config.py).
"""
import hf_env
hf_env.set_env('202111')

import os
import sys

os.environ.setdefault("CUDA_VISIBLE_DEVICES", "0")     # ref:resnet_single_gpu.py:14 (quirk Q13)

from pytorch_distributed_b200 import config, harness


def main(argv=None):
    cfg = config.from_args("single", argv)
    return harness.main(cfg, local_rank=0)


if __name__ == '__main__':
    main(sys.argv[1:])
