"""ResNet-50 under single-process DataParallel — entry point with the reference's name.

Counterpart of ref:resnet_dp.py (132 lines): one process drives every local GPU, global batch =
400 × #GPUs assembled on GPU 0 (ref:resnet_dp.py:69,77-82), checkpoints under output/resnet_dp.
"""
import hf_env
hf_env.set_env('202111')

import sys

from pytorch_distributed_b200 import config, harness


def main(argv=None):
    cfg = config.from_args("dp", argv)
    return harness.main(cfg, local_rank=0)


if __name__ == '__main__':
    main(sys.argv[1:])
