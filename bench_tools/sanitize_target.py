"""Tiny single-GPU workload for compute-sanitizer (memcheck / racecheck / synccheck): one native training step
at batch 8, 64x64 plus the standalone kernels.   compute-sanitizer --tool memcheck python bench_tools/sanitize_target.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from pytorch_distributed_b200 import amp, harness, ops  # noqa: E402
from pytorch_distributed_b200.config import TrainConfig  # noqa: E402
from pytorch_distributed_b200.nn import CrossEntropyLoss, topk_correct  # noqa: E402
from pytorch_distributed_b200.optim import SGD  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    cfg = TrainConfig(mode="ddp_amp", batch_size=8, dtype="bf16", device="cuda")
    model = harness.build_model(cfg, dev)
    model.enable_native(batch_size=8, dtype=torch.bfloat16)
    opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    sc = amp.GradScaler(device=dev)
    crit = CrossEntropyLoss()
    x = torch.randn(8, 3, 64, 64, device=dev)
    y = torch.randint(0, 1000, (8,), device=dev)
    model.train()
    loss = harness.train_step(model, crit, opt, x, y, cfg, dev, sc, stock_autocast=False)
    model.eval()
    with torch.no_grad():
        out = model(x)
        topk_correct(out, y)
    sums = torch.empty(128, device=dev)
    ops.bn_stats(torch.randn(1000, 64, device=dev).bfloat16(), sums)
    torch.cuda.synchronize()
    print("sanitize target ok, loss", float(loss))


if __name__ == "__main__":
    main()
