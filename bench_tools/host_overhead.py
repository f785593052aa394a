"""Is the training step host-bound?  Times how long the host needs to ENQUEUE a step (no synchronisation) against
the device time of the step.  python bench_tools/host_overhead.py [--steps 30]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from pytorch_distributed_b200 import amp, harness
from pytorch_distributed_b200.config import TrainConfig
from pytorch_distributed_b200.nn import CrossEntropyLoss
from pytorch_distributed_b200.optim import SGD

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--batch", type=int, default=400)
args = ap.parse_args()
dev = torch.device("cuda", 0)
cfg = TrainConfig(mode="ddp_amp", batch_size=args.batch, dtype="bf16", device="cuda")
model = harness.build_model(cfg, dev)
model.enable_native(batch_size=args.batch, dtype=torch.bfloat16)
opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
sc = amp.GradScaler(device=dev)
crit = CrossEntropyLoss()
x = torch.randn(args.batch, 3, 224, 224, device=dev)
y = torch.randint(0, 1000, (args.batch,), device=dev)
model.train()
for _ in range(5):
    harness.train_step(model, crit, opt, x, y, cfg, dev, sc, stock_autocast=False)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(args.steps):
    harness.train_step(model, crit, opt, x, y, cfg, dev, sc, stock_autocast=False)
e1.record()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host enqueue {1e3 * (t1 - t0) / args.steps:.2f} ms/step   device {e0.elapsed_time(e1) / args.steps:.2f} ms/step   "
      f"wall {1e3 * (t2 - t0) / args.steps:.2f} ms/step")
import cProfile, pstats
pr = cProfile.Profile()
pr.enable()
for _ in range(10):
    harness.train_step(model, crit, opt, x, y, cfg, dev, sc, stock_autocast=False)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
