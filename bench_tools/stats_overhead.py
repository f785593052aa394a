"""How much do the per-channel statistics atomics cost the conv epilogue?  (with vs without stats)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from pytorch_distributed_b200 import ops
from pytorch_distributed_b200.ops import ConvShape
from bench_tools.conv_tune import timeit

for (H, C, K, R) in [(14, 256, 1024, 1), (14, 1024, 256, 1), (14, 256, 256, 3), (28, 128, 512, 1), (56, 64, 256, 1), (56, 64, 64, 3), (7, 512, 2048, 1)]:
    N, pad = 400, (1 if R == 3 else 0)
    x = torch.randn(N, H, H, C, device="cuda").bfloat16()
    w = torch.randn(K, R, R, C, device="cuda").bfloat16()
    y = torch.empty(N, H, H, K, device="cuda", dtype=torch.bfloat16)
    st = torch.zeros(2 * K, device="cuda")
    sh = ConvShape(N, H, H, C, K, R, R, 1, pad)
    a = timeit(lambda: ops.conv_fprop(x, w, y, sh, stats=st))
    b = timeit(lambda: ops.conv_fprop(x, w, y, sh, stats=None))
    scr = torch.zeros(16 * 2 * K + 32, device="cuda")
    tick = scr[16 * 2 * K:]
    c = timeit(lambda: (tick.zero_(), ops.conv_fprop(x, w, y, sh, stats=st, stats_scratch=scr)))
    print(f"{(H, C, K, R)}: stats(atomics) {a*1e3:7.1f} us   stats(replicas, incl. 128 B memset) {c*1e3:7.1f} us   "
          f"no stats {b*1e3:7.1f} us", flush=True)
