"""Device time of the stem kernels at batch 400: space-to-depth TMA path vs the explicit im2col matrix path."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from pytorch_distributed_b200 import ops
from pytorch_distributed_b200.ops import ConvShape
from bench_tools.conv_tune import timeit

N, H, dt = 400, 224, torch.bfloat16
x = torch.randn(N, 3, H, H, device="cuda")
w16 = (torch.randn(64, 7, 7, 3, device="cuda") * 0.08).to(dt)
xs = torch.empty(N, 112, 112, 16, device="cuda", dtype=dt)
ws = torch.empty(64, 256, device="cuda", dtype=dt)
y = torch.empty(N, 112, 112, 64, device="cuda", dtype=dt)
dy = torch.randn(N, 112, 112, 64, device="cuda").to(dt)
st, scr = torch.zeros(128, device="cuda"), torch.zeros(16 * 128 + 32, device="cuda")
dws, dw = torch.zeros(64, 256, device="cuda"), torch.zeros(64, 147, device="cuda")
ops.stem_weight_s2d(w16, ws)
r = {}
r["stem_s2d"] = timeit(lambda: ops.stem_s2d(x, xs))
r["stem_fprop"] = timeit(lambda: ops.stem_fprop(xs, ws, y, stats=st, stats_scratch=scr))
r["stem_wgrad"] = timeit(lambda: ops.stem_wgrad(xs, dy, dws))
rows = N * 112 * 112
col = torch.empty(rows, 152, device="cuda", dtype=dt)
wp = torch.zeros(64, 152, device="cuda", dtype=dt)
r["im2col"] = timeit(lambda: ops.stem_im2col(x, col))
sh = ConvShape(rows, 1, 1, 152, 64, 1, 1, 1, 0)
r["im2col_fprop"] = timeit(lambda: ops.conv_fprop(col, wp, y, sh, stats=st, stats_scratch=scr))
r["im2col_wgrad"] = timeit(lambda: ops.conv_wgrad(col, dy.view(rows, 64), dw, sh, ldw=147, kvalid=147))
print({k: round(v * 1e3, 1) for k, v in r.items()}, "us;  s2d path total",
      round((r["stem_s2d"] + r["stem_fprop"] + r["stem_wgrad"]) * 1e3), "us vs im2col path",
      round((r["im2col"] + r["im2col_fprop"] + r["im2col_wgrad"]) * 1e3), "us")
