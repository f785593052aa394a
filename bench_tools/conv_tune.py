"""Sweep tile / split choices of the tcgen05 conv kernels over the ResNet-50 shapes at batch 400.

  python bench_tools/conv_tune.py --out gpurun_out/conv_tune.json
Prints, per shape, the device time of every variant (CUDA events, 3 warm-ups, mean of 10) and the best.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from pytorch_distributed_b200 import ops  # noqa: E402
from pytorch_distributed_b200.ops import ConvShape  # noqa: E402

# (H, C, K, R, stride) at batch 400; 1x1 convs appear both ways (fprop C->K, dgrad K->C)
FPROP = [(56, 64, 64, 1, 1), (56, 64, 64, 3, 1), (56, 64, 256, 1, 1), (56, 256, 64, 1, 1), (56, 256, 128, 1, 1),
         (56, 128, 128, 3, 2), (56, 128, 128, 3, 1), (28, 128, 128, 3, 1), (28, 128, 512, 1, 1), (28, 512, 128, 1, 1),
         (56, 256, 512, 1, 2), (28, 512, 256, 1, 1), (28, 256, 256, 3, 2), (14, 256, 256, 3, 1), (14, 256, 1024, 1, 1),
         (14, 1024, 256, 1, 1), (14, 1024, 512, 1, 1), (14, 512, 512, 3, 2), (7, 512, 512, 3, 1), (7, 512, 2048, 1, 1),
         (7, 2048, 512, 1, 1), (14, 1024, 2048, 1, 2)]
WGRAD = [(56, 64, 64, 1, 1), (56, 64, 64, 3, 1), (56, 64, 256, 1, 1), (56, 256, 64, 1, 1), (56, 128, 128, 3, 2),
         (28, 128, 128, 3, 1), (28, 128, 512, 1, 1), (28, 512, 128, 1, 1), (28, 256, 256, 3, 2), (14, 256, 256, 3, 1),
         (14, 256, 1024, 1, 1), (14, 1024, 256, 1, 1), (14, 512, 512, 3, 2), (7, 512, 512, 3, 1), (7, 512, 2048, 1, 1),
         (7, 2048, 512, 1, 1)]


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/conv_tune.json")
    ap.add_argument("--batch", type=int, default=400)
    ap.add_argument("--pair", action="store_true", help="fprop only: compare the 256-wide tile on one CTA vs a CTA pair")
    ap.add_argument("--fprop-only", action="store_true")
    args = ap.parse_args()
    only_pair = args.pair
    N, dt = args.batch, torch.bfloat16
    out = {"fprop": [], "wgrad": []}
    for (H, C, K, R, stride) in FPROP:
        pad = 1 if R == 3 else 0
        P = (H + 2 * pad - R) // stride + 1
        x = torch.randn(N, H, H, C, device="cuda", dtype=dt)
        w = torch.randn(K, R, R, C, device="cuda", dtype=dt)
        y = torch.empty(N, P, P, K, device="cuda", dtype=dt)
        st = torch.zeros(2 * K, device="cuda")
        sh = ConvShape(N, H, H, C, K, R, R, stride, pad)
        flops = 2.0 * N * P * P * K * C * R * R
        rec = {"shape": [H, C, K, R, stride], "variants": {}}
        for bn in (64, 128, 256, 512):            # 512 = 256-wide tile on a CTA pair (cta_group::2)
            if (bn if bn < 512 else 256) > max(64, K) or (only_pair and bn < 256):
                continue
            ms = timeit(lambda: ops.conv_fprop(x, w, y, sh, stats=st, block_n=bn))
            rec["variants"][f"bn{bn}"] = {"ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1)}
        if not rec["variants"]:
            continue
        best = min(rec["variants"].items(), key=lambda kv: kv[1]["ms"])
        rec["best"] = best[0]
        out["fprop"].append(rec)
        print("fprop", rec, flush=True)
        del x, w, y
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    for (H, C, K, R, stride) in ([] if (only_pair or args.fprop_only) else WGRAD):
        pad = 1 if R == 3 else 0
        P = (H + 2 * pad - R) // stride + 1
        x = torch.randn(N, H, H, C, device="cuda", dtype=dt)
        dy = torch.randn(N, P, P, K, device="cuda", dtype=dt)
        dw = torch.zeros(K, R, R, C, device="cuda")
        sh = ConvShape(N, H, H, C, K, R, R, stride, pad)
        flops = 2.0 * N * P * P * K * C * R * R
        rec = {"shape": [H, C, K, R, stride], "variants": {}}
        ktot = R * R * C
        for bn in (64, 128, 256):
            if bn > max(64, ktot):
                continue
            tiles = ((K + 127) // 128) * ((ktot + bn - 1) // bn)
            for waves in (0.5, 1, 2, 4):
                splits = max(1, int(round(waves * sms / tiles)))
                ms = timeit(lambda: ops.conv_wgrad(x, dy, dw, sh, splits=splits, block_n=bn))
                rec["variants"][f"bn{bn}_w{waves}_s{splits}"] = {"ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1)}
        best = min(rec["variants"].items(), key=lambda kv: kv[1]["ms"])
        rec["best"] = best[0]
        out["wgrad"].append(rec)
        print("wgrad", {"shape": rec["shape"], "best": best}, flush=True)
        del x, dy, dw
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
