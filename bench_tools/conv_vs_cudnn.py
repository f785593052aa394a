"""Every ResNet-50 convolution shape at batch 400: this library's tcgen05 kernels vs cuDNN on the same box.

SURVEY.md section 7.2 step 7 names this comparison as the exit test of the convolution work: all distinct shapes x
{fprop, dgrad, wgrad}, our kernel against `torch.nn.functional.conv2d` / its backward on channels_last tensors (cuDNN
9.x picks its own Blackwell kernels), same dtype, CUDA events, 3 warm-ups + mean of 10, clocks recorded.  Notes on
what is compared: our fprop time INCLUDES the BatchNorm statistics (sum / sum of squares per channel) that its epilogue
produces; cuDNN's does not (the stock model needs a separate statistics pass for them).  Our stride-2 3x3 dgrad runs
on a zero-upsampled dY and is charged for the upsample kernel too.

  python bench_tools/conv_vs_cudnn.py [--dtype bf16|fp16|fp32] [--out gpurun_out/conv_vs_cudnn.json] [--md profiles/CONV_VS_CUDNN.md]
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from pytorch_distributed_b200 import ops  # noqa: E402
from pytorch_distributed_b200.ops import ConvShape  # noqa: E402

# (H, C, K, R, stride): the 23 distinct convolutions of torchvision's resnet50 at 224x224 (stem listed separately)
SHAPES = [(56, 64, 64, 1, 1), (56, 64, 64, 3, 1), (56, 64, 256, 1, 1), (56, 256, 64, 1, 1), (56, 256, 128, 1, 1),
          (56, 128, 128, 3, 2), (28, 128, 512, 1, 1), (56, 256, 512, 1, 2), (28, 512, 128, 1, 1), (28, 128, 128, 3, 1),
          (28, 512, 256, 1, 1), (28, 256, 256, 3, 2), (14, 256, 1024, 1, 1), (28, 512, 1024, 1, 2), (14, 1024, 256, 1, 1),
          (14, 256, 256, 3, 1), (14, 1024, 512, 1, 1), (14, 512, 512, 3, 2), (7, 512, 2048, 1, 1), (14, 1024, 2048, 1, 2),
          (7, 2048, 512, 1, 1), (7, 512, 512, 3, 1)]
DT = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}


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


def clocks():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active",
                              "--format=csv,noheader", "-i", "0"], capture_output=True, text=True, timeout=10).stdout.strip()
        return out
    except Exception as e:
        return f"unavailable: {e}"


def measure(N, H, C, K, R, stride, dt):
    pad = 1 if R == 3 else 0
    P = (H + 2 * pad - R) // stride + 1
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(N, H, H, C, device="cuda", generator=g).to(dt)
    w = (torch.randn(K, R, R, C, device="cuda", generator=g) / (R * R * C) ** 0.5).to(dt)
    dy = torch.randn(N, P, P, K, device="cuda", generator=g).to(dt)
    flops = 2.0 * N * P * P * K * C * R * R
    rec = {"shape": f"{C}->{K} {R}x{R}/{stride} @{H}", "gflop": round(flops / 1e9, 1)}
    # ---- ours
    y = torch.empty(N, P, P, K, device="cuda", dtype=dt)
    st = torch.zeros(2 * K, device="cuda")
    scr = torch.zeros(int(ops.require().pdb_conv_stats_scratch_floats(K)), device="cuda")
    shape = ConvShape(N, H, H, C, K, R, R, stride, pad)

    def ours_fprop():                      # as the engine runs it: BatchNorm statistics in the epilogue
        scr.zero_()
        ops.conv_fprop(x, w, y, shape, stats=st, stats_scratch=scr)

    def ours_fprop_plain():                # the convolution alone, which is all cuDNN's kernel computes
        ops.conv_fprop(x, w, y, shape)
    wt = torch.empty(C, R, R, K, device="cuda", dtype=dt)
    ops.weight_dgrad_layout(w, wt, K, R, R, C)
    dx = torch.empty(N, H, H, C, device="cuda", dtype=dt)
    if R == 1:
        dxc = dx if stride == 1 else torch.empty(N, P, P, C, device="cuda", dtype=dt)

        def ours_dgrad():
            ops.conv_fprop(dy, wt, dxc, ConvShape(N, P, P, K, C, 1, 1, 1, 0))
            if stride == 2:
                ops.upsample2_zero(dxc, dx)
    else:
        up = dy if stride == 1 else torch.empty(N, H, H, K, device="cuda", dtype=dt)

        def ours_dgrad():
            if stride == 2:
                ops.upsample2_zero(dy, up)
            ops.conv_fprop(up, wt, dx, ConvShape(N, H, H, K, C, R, R, 1, R - 1 - pad))
    dw = torch.zeros(K, R, R, C, device="cuda")

    def ours_wgrad():
        ops.conv_wgrad(x, dy, dw, shape)
    # ---- cuDNN through torch (channels_last views of the same NHWC data)
    xc, wc, dyc = x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), dy.permute(0, 3, 1, 2)

    def cudnn_fprop():
        F.conv2d(xc, wc, stride=stride, padding=pad)

    def cudnn_bwd(mask):
        return lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [stride, stride], [pad, pad], [1, 1], False,
                                                           [0, 0], 1, mask)
    for name, a, b in (("fprop", ours_fprop, cudnn_fprop), ("dgrad", ours_dgrad, cudnn_bwd([True, False, False])),
                       ("wgrad", ours_wgrad, cudnn_bwd([False, True, False]))):
        to, tc = timeit(a), timeit(b)
        rec[name] = {"ours_us": round(1e3 * to, 1), "cudnn_us": round(1e3 * tc, 1), "ours_tflops": round(flops / to / 1e9, 1),
                     "cudnn_tflops": round(flops / tc / 1e9, 1), "speedup": round(tc / to, 3)}
    rec["fprop"]["ours_plain_us"] = round(1e3 * timeit(ours_fprop_plain), 1)
    # what the stock model pays for the same result: cuDNN's convolution + a statistics pass over its output
    sums = torch.empty(2 * K, device="cuda")
    rec["fprop"]["stats_pass_us"] = round(1e3 * timeit(lambda: ops.bn_stats(y.view(-1, K), sums)), 1)
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", choices=tuple(DT), default="bf16")
    ap.add_argument("--batch", type=int, default=400)
    ap.add_argument("--out", default="gpurun_out/conv_vs_cudnn.json")
    ap.add_argument("--md", default=None, help="also write a markdown table here")
    args = ap.parse_args()
    dt = DT[args.dtype]
    torch.backends.cudnn.benchmark = True              # let cuDNN pick its best algorithm per shape
    torch.backends.cudnn.allow_tf32 = True             # fp32: TF32 on both sides (torch's default, what the reference runs)
    torch.backends.cuda.matmul.allow_tf32 = True
    out = {"dtype": args.dtype, "batch": args.batch, "gpu": torch.cuda.get_device_name(0), "cudnn": torch.backends.cudnn.version(),
           "clocks_before": clocks(), "rows": []}
    for (H, C, K, R, stride) in SHAPES:
        rec = measure(args.batch, H, C, K, R, stride, dt)
        out["rows"].append(rec)
        print(rec["shape"], {k: (v["ours_us"], v["cudnn_us"]) for k, v in rec.items() if isinstance(v, dict)}, flush=True)
        torch.cuda.empty_cache()
    out["clocks_after"] = clocks()
    tot = {k: [sum(r[k]["ours_us"] for r in out["rows"]), sum(r[k]["cudnn_us"] for r in out["rows"])] for k in ("fprop", "dgrad", "wgrad")}
    out["totals_us"] = tot
    out["slower_than_cudnn"] = [(r["shape"], k, r[k]["speedup"]) for r in out["rows"] for k in ("fprop", "dgrad", "wgrad") if r[k]["speedup"] < 1.0]
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)
    if args.md:
        lines = [f"## {args.dtype}, batch {args.batch}, {out['gpu']}, cuDNN {out['cudnn']}", "",
                 f"clocks before: `{out['clocks_before']}`; after: `{out['clocks_after']}`", "",
                 "| shape | GFLOP | fprop+stats ours / cuDNN conv only us (x) | fprop alone ours us | a separate statistics pass us | dgrad ours / cuDNN us (x) | wgrad ours / cuDNN us (x) | ours TFLOPS f/d/w |",
                 "|---|---|---|---|---|---|---|---|"]
        for r in out["rows"]:
            cell = lambda k: f"{r[k]['ours_us']} / {r[k]['cudnn_us']} ({r[k]['speedup']}x)"          # noqa: E731
            lines.append(f"| {r['shape']} | {r['gflop']} | {cell('fprop')} | {r['fprop']['ours_plain_us']} | {r['fprop']['stats_pass_us']} | {cell('dgrad')} | {cell('wgrad')} | "
                         f"{r['fprop']['ours_tflops']} / {r['dgrad']['ours_tflops']} / {r['wgrad']['ours_tflops']} |")
        plain = sum(r["fprop"]["ours_plain_us"] for r in out["rows"])
        stp = sum(r["fprop"]["stats_pass_us"] for r in out["rows"])
        lines.append(f"| **sum** | | {tot['fprop'][0]:.0f} / {tot['fprop'][1]:.0f} | {plain:.0f} | {stp:.0f} | {tot['dgrad'][0]:.0f} / {tot['dgrad'][1]:.0f} | "
                     f"{tot['wgrad'][0]:.0f} / {tot['wgrad'][1]:.0f} | |")
        lines.append("")
        lines.append(f"slower than cuDNN: {out['slower_than_cudnn'] or 'none'}")
        with open(args.md, "a") as f:
            f.write("\n".join(lines) + "\n\n")
    print(json.dumps({"totals_us": tot, "n_slower": len(out["slower_than_cudnn"])}))


if __name__ == "__main__":
    main()
