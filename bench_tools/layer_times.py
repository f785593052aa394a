"""Per-kernel-class device time of one native training step (CUDA events around every launch).

Serialises the step (an event pair per kernel), so compare SHARES, not the absolute total.
  python bench_tools/layer_times.py [--batch 400] [--out gpurun_out/layer_times.json]
"""
import argparse
import collections
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["PDB200_GRAPHS"] = "0"     # every kernel must be an eager launch to get its own event pair

import torch  # noqa: E402

from pytorch_distributed_b200 import amp, harness, ops  # noqa: E402
from pytorch_distributed_b200.config import TrainConfig  # noqa: E402
from pytorch_distributed_b200.nn import CrossEntropyLoss  # noqa: E402
from pytorch_distributed_b200.ops import elementwise as E  # noqa: E402
from pytorch_distributed_b200.optim import SGD  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=400)
    ap.add_argument("--out", default="gpurun_out/layer_times.json")
    ap.add_argument("--dtype", choices=("bf16", "fp16", "fp32"), default="bf16")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    cfg = TrainConfig(mode="ddp_amp" if args.dtype != "fp32" else "single", batch_size=args.batch, dtype=args.dtype, device="cuda")
    model = harness.build_model(cfg, dev)
    model.enable_native(batch_size=args.batch, dtype=harness.torch_dtype(cfg))
    opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    sc = amp.GradScaler(device=dev)
    crit = CrossEntropyLoss()
    x = torch.randn(args.batch, 3, 224, 224, device=dev)
    y = torch.randint(0, 1000, (args.batch,), device=dev)
    model.train()
    for _ in range(3):
        harness.train_step(model, crit, opt, x, y, cfg, dev, sc, stock_autocast=False)
    torch.cuda.synchronize()

    records = []
    names = [n for n in E.__all__ if callable(getattr(ops, n, None)) and n not in ("dtype_tag", "ConvShape")]
    originals = {}

    def wrap(name, fn):
        def inner(*a, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*a, **k)
            e1.record()
            tag = name
            if name in ("conv_fprop", "conv_wgrad"):
                s = a[3]
                tag = f"{name} {s.C}->{s.K} {s.R}x{s.S}/{s.stride} M={s.N * ((s.H + 2 * s.pad - s.R) // s.stride + 1) * ((s.W + 2 * s.pad - s.S) // s.stride + 1)}"
                flops = 2.0 * s.N * ((s.H + 2 * s.pad - s.R) // s.stride + 1) * ((s.W + 2 * s.pad - s.S) // s.stride + 1) * s.K * s.C * s.R * s.S
            else:
                flops = 0.0
                if name.startswith("bn_") or name in ("maxpool_fwd", "maxpool_bwd", "add_strided2", "avgpool_bwd"):
                    t = a[2] if name in ("bn_bwd_reduce", "bn_bwd_apply") else a[0]
                    tag = f"{name} {tuple(t.shape)}" + (" +res" if name == "bn_apply" and a[1] is not None else "") \
                        + (" +dres" if name == "bn_bwd_apply" and a[4] is not None else "")
                    # bytes moved per activation stream: for a GB/s column
                    n = t.numel() * t.element_size() // 2
                    streams = {"bn_apply": 2 + (a[1] is not None), "bn_bwd_reduce": 2.06,
                               "bn_bwd_apply": 3.06 + (a[4] is not None if name == "bn_bwd_apply" else 0)}.get(name, 0)
                    flops = -2.0 * n * streams          # negative => bytes, reported as GB/s
            records.append((name, tag, e0, e1, flops))
            return r
        return inner

    import pytorch_distributed_b200.models.native_resnet as NR
    import pytorch_distributed_b200.optim.sgd as SG
    for n in names:
        originals[n] = getattr(ops, n)
        setattr(ops, n, wrap(n, originals[n]))
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record()
    harness.train_step(model, crit, opt, x, y, cfg, dev, sc, stock_autocast=False)
    a1.record()
    torch.cuda.synchronize()
    for n in names:
        setattr(ops, n, originals[n])
    by_class = collections.OrderedDict()
    by_tag = collections.OrderedDict()
    for name, tag, e0, e1, flops in records:
        ms = e0.elapsed_time(e1)
        c = by_class.setdefault(name, [0, 0.0, 0.0])
        c[0] += 1; c[1] += ms; c[2] += flops
        t = by_tag.setdefault(tag, [0, 0.0, 0.0])
        t[0] += 1; t[1] += ms; t[2] += flops
    total = a0.elapsed_time(a1)
    out = {"batch": args.batch, "dtype": args.dtype, "step_ms_serialised": total, "launches": len(records),
           "by_class": {k: {"n": v[0], "ms": round(v[1], 3), "tflops": round(v[2] / v[1] / 1e9, 1) if v[1] > 0 and v[2] > 0 else None}
                        for k, v in sorted(by_class.items(), key=lambda kv: -kv[1][1])},
           "by_shape": {k: {"n": v[0], "ms": round(v[1], 3),
                            "tflops": round(v[2] / v[1] / 1e9, 1) if v[1] > 0 and v[2] > 0 else None,
                            "GBps": round(-v[2] / v[1] / 1e6, 0) if v[1] > 0 and v[2] < 0 else None}
                        for k, v in sorted(by_tag.items(), key=lambda kv: -kv[1][1])[:90]}}
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps({"step_ms_serialised": total, "launches": len(records)}))
    for k, v in list(out["by_class"].items())[:14]:
        print(f"{k:22s} n={v['n']:4d} ms={v['ms']:9.3f} tflops={v['tflops']}")


if __name__ == "__main__":
    main()
