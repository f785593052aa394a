"""profiles/ROOFLINE.md from measured files: per-kernel roofline fractions of one training step + the collectives.

  python bench_tools/roofline.py --layers profiles/layer_times_r2_fp16.json [profiles/layer_times_r2_tf32.json] \
      [--comm profiles/comm_sweep_8gpu_r2.json] [--steps profiles/step_profile_8gpu_r2.jsonl] > profiles/ROOFLINE.md

Denominators (MEASURED_PEAKS.json, driver-written): HBM copy bandwidth and the cuBLAS bf16 matmul rate.  A kernel that
bench_tools/layer_times.py times in isolation (an event pair per launch, eager) is judged against the BURST matmul rate,
not the sustained one: a 0.2 ms kernel never reaches the power-capped steady state that "sustained" describes.  TF32
has half the tensor-core rate of bf16 (B200: 1.1 vs 2.25 PFLOP/s nominal), so fp32-storage GEMMs are held to half the
measured bf16 figure.  Bytes are the algorithmic minimum of the pass in the layer's storage type; a stride-2 1x1
convolution reads only the pixels it uses (a quarter of the input).
"""
import argparse
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), float(p["bf16_tflops"]), float(p.get("bf16_tflops_sustained", p["bf16_tflops"]))
    except Exception:
        return 6546.9, 1703.2, 1422.7          # the values of this pool's MEASURED_PEAKS.json (fallback)


CONV = re.compile(r"conv_(fprop|wgrad) (\d+)->(\d+) (\d)x\d/(\d) M=(\d+)")
BN = re.compile(r"(bn_apply|bn_bwd_reduce|bn_bwd_apply) \((\d+), (\d+)\)( \+res| \+dres)?")


def conv_bytes(kind, C, K, R, stride, M, es):
    """Algorithmic bytes of one launch.  M = output pixels of the GEMM as launched (dgrad launches carry C/K swapped
    and M = the pixels they write)."""
    if R == 1:
        a = M * C * es                     # exactly the rows that are read (stride-2: the compact quarter)
    else:
        a = M * stride * stride * C * es   # every input pixel of a 3x3 window sweep is read once
    w = K * C * R * R * es
    if kind == "fprop":
        return a + w + M * K * es
    return a + M * K * es + K * C * R * R * 4          # wgrad: both operands + the fp32 gradient tile


def layer_tables(path, hbm, burst):
    d = json.load(open(path))
    fp32 = "tf32" in os.path.basename(path) or "fp32" in os.path.basename(path) or d.get("dtype") == "fp32"
    es, peak = (4, burst / 2) if fp32 else (2, burst)
    out = [f"## One step, batch {d.get('batch', 400)}, {'fp32 storage / TF32 GEMMs' if fp32 else '16-bit storage'} (`{os.path.relpath(path, ROOT)}`)", "",
           f"Serialised step {d['step_ms_serialised']:.1f} ms over {d['launches']} launches (an event pair per launch: compare shares).  "
           f"GEMM peak used: {peak:.0f} TFLOP/s ({'half of ' if fp32 else ''}cuBLAS bf16 burst {burst:.0f}); HBM copy peak {hbm:.0f} GB/s.", "",
           "### Convolutions (tcgen05 implicit GEMM): the larger of the two fractions is the roofline the layer sits on", "",
           "| kernel, shape | launches | ms | TFLOP/s | of GEMM peak | algorithmic GB/s | of HBM copy peak | bound |", "|---|---|---|---|---|---|---|---|"]
    bn_rows = []
    over = []
    for tag, v in d["by_shape"].items():
        m = CONV.match(tag)
        if m:
            kind, C, K, R, stride, M = m.group(1), *map(int, m.groups()[1:])
            n, ms = v["n"], v["ms"]
            tfl = 2.0 * M * K * C * R * R * n / ms / 1e9
            gbs = conv_bytes(kind, C, K, R, stride, M, es) * n / ms / 1e6
            fc, fm = tfl / peak, gbs / hbm
            if max(fc, fm) > 1.0:
                over.append((tag, round(fc, 2), round(fm, 2)))
            out.append(f"| `{tag}` | {n} | {ms:.3f} | {tfl:.0f} | {fc:.2f} | {gbs:.0f} | {fm:.2f} | {'compute' if fc >= fm else 'memory'} |")
            continue
        m = BN.match(tag)
        if m:
            name, rows, C, extra = m.group(1), int(m.group(2)), int(m.group(3)), m.group(4) or ""
            streams = {"bn_apply": 2 + (extra == " +res"), "bn_bwd_reduce": 2, "bn_bwd_apply": 3 + (extra == " +dres")}[name]
            mask = rows * C / 8.0                       # the 1-bit ReLU mask, read or written once
            gbs = (rows * C * es * streams + mask) * v["n"] / v["ms"] / 1e6
            bn_rows.append(f"| `{tag}` | {v['n']} | {v['ms']:.3f} | {gbs:.0f} | {gbs / hbm:.2f} |")
    out += ["", "### BatchNorm passes (memory-bound; bytes = activation streams read + written, plus the 1-bit mask)", "",
            "| kernel, tensor | launches | ms | GB/s | of HBM copy peak |", "|---|---|---|---|---|"] + bn_rows
    out += ["", "### Whole classes", "", "| class | launches | ms | TFLOP/s |", "|---|---|---|---|"]
    for k, v in d["by_class"].items():
        out.append(f"| {k} | {v['n']} | {v['ms']:.3f} | {v['tflops'] or ''} |")
    if over:
        out += ["", f"Fractions above 1.0: {over}.  Memory fractions: reads served by the 126 MB L2 (e.g. a tensor the previous kernel just "
                "wrote).  GEMM fractions: the denominator is the rate cuBLAS reached in the driver's 8192^3 measurement, not the hardware's "
                f"nominal {'1,125 TF32' if fp32 else '2,250 bf16'} TFLOP/s, and a 0.2-0.4 ms kernel runs at boost clocks the 8192^3 GEMM does not hold."]
    return out + [""]


def comm_table(path, link=770.0):
    d = json.load(open(path))
    w = d["world"]
    out = [f"## Collectives ({w} GPUs, idle-GPU sweep, `{os.path.relpath(path, ROOT)}`)", "",
           f"Bus bandwidth = 2(w-1)/w x bytes / device time (max over ranks), against the measured {link:.0f} GB/s per NVLink direction.  "
           "`default`: the algorithm `NativeComm.allreduce_range` picks for that size.", "",
           "| message | kind | one-shot us | P2P two-shot us | NVLS us | NCCL us | default | default kernel alone us | best bus GB/s | of link roofline |", "|---|---|---|---|---|---|---|---|---|---|"]
    for r in d["rows"]:
        best = max(r.get("p2p_busbw_GBs", 0), r.get("nvls_busbw_GBs", 0), r.get("oneshot_busbw_GBs", 0), r.get("default_kernel_busbw_GBs", 0))
        out.append(f"| {r['bytes'] / 1e6:.2f} MB | {'DDP bucket' if r['bucket'] else 'sweep'} | {r.get('oneshot_us', '')} | {r.get('p2p_us', '')} | "
                   f"{r.get('nvls_us', '')} | {r['nccl_us']} | {r.get('default_algo', '')} | {r.get('default_kernel_us', '')} | {best:.0f} | {best / link:.2f} |")
    return out + [""]


def step_table(path, link=770.0):
    recs = [json.loads(l) for l in open(path) if l.strip()]
    steps = [r for r in recs if r.get("event") == "step_time" and r.get("buckets")]
    if not steps:
        return []
    steps = steps[len(steps) // 2:]                    # the second half: CUDA graphs are warm
    nb = len(steps[0]["buckets"])
    out = [f"## Gradient buckets inside the training step (`{os.path.relpath(path, ROOT)}`, mean of {len(steps)} warm steps)", "",
           "Every bucket's all-reduce runs on the comm stream while backward keeps running on the compute stream.  `event us`: CUDA "
           "events around the launch.  The kernel also stamps the GPU's global timer at entry, after its first cross-rank meeting and "
           "at exit: `wait us` is the time until the SLOWEST rank's backward has reached this bucket (not communication), `collective "
           "us` is reduce-scatter + all-gather + the closing meeting, sharing SMs and HBM with the backward kernels; the roofline "
           "fraction is taken on that part.", "",
           "| bucket | MB | event us | wait us | collective us | collective bus GB/s | of link roofline |", "|---|---|---|---|---|---|---|"]
    for b in range(nb):
        mean = lambda k: sum(s["buckets"][b].get(k, 0.0) for s in steps) / len(steps)          # noqa: E731
        us = mean("us")
        by = steps[0]["buckets"][b]["bytes"]
        world = steps[0].get("world", 8)
        if "xfer_us" in steps[0]["buckets"][b]:
            xf = mean("xfer_us")
            bus = 2.0 * (world - 1) / world * by / xf / 1e3
            out.append(f"| {b} | {by / 1e6:.2f} | {us:.1f} | {mean('wait_us'):.1f} | {xf:.1f} | {bus:.0f} | {bus / link:.2f} |")
        else:
            bus = 2.0 * (world - 1) / world * by / us / 1e3
            out.append(f"| {b} | {by / 1e6:.2f} | {us:.1f} | | | {bus:.0f} (whole event) | {bus / link:.2f} |")
    ms = sum(s["ms"] for s in steps) / len(steps)
    out += ["", f"Step time (device, max over ranks) in the same run: {ms:.2f} ms.", ""]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", nargs="*", default=[])
    ap.add_argument("--comm", default=None)
    ap.add_argument("--steps", default=None)
    ap.add_argument("--peaks", default=None, help="hbm_gbs,bf16_burst_tflops,bf16_sustained_tflops instead of MEASURED_PEAKS.json")
    args = ap.parse_args()
    hbm, burst, sustained = map(float, args.peaks.split(",")) if args.peaks else peaks()
    out = ["# Roofline fractions of the step's kernels", "",
           f"Denominators from `MEASURED_PEAKS.json` (driver-written on this pool): HBM copy bandwidth **{hbm:.1f} GB/s**, cuBLAS bf16 "
           f"**{burst:.1f} TFLOP/s burst** / {sustained:.1f} sustained.  Generated by `bench_tools/roofline.py` from the files named in each "
           "section; isolated kernels are judged against the burst rate, the whole step (power-capped) against the sustained one.", ""]
    for p in args.layers:
        out += layer_tables(p, hbm, burst)
    if args.comm:
        out += comm_table(args.comm)
    if args.steps:
        out += step_table(args.steps)
    sys.stdout.write("\n".join(out) + "\n")


if __name__ == "__main__":
    main()
