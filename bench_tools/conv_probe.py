"""GPU probe: runs every conv/GEMM configuration in its own subprocess (a trap or a hang in one
tcgen05/TMA configuration must not poison the others) and prints a table of relative errors.

  python bench_tools/conv_probe.py            # all cases
  python bench_tools/conv_probe.py one <json> # internal: run one case
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [
    # kind, N, H, W, C, K, R, S, stride, pad, block_n
    ("fprop", 2, 16, 16, 64, 64, 1, 1, 1, 0, 64),       # plain GEMM, M=512
    ("fprop", 2, 16, 16, 64, 128, 1, 1, 1, 0, 128),
    ("fprop", 2, 16, 16, 128, 256, 1, 1, 1, 0, 256),
    ("fprop", 3, 14, 14, 256, 1024, 1, 1, 1, 0, 256),   # M = 588: tail tile
    ("fprop", 2, 56, 56, 64, 64, 3, 3, 1, 1, 64),       # im2col 3x3
    ("fprop", 2, 28, 28, 128, 128, 3, 3, 1, 1, 128),
    ("fprop", 3, 14, 14, 256, 256, 3, 3, 1, 1, 256),
    ("fprop", 2, 56, 56, 128, 128, 3, 3, 2, 1, 128),    # stride 2
    ("fprop", 2, 56, 56, 256, 512, 1, 1, 2, 0, 256),    # 1x1 stride 2 (downsample)
    ("fprop", 5, 7, 7, 512, 512, 3, 3, 1, 1, 256),      # 7x7 maps: tiles span images
    ("fprop", 400, 1, 1, 2048, 1000, 1, 1, 1, 0, 128),  # fc
    ("fprop", 2, 112, 112, 152, 64, 1, 1, 1, 0, 64),    # stem as GEMM over the im2col buffer (K tail)
    ("fprop", 3, 14, 14, 256, 1024, 1, 1, 1, 0, 512),   # 512 = 256-wide tile on a CTA pair (cta_group::2); 5 M blocks (odd)
    ("fprop", 3, 14, 14, 256, 256, 3, 3, 1, 1, 512),
    ("fprop", 5, 7, 7, 512, 512, 3, 3, 1, 1, 512),
    ("fprop", 2, 56, 56, 256, 512, 1, 1, 2, 0, 512),
    ("fprop", 16, 14, 14, 1024, 256, 1, 1, 1, 0, 512),
    ("dgrad", 2, 28, 28, 128, 128, 3, 3, 1, 1, 0),
    ("dgrad", 2, 14, 14, 256, 1024, 1, 1, 1, 0, 0),
    ("wgrad", 2, 16, 16, 64, 64, 1, 1, 1, 0, 64),
    ("wgrad", 2, 16, 16, 256, 128, 1, 1, 1, 0, 256),
    ("wgrad", 2, 28, 28, 64, 64, 3, 3, 1, 1, 256),
    ("wgrad", 2, 28, 28, 128, 256, 3, 3, 1, 1, 256),
    ("wgrad", 2, 56, 56, 128, 128, 3, 3, 2, 1, 256),
    ("stem", 4, 64, 64, 3, 64, 7, 7, 2, 3, 0),          # space-to-depth stem: fprop + statistics + wgrad
    ("stem", 3, 224, 224, 3, 64, 7, 7, 2, 3, 0),
    ("wgrad", 2, 56, 56, 256, 512, 1, 1, 2, 0, 256),
    ("wgrad", 400, 1, 1, 2048, 1000, 1, 1, 1, 0, 256),  # fc
    ("wgrad", 3, 7, 7, 512, 512, 3, 3, 1, 1, 256),
]


def run_one(case):
    import torch  # noqa: F401
    from tests._gpu_helpers import conv_case, dgrad_case, stem_case, wgrad_case
    kind, N, H, W, C, K, R, S, stride, pad, bn = case
    if kind == "fprop":
        e = conv_case(N, H, W, C, K, R, S, stride, pad, block_n=bn)
        return {"y": e[0], "stats": e[1]}
    if kind == "stem":
        e = stem_case(N, H, W)
        return {"y": e[0], "stats": e[1], "dw": e[2]}
    if kind == "dgrad":
        return {"dx": dgrad_case(N, H, W, C, K, R, S, pad)}
    return {"dw": wgrad_case(N, H, W, C, K, R, S, stride, pad, block_n=bn)}


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "one":
        case = json.loads(sys.argv[2])
        print("RESULT " + json.dumps(run_one(case)))
        return 0
    only = sys.argv[1] if len(sys.argv) > 1 else None
    bad = 0
    for case in CASES:
        if only == "pair":
            if case[-1] != 512:
                continue
        elif only and case[0] != only:
            continue
        try:
            r = subprocess.run([sys.executable, __file__, "one", json.dumps(case)], capture_output=True, text=True,
                               timeout=180, cwd=ROOT)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            if r.returncode == 0 and line:
                res = json.loads(line[0][7:])
                ok = all(v == v and v < 2e-2 for v in res.values())
                bad += 0 if ok else 1
                print(("ok  " if ok else "BAD ") + f"{case} -> {res}", flush=True)
            else:
                bad += 1
                tail = (r.stderr or r.stdout).strip().splitlines()[-3:]
                print(f"FAIL {case} rc={r.returncode} :: {' | '.join(tail)}", flush=True)
        except subprocess.TimeoutExpired:
            bad += 1
            print(f"HANG {case} (killed after 180 s)", flush=True)
    print(f"conv_probe: {len(CASES) - bad}/{len(CASES)} ok", flush=True)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
