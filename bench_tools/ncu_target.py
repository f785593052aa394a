"""Short single-GPU program for ncu captures: runs one kernel family a few times on a real ResNet-50 shape
at batch 400.   python bench_tools/ncu_target.py {fprop|wgrad|bn|step|fprop_tf32|wgrad_tf32|comm}

`comm`: the cross-GPU kernels launched with world = 1 (one buffer, one signal pad: the meetings are trivially met, the
data path is the real one) — ncu replays every kernel ~40 times, which a kernel that waits for a peer cannot survive,
but a loopback launch can; it also puts allreduce_kernel & co. into a single-GPU launch census."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from pytorch_distributed_b200 import ops  # noqa: E402
from pytorch_distributed_b200.ops import ConvShape  # noqa: E402


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "fprop"
    N, dt = 400, torch.bfloat16
    if what == "fprop1x1":
        H, C, K = 14, 256, 1024
        x = torch.randn(N, H, H, C, device="cuda", dtype=dt)
        w = torch.randn(K, 1, 1, C, device="cuda", dtype=dt) * 0.05
        y = torch.empty(N, H, H, K, device="cuda", dtype=dt)
        st = torch.zeros(2 * K, device="cuda")
        bn = int(sys.argv[2]) if len(sys.argv) > 2 else 0
        for _ in range(6):
            ops.conv_fprop(x, w, y, ConvShape(N, H, H, C, K, 1, 1, 1, 0), stats=st, block_n=bn)
    elif what == "stem":
        x = torch.randn(N, 3, 224, 224, device="cuda")
        w16 = (torch.randn(64, 7, 7, 3, device="cuda") * 0.08).to(dt)
        xs = torch.empty(N, 112, 112, 16, device="cuda", dtype=dt)
        ws = torch.empty(64, 256, device="cuda", dtype=dt)
        y = torch.empty(N, 112, 112, 64, device="cuda", dtype=dt)
        dy = torch.randn(N, 112, 112, 64, device="cuda").to(dt)
        st, scr = torch.zeros(128, device="cuda"), torch.zeros(16 * 128 + 32, device="cuda")
        dws = torch.zeros(64, 256, device="cuda")
        ops.stem_s2d(x, xs)
        ops.stem_weight_s2d(w16, ws)
        for _ in range(4):
            ops.stem_fprop(xs, ws, y, stats=st, stats_scratch=scr)
            ops.stem_wgrad(xs, dy, dws)
    elif what == "comm":
        import ctypes
        from pytorch_distributed_b200.ops._lib import check, require
        L = require()
        n = 7875584                                   # the largest ResNet-50 DDP bucket (31.5 MB)
        buf = torch.randn(n, device="cuda")
        pad = torch.zeros(64 * 16, dtype=torch.int32, device="cuda")
        flag = torch.zeros(64, device="cuda")
        stage = torch.zeros(2 * 8192, device="cuda")
        arr = lambda t: (ctypes.c_void_p * 1)(t.data_ptr())        # noqa: E731
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        inv = torch.tensor(1.0 / 1024, device="cuda")
        small = torch.randn(4096, device="cuda")
        for _ in range(3):
            check(L.pdb_allreduce_inplace(arr(buf), arr(pad), arr(flag), None, 0, 1, 0, n, 1.0, ctypes.c_void_p(inv.data_ptr()), 32, st), "allreduce")
            check(L.pdb_allreduce_oneshot(arr(buf), arr(pad), arr(flag), 0, 1, 0, 65536, 1.0, ctypes.c_void_p(inv.data_ptr()), st), "oneshot")
            check(L.pdb_broadcast_inplace(arr(buf), arr(pad), 0, 1, 0, 0, 53120, 4, st), "broadcast")
            check(L.pdb_small_allreduce(arr(stage), arr(pad), 0, 1, ctypes.c_void_p(small.data_ptr()), 4096, 8192, st), "small")
    elif what in ("fprop", "wgrad", "fprop_tf32", "wgrad_tf32"):
        if what.endswith("_tf32"):
            what, dt = what[:-5], torch.float32
        H, C, K, R = 14, 256, 256, 3
        x = torch.randn(N, H, H, C, device="cuda", dtype=dt)
        sh = ConvShape(N, H, H, C, K, R, R, 1, 1)
        if what == "fprop":
            w = torch.randn(K, R, R, C, device="cuda", dtype=dt) * 0.02
            y = torch.empty(N, H, H, K, device="cuda", dtype=dt)
            st = torch.zeros(2 * K, device="cuda")
            for _ in range(6):
                ops.conv_fprop(x, w, y, sh, stats=st)
        else:
            dy = torch.randn(N, H, H, K, device="cuda", dtype=dt)
            dw = torch.zeros(K, R, R, C, device="cuda")
            for _ in range(6):
                ops.conv_wgrad(x, dy, dw, sh)
    elif what == "bn":
        rows, C = N * 56 * 56, 256
        x = torch.randn(rows, C, device="cuda", dtype=dt)
        z = torch.relu(x)
        dz = torch.randn(rows, C, device="cuda", dtype=dt)
        mean, rstd = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
        gamma = torch.ones(C, device="cuda")
        ds = torch.zeros(2 * C, device="cuda")
        dx, dres = torch.empty_like(x), torch.empty_like(x)
        sums = torch.zeros(2 * C, device="cuda")
        y = torch.empty_like(x)
        for _ in range(4):
            ops.bn_bwd_reduce(dz, z, x, mean, rstd, ds)
            ops.bn_bwd_apply(dz, z, x, dx, dres, mean, rstd, gamma, ds, rows, None, None)
            ops.bn_apply(x, z, y, sums + 1.0, rows, gamma, mean, 1e-5, 0.1, None, None, None, None, None, True)
    else:
        from pytorch_distributed_b200 import amp, harness
        from pytorch_distributed_b200.config import TrainConfig
        from pytorch_distributed_b200.nn import CrossEntropyLoss
        from pytorch_distributed_b200.optim import SGD
        dev = torch.device("cuda", 0)
        name = {"step_fp16": "fp16", "step_fp32": "fp32"}.get(what, "bf16")          # step | step_fp16 | step_fp32
        cfg = TrainConfig(mode="single" if name == "fp32" else "ddp_amp", batch_size=N, dtype=name, device="cuda")
        model = harness.build_model(cfg, dev)
        os.environ["PDB200_GRAPHS"] = "0"            # eager launches: every kernel shows up by name
        model.enable_native(batch_size=N, dtype=harness.torch_dtype(cfg))
        opt = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
        sc = amp.GradScaler(device=dev)
        crit = CrossEntropyLoss()
        xb = torch.randn(N, 3, 224, 224, device=dev)
        yb = torch.randint(0, 1000, (N,), device=dev)
        model.train()
        for _ in range(3):
            harness.train_step(model, crit, opt, xb, yb, cfg, dev, sc, stock_autocast=False)
    torch.cuda.synchronize()
    print("ok", what)


if __name__ == "__main__":
    main()
