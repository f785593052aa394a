#!/usr/bin/env python
"""Headline benchmark: ResNet-50 training images/sec, batch 400 per GPU, synthetic ImageNet.

Contract (driver): ``python bench.py --gpus N --steps K --warmup W`` for N=1, or the same under
``python -m torch.distributed.run --nproc-per-node N`` for N>1 (one rank per GPU, NCCL bootstrap).
Rank 0 prints ONE JSON line.  ``--impl reference`` runs the UNMODIFIED reference training loop
(baseline/_ref/*.py ``train()``) on stock torch/torchvision/cuDNN/NCCL for the same config.

Metric/config come from BASELINE.json: "ResNet-50 train images/sec (device-timed, max over
ranks)", batch 400/GPU, SGD(0.1, 0.9, 1e-4); the flagship mode is the reference's fastest one,
DDP + mixed precision ("Apex", ref:resnet_ddp_apex.py).  ``--mode ddp|single`` selects the others.

Timing: W>=3 warm-up steps, then exactly K steps between CUDA events on the compute stream with a
barrier + synchronize on both sides, max over ranks.  The per-step input (240.8 MB fp32) and the
activations (GBs) are far larger than the 126 MB L2, so no explicit L2 flush is needed.
``value`` times the step with the batch already on the device; ``e2e`` goes through the public
harness loop: every step copies its batch from pinned host memory and reads the loss back.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

# Derived from the reference's chart: train img/s per GPU.
PUBLISHED_PER_GPU = {"ddp_amp": 693.3, "ddp": 669.0, "single": 717.0, "dp": 162.7}


# ------------------------------------------------------------------------------------ clocks
class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,"
             "utilization.gpu,memory.used")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="pdb200_clocks_", suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power, util, mem = [], [], set(), [], [], []
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])), mx.append(float(f[2])), power.append(float(f[3]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
                try:        # the other two panels of the reference's chart (README.md:33-34): utilisation and memory
                    util.append(float(f[9])), mem.append(float(f[10]))
                except (ValueError, IndexError):
                    pass
            os.remove(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm), power_w_max=max(power))
        if util:
            out.update(gpu_util_pct_mean=round(statistics.mean(util), 1), mem_used_gb_max=round(max(mem) / 1024, 2))
        return out


# ------------------------------------------------------------------------------------ common
def dist_setup(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.mode == "dp":                      # single process drives every GPU (ref:resnet_dp.py)
        torch.cuda.set_device(0)
        args.gpus = min(args.gpus, torch.cuda.device_count())
        return 1, 0, 0
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N>1 must be launched with python -m torch.distributed.run --nproc-per-node N")
        args.gpus = world
    torch.cuda.set_device(local)
    if args.mode != "single" or world > 1:
        if "MASTER_ADDR" not in os.environ:
            os.environ["MASTER_ADDR"] = "127.0.0.1"
            os.environ.setdefault("MASTER_PORT", str(29400 + os.getpid() % 500))
        dist.init_process_group("nccl", init_method="env://", world_size=world, rank=rank,
                                device_id=torch.device("cuda", local))
    return world, rank, local


def timed_region(fn, steps, device):
    """barrier + sync, events around exactly `steps` calls, sync; returns max-over-ranks ms."""
    import torch
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize(device)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        fn(i)
    b.record()
    torch.cuda.synchronize(device)
    ms = a.elapsed_time(b)
    if dist.is_initialized():
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        ms = float(t.item())
    return ms


def synthetic_device_batch(batch, device, seed):
    import torch
    g = torch.Generator(device=device).manual_seed(seed)
    x = torch.randn(batch, 3, 224, 224, device=device, generator=g)
    y = torch.randint(0, 1000, (batch,), device=device, generator=g)
    return x, y


class PinnedRing:
    """Ring of pinned host batches: the per-step cost is exactly the H2D copy."""

    def __init__(self, batch, n, ring=3, seed=0):
        import torch
        g = torch.Generator().manual_seed(seed)
        self.items = []
        for _ in range(ring):
            x = torch.empty(batch, 3, 224, 224, pin_memory=True)
            x.normal_(generator=g)
            y = torch.randint(0, 1000, (batch,), generator=g).pin_memory()
            self.items.append((x, y))
        self.n = n

    def __len__(self):
        return self.n

    def __iter__(self):
        for i in range(self.n):
            yield self.items[i % len(self.items)]

    def bytes_per_step(self):
        x, y = self.items[0]
        return x.numel() * x.element_size() + y.numel() * y.element_size()


def emit(args, world, ms, e2e, clocks, launches, extra_cfg, impl, peak_mem_gb=None):
    gb = args.batch * world
    value = gb * args.steps / (ms / 1e3)
    pub = PUBLISHED_PER_GPU[args.mode] * world
    rec = {
        "metric": "ResNet-50 train images/sec (device-timed, max over ranks)",
        "value": round(value, 2), "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": round(value / pub, 3), "dtype": args.dtype,
        "data": "synthetic ImageNet-shaped (random 3x224x224 fp32, 1000 classes), random-init weights",
        "impl": impl,
        "config": {"model": "resnet50", "global_batch": gb, "per_gpu_batch": args.batch, "image": 224,
                   "mode": args.mode, "parallelism": f"dp{world}", "optimizer": "SGD(0.1,0.9,1e-4)",
                   "l2": "per-step inputs (240.8 MB) and activations exceed the 126 MB L2; no flush needed",
                   "published_baseline_img_s": pub, **extra_cfg},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        # the reference reports three panels per mode (result.png): epoch time, average GPU utilisation, average
        # GPU memory — the last two as sampled by nvidia-smi on GPU 0 during the timed regions, plus torch's peak
        "gpu_util_pct": clocks.get("gpu_util_pct_mean") if clocks else None,
        "mem_used_gb": clocks.get("mem_used_gb_max") if clocks else None,
        "peak_mem_gb": peak_mem_gb,
    }
    print(json.dumps(rec), flush=True)


# ------------------------------------------------------------------------------------ ours
def run_ours(args):
    import torch
    world, rank, local = dist_setup(args)
    device = torch.device("cuda", local)
    from pytorch_distributed_b200 import amp, harness, ops
    from pytorch_distributed_b200.config import TrainConfig
    from pytorch_distributed_b200.nn import CrossEntropyLoss
    from pytorch_distributed_b200.optim import SGD, StepLR
    from pytorch_distributed_b200.parallel import DistributedDataParallel

    cfg = TrainConfig(mode=args.mode, batch_size=args.batch, dtype=args.dtype or "auto", engine=args.engine,
                      backend=args.backend, sync_bn=args.sync_bn, device="cuda", seed=rank)
    model = harness.build_model(cfg, device)
    native = harness._want_native(cfg, device)
    if native:
        model.enable_native(batch_size=args.batch, dtype=harness.torch_dtype(cfg), sync_bn=args.sync_bn)
    elif args.engine != "torch":
        raise SystemExit("bench.py: the native sm_100a engine is not available on this machine (libpdb200.so missing "
                         "or not loadable); pass --engine torch to time the stock-op path on purpose")
    args.dtype = cfg.resolved_dtype() + ("(tf32)" if cfg.resolved_dtype() == "fp32" and native else "")
    ndp = args.gpus if args.mode == "dp" else 1
    if args.mode in ("ddp", "ddp_amp"):
        wrapped = DistributedDataParallel(model, device_ids=[local], backend=args.backend)
    elif args.mode == "dp":
        from pytorch_distributed_b200.parallel import DataParallel
        wrapped = DataParallel(model, device_ids=list(range(ndp)), output_device=0)
    else:
        harness.ensure_arenas(model, device)
        wrapped = model
    criterion = CrossEntropyLoss()
    optimizer = SGD(wrapped.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    scheduler = StepLR(optimizer, 30, 0.1)
    scaler = None
    if args.mode == "ddp_amp":
        scaler = amp.GradScaler(enabled=True, device=device)
        wrapped.attach_grad_scaler(scaler)
    stock_autocast = not native
    wrapped.train()
    comm_name = getattr(getattr(wrapped, "comm", None), "name", None)
    if world > 1 and args.mode in ("ddp", "ddp_amp") and args.backend != "nccl" and comm_name != "native":
        raise SystemExit(f"bench.py: {world} ranks but the gradient path is {comm_name!r}, not the native peer-memory "
                         "kernels; pass --backend nccl to time the library path on purpose")

    x, y = synthetic_device_batch(args.batch * ndp, device, 1000 + rank)

    def step(_i):
        harness.train_step(wrapped, criterion, optimizer, x, y, cfg, device, scaler, stock_autocast)

    for i in range(args.warmup):
        step(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ops.launch_count() if ops.available() else 0
    torch.cuda.reset_peak_memory_stats(device)
    ms = timed_region(step, args.steps, device)
    launches = (ops.launch_count() - l0) if ops.available() else 0
    peak_mem_gb = round(torch.cuda.max_memory_allocated(device) / 2**30, 2)

    # ---- end to end through the public harness loop: pinned host batch -> H2D, loss -> D2H, each step
    ring = PinnedRing(args.batch * ndp, args.warmup + args.steps, seed=rank)
    losses = []

    class _E2E:
        def __init__(self):
            self.a = torch.cuda.Event(enable_timing=True)
            self.b = torch.cuda.Event(enable_timing=True)
            self.n = 0

        def on_step(self, s, loss):
            losses.append(float(loss.detach()))          # device -> host read of the step's result
            self.n += 1
            if self.n == args.warmup:
                import torch.distributed as dist
                if dist.is_initialized():
                    dist.barrier()
                torch.cuda.synchronize(device)
                self.a.record()

    e = _E2E()
    cfg.max_steps = args.warmup + args.steps
    harness.train(ring, wrapped, criterion, optimizer, scheduler, 0, 0, 0.0, Path(tempfile.gettempdir()), cfg,
                  device, local, scaler, None, on_step=e.on_step)
    e.b.record()
    torch.cuda.synchronize(device)
    e2e_ms = e.a.elapsed_time(e.b)
    import torch.distributed as dist
    if dist.is_initialized():
        t = torch.tensor([e2e_ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    clocks = sampler.stop() if rank == 0 else {}
    world = world * ndp                        # GPUs doing the work (DP mode: one process, ndp devices)
    e2e = {"value": round(args.batch * world * args.steps / (e2e_ms / 1e3), 2), "unit": "images/s",
           "ms_per_step": round(e2e_ms / args.steps, 3), "h2d_bytes_per_step": ring.bytes_per_step(),
           "d2h_bytes_per_step": 4, "api": "pytorch_distributed_b200.harness.train", "last_loss": losses[-1]}
    if rank == 0:
        emit(args, world, ms, e2e, clocks, launches,
             {"engine": "native-sm100a" if native else "torch-ops", "comm": comm_name,
              "sync_bn": args.sync_bn}, "ours", peak_mem_gb)
    if dist.is_initialized():
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------ reference
def run_reference(args):
    ref_dir = ROOT / "baseline" / "_ref"
    script = {"ddp_amp": "resnet_ddp_apex", "ddp": "restnet_ddp", "single": "resnet_single_gpu", "dp": "resnet_dp"}[args.mode]
    if not (ref_dir / f"{script}.py").exists():
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref missing: the reference has no "
                          "setup.py/pyproject (pip install fails); run python baseline/install_ref.py"}))
        return
    # stock shims FIRST so `import hfai` inside the reference resolves to stock torch, not to our package
    sys.path.insert(0, str(ROOT / "baseline" / "stock_shims"))
    sys.path.insert(0, str(ref_dir))
    for m in [k for k in sys.modules if k == "hfai" or k.startswith("hfai.") or k == "hf_env"]:
        del sys.modules[m]
    import torch
    cvd = os.environ.get("CUDA_VISIBLE_DEVICES")
    world, rank, local = dist_setup(args)
    device = torch.device("cuda", local)
    import importlib
    ref = importlib.import_module(script)                # unmodified reference module
    if cvd is None:
        os.environ.pop("CUDA_VISIBLE_DEVICES", None)     # resnet_single_gpu.py sets it at import (quirk Q13)
    else:
        os.environ["CUDA_VISIBLE_DEVICES"] = cvd
    from torch import nn
    from torch.optim import SGD
    from torch.optim.lr_scheduler import StepLR
    from torchvision import models

    torch.manual_seed(rank)
    model = models.resnet50().cuda()
    ndp = args.gpus if args.mode == "dp" else 1
    if args.mode in ("ddp", "ddp_amp"):
        kw = {} if args.mode == "ddp_amp" else {"device_ids": [local]}      # as in the two scripts
        model = torch.nn.parallel.DistributedDataParallel(model, **kw)
    elif args.mode == "dp":
        model = nn.DataParallel(model, device_ids=list(range(ndp)), output_device=0)     # ref:resnet_dp.py:82
    criterion = nn.CrossEntropyLoss()
    optimizer = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    scheduler = StepLR(optimizer, step_size=30, gamma=0.1)
    scaler = torch.cuda.amp.GradScaler(enabled=True) if args.mode == "ddp_amp" else None
    save = Path(tempfile.mkdtemp(prefix="pdb200_ref_"))

    def call_train(loader):
        if args.mode == "ddp_amp":
            ref.train(loader, model, criterion, optimizer, scheduler, scaler, 0, local, 0, 0.0, save)
        elif args.mode == "ddp":
            ref.train(loader, model, criterion, optimizer, scheduler, 0, local, 0, 0.0, save)
        elif args.mode == "dp":
            ref.train(loader, model, criterion, optimizer, scheduler, 0, 0, 0.0, save)
        else:
            import contextlib
            import io
            with contextlib.redirect_stdout(io.StringIO()):          # its step ticker
                ref.train(loader, model, criterion, optimizer, scheduler, 0, 0, 0.0, save)

    x, y = synthetic_device_batch(args.batch * ndp, device, 1000 + rank)
    call_train([(x, y)] * args.warmup)
    # Extra warm-up for the REFERENCE only (it helps the reference): cuDNN autotuning and the allocator growth that
    # follows every empty_cache() leaked into round 1's timed region (its device-timed value came out slower than its
    # own end-to-end number).  Keep stepping until two consecutive steps agree within 5 %, at most 12 more.
    prev, extra_warm = None, 0
    while extra_warm < 12:
        t = timed_region(lambda i: call_train([(x, y)]), 1, device)
        extra_warm += 1
        if prev is not None and abs(t - prev) <= 0.05 * prev:
            break
        prev = t
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    torch.cuda.reset_peak_memory_stats(device)
    ms = timed_region(lambda i: call_train([(x, y)]), args.steps, device)
    peak_mem_gb = round(torch.cuda.max_memory_allocated(device) / 2**30, 2)
    ring = PinnedRing(args.batch * ndp, args.steps, seed=rank)
    call_train(list(ring)[:2])
    world = world * ndp
    e2e_ms = timed_region(lambda i: call_train([ring.items[i % len(ring.items)]]), args.steps, device)
    clocks = sampler.stop() if rank == 0 else {}
    e2e = {"value": round(args.batch * world * args.steps / (e2e_ms / 1e3), 2), "unit": "images/s",
           "ms_per_step": round(e2e_ms / args.steps, 3), "h2d_bytes_per_step": ring.bytes_per_step(),
           "d2h_bytes_per_step": 0, "api": f"baseline/_ref/{script}.py::train (unmodified)"}
    if rank == 0:
        # same labels as the other arm uses for the same precision; the detail of how the reference gets there is in config
        args.dtype = "fp16" if args.mode == "ddp_amp" else "fp32(tf32)"
        detail = ("torch.cuda.amp.autocast() fp16 + GradScaler (ref:resnet_ddp_apex.py:27,107)" if args.mode == "ddp_amp" else
                  "fp32 tensors, cuDNN convolutions in TF32 (torch default), fc matmul in fp32")
        emit(args, world, ms, e2e, clocks, None,
             {"engine": "stock torch + cuDNN + NCCL", "precision_detail": detail,
              "cudnn_allow_tf32": torch.backends.cudnn.allow_tf32,
              "per_step_empty_cache": args.mode != "single", "extra_warmup_steps": extra_warm}, "reference", peak_mem_gb)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", choices=("ours", "reference"), default="ours")
    p.add_argument("--mode", choices=("ddp_amp", "ddp", "single", "dp"), default="ddp_amp")
    p.add_argument("--batch", type=int, default=400)
    p.add_argument("--dtype", choices=("bf16", "fp16", "fp32"), default=None,
                   help="default: the reference's precision for the mode — fp16 for ddp_amp, fp32 (TF32 GEMMs) otherwise")
    p.add_argument("--engine", choices=("auto", "native", "torch"), default="auto")
    p.add_argument("--backend", choices=("auto", "native", "nccl"), default="auto")
    p.add_argument("--sync-bn", action="store_true")
    args = p.parse_args()
    args.warmup = max(3, args.warmup)
    import torch
    if not torch.cuda.is_available():
        # the metric is device-timed on B200s; without a GPU there is nothing to measure — say so in one line
        if int(os.environ.get("RANK", "0")) == 0:
            if args.impl == "reference":
                print(json.dumps({"impl": "reference", "unavailable": "no CUDA device on this machine"}))
            else:
                print(json.dumps({"impl": "ours", "error": "no CUDA device on this machine: bench.py times the training "
                                  "step on B200 GPUs (run it through gpurun)"}))
        sys.exit(0 if args.impl == "reference" else 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
