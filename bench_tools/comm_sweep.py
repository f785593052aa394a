"""All-reduce bus bandwidth: our peer-memory kernel (P2P two-shot and NVLS multimem) vs NCCL, on the five
ResNet-50 DDP bucket sizes and a 4 KB .. 128 MB sweep.  Launch with torchrun (one rank per GPU):

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
      bench_tools/comm_sweep.py --out gpurun_out/comm_sweep.json

Device-timed with CUDA events, max over ranks.  busbw = 2 (N-1)/N * bytes / time (NCCL convention).
Roofline: NVLink 5 at the measured 770 GB/s per direction per GPU (B200_PROFILING.md).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

BUCKETS = [2049024, 7875584, 6563840, 6637568, 2431040]


def timed(fn, iters, dev):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([a.elapsed_time(b) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/comm_sweep.json")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm_mem
    from pytorch_distributed_b200.parallel.symm import NativeComm
    comm = NativeComm()
    nmax = 32 << 20
    buf = symm_mem.empty(nmax, dtype=torch.float32, device=dev)
    buf.zero_()
    plain = torch.zeros(nmax, device=dev)
    has_mc = bool(comm._map(buf).multicast)
    sizes = sorted(set(BUCKETS + [1 << k for k in range(10, 26)]))
    rows = []
    for n in sizes:
        iters = 50 if n < (1 << 20) else 20
        rec = {"elems": n, "bytes": n * 4, "bucket": n in BUCKETS}

        def run_native():
            comm.allreduce_range(buf, 0, n, average=True)
            comm.wait_all()
        one_shot_default = 4 * n <= comm.oneshot_bytes
        if 4 * n <= (1 << 20):            # the one-shot kernel (the default up to 256 KB) against the two-shot ones
            keep = comm.oneshot_bytes
            comm.oneshot_bytes = 1 << 20
            ms = timed(run_native, iters, dev)
            comm.oneshot_bytes = keep
            rec["oneshot_us"] = round(ms * 1e3, 2)
            rec["oneshot_busbw_GBs"] = round(2 * (world - 1) / world * n * 4 / (ms * 1e-3) / 1e9, 1)
        rec["default_algo"] = "oneshot" if one_shot_default else ("nvls" if has_mc else "p2p")
        keep = comm.oneshot_bytes
        comm.oneshot_bytes = 0            # force the two-shot kernels for the p2p / nvls columns
        for mode in (["p2p", "nvls"] if has_mc else ["p2p"]):
            comm.use_multicast = mode == "nvls"
            ms = timed(run_native, iters, dev)
            rec[f"{mode}_us"] = round(ms * 1e3, 2)
            rec[f"{mode}_busbw_GBs"] = round(2 * (world - 1) / world * n * 4 / (ms * 1e-3) / 1e9, 1)
        comm.oneshot_bytes = keep
        # device time of the kernel alone (events on the comm stream around the launch): the columns above also pay for
        # the two cross-stream hand-offs (main -> comm stream -> main) of allreduce_range + wait_all, which in a
        # training step are off the critical path
        comm.profile = True
        comm.bucket_times()
        for _ in range(10):
            run_native()
        torch.cuda.synchronize(dev)
        ks = [b["us"] for b in comm.bucket_times()][2:]
        comm.profile = False
        t = torch.tensor([sum(ks) / max(len(ks), 1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rec["default_kernel_us"] = round(float(t), 2)
        rec["default_kernel_busbw_GBs"] = round(2 * (world - 1) / world * n * 4 / (float(t) * 1e-6) / 1e9, 1)
        ms = timed(lambda: dist.all_reduce(plain[:n]), iters, dev)
        rec["nccl_us"] = round(ms * 1e3, 2)
        rec["nccl_busbw_GBs"] = round(2 * (world - 1) / world * n * 4 / (ms * 1e-3) / 1e9, 1)
        best = max(rec.get("p2p_busbw_GBs", 0), rec.get("nvls_busbw_GBs", 0), rec.get("oneshot_busbw_GBs", 0),
                   rec.get("default_kernel_busbw_GBs", 0))
        rec["roofline_frac_of_770GBs"] = round(best / 770.0, 3)
        rows.append(rec)
        if rank == 0:
            print(rec, flush=True)
    if rank == 0:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        json.dump({"world": world, "multicast": has_mc, "rows": rows}, open(args.out, "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
