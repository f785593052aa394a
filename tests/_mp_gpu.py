"""Spawned workers for the multi-GPU tests (one process per GPU, NCCL bootstrap on 127.0.0.1)."""
import json
import os

import torch


def _init(rank, world, port):
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", world_size=world, rank=rank,
                            device_id=torch.device("cuda", rank))
    return dist


BUCKET_ELEMS = [2049024, 7875584, 6563840, 6637568, 2431040]      # the five ResNet-50 DDP buckets


def comm_worker(rank, world, port, out_dir):
    dist = _init(rank, world, port)
    import torch.distributed._symmetric_memory as symm_mem
    from pytorch_distributed_b200.parallel.symm import NativeComm
    dev = torch.device("cuda", rank)
    res = {}
    comm = NativeComm()
    n_total = sum(BUCKET_ELEMS)
    buf = symm_mem.empty(n_total, dtype=torch.float32, device=dev)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    inv = torch.tensor(1.0 / 512.0, device=dev)
    for use_mc in ([True, False] if comm._map(buf).multicast else [False]):
        comm.use_multicast = use_mc
        buf.copy_(torch.randn(n_total, device=dev, generator=g) * 512)
        ref = buf.clone()
        dist.all_reduce(ref)
        ref.mul_(1.0 / world / 512.0)
        found = torch.zeros((), device=dev)
        off = 0
        for n in BUCKET_ELEMS:
            comm.allreduce_range(buf, off, off + n, average=True, inv_scale=inv, found_inf=found)
            off += n
        comm.wait_all()
        torch.cuda.synchronize()
        err = float((buf - ref).abs().max() / ref.abs().max())
        res[f"allreduce_err_mc{int(use_mc)}"] = err
        res[f"found_inf_clean_mc{int(use_mc)}"] = float(found)
        # inf injection on ONE rank must raise the flag on ALL ranks
        buf.normal_()
        if rank == world - 1:
            buf[12345] = float("inf")
        found.zero_()
        comm.allreduce_range(buf, 0, BUCKET_ELEMS[0], average=True, inv_scale=inv, found_inf=found)
        comm.wait_all()
        torch.cuda.synchronize()
        res[f"found_inf_injected_mc{int(use_mc)}"] = float(found)
    res["multicast_available"] = bool(comm._map(buf).multicast)
    # 16-bit wire (bf16_compress / fp16_compress semantics)
    for wd, tag in ((torch.bfloat16, "bf16"), (torch.float16, "fp16")):
        buf.copy_(torch.randn(n_total, device=dev, generator=g))
        ref = buf.clone()
        dist.all_reduce(ref)
        ref.mul_(1.0 / world)
        n = BUCKET_ELEMS[0]
        comm.allreduce_range(buf, 0, n, average=True, wire_dtype=wd)
        comm.wait_all()
        torch.cuda.synchronize()
        res[f"wire_{tag}_err"] = float((buf[:n] - ref[:n]).abs().max() / ref[:n].abs().max())
    # K6: statistics kernel with the cross-GPU reduction inside
    xs = (torch.randn(3000, 256, device=dev, generator=g) + rank).to(torch.bfloat16)
    sums = torch.empty(512, device=dev)
    comm.bn_stats_sync(xs, sums)
    comm.bn_stats_sync(xs, sums)                 # twice: slots and tickets are reusable
    want = torch.cat([xs.float().sum(0), (xs.float() ** 2).sum(0)])
    dist.all_reduce(want)
    torch.cuda.synchronize()
    res["bn_stats_sync_err"] = float((sums - want).abs().max() / want.abs().max())
    # broadcast
    buf.fill_(float(rank + 1))
    comm.broadcast_(buf, 0)
    torch.cuda.synchronize()
    res["broadcast_ok"] = bool((buf == 1.0).all())
    # small all-reduce, repeated (slot reuse)
    ok = True
    for it in range(20):
        v = torch.arange(516, device=dev, dtype=torch.float32) * (rank + 1) + it
        comm.allreduce_small_(v)
        want = torch.arange(516, device=dev, dtype=torch.float32) * (world * (world + 1) / 2) + it * world
        ok &= bool(torch.allclose(v, want))
    torch.cuda.synchronize()
    res["small_allreduce_ok"] = ok
    # soak: many back-to-back all-reduces shake out signal-pad reuse bugs
    buf.fill_(1.0)
    for _ in range(200):
        comm.allreduce_range(buf, 0, 4096, average=True)
    comm.wait_all()
    torch.cuda.synchronize()
    res["soak_ok"] = bool(torch.allclose(buf[:4096], torch.ones(4096, device=dev)))
    with open(os.path.join(out_dir, f"comm_{rank}.json"), "w") as f:
        json.dump(res, f)
    dist.barrier()
    dist.destroy_process_group()


def ddp_native_worker(rank, world, port, out_dir, backend, sync_bn, precise=False):
    """precise: the `ddp` mode as the reference runs it (fp32, no loss scaling) on the DETERMINISTIC engine — the two
    comm backends then differ only in the association of the cross-rank sum, so the runs can be held to 1e-2."""
    dist = _init(rank, world, port)
    from pytorch_distributed_b200 import amp, harness
    from pytorch_distributed_b200.config import TrainConfig
    from pytorch_distributed_b200.nn import CrossEntropyLoss
    from pytorch_distributed_b200.optim import SGD
    from pytorch_distributed_b200.parallel import DistributedDataParallel
    dev = torch.device("cuda", rank)
    cfg = TrainConfig(mode="ddp" if precise else "ddp_amp", batch_size=16, dtype="fp32" if precise else "bf16", device="cuda",
                      seed=7 + rank, sync_bn=sync_bn)
    model = harness.build_model(cfg, dev)            # different init per rank: the ctor broadcast must fix it
    model.enable_native(batch_size=16, dtype=harness.torch_dtype(cfg), sync_bn=sync_bn, deterministic=precise)
    ddp = DistributedDataParallel(model, device_ids=[rank], backend=backend)
    opt = SGD(ddp.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    sc = None
    if not precise:
        sc = amp.GradScaler(device=dev)
        ddp.attach_grad_scaler(sc)
    crit = CrossEntropyLoss()
    g = torch.Generator(device=dev).manual_seed(5)          # same stream on all ranks; each takes its slice
    steps = 5              # two eager steps, then the captured forward / backward graphs are replayed
    xs = torch.randn(steps, world, 16, 3, 64, 64, device=dev, generator=g)
    ys = torch.randint(0, 1000, (steps, world, 16), device=dev, generator=g)
    ddp.train()
    losses = []
    if hasattr(ddp.comm, "bucket_times"):
        ddp.comm.profile = True                 # event pair around every bucket all-reduce
    for s in range(steps):
        losses.append(float(harness.train_step(ddp, crit, opt, xs[s, rank], ys[s, rank], cfg, dev, sc, False)))
    torch.cuda.synchronize()
    arena = model._pdb200_arena
    buckets = ddp.comm.bucket_times() if hasattr(ddp.comm, "bucket_times") else None
    torch.save({"losses": losses, "params": arena.params.detach().cpu().clone(), "comm": ddp.comm.name, "buckets": buckets,
                "bufs": model._pdb200_buffer_arena.floats.detach().cpu().clone(), "scale": sc.get_scale() if sc else None},
               os.path.join(out_dir, f"ddp_{backend}_{int(sync_bn)}{'_precise' if precise else ''}_{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()
