"""Top-level functions for spawned children (must be importable by name)."""
import os
import time

import torch


def tiny_cfg(mode, **kw):
    from pytorch_distributed_b200.config import TrainConfig
    base = dict(mode=mode, epochs=1, batch_size=4, max_steps=3, max_val_steps=1, image_size=32, num_classes=10,
                train_samples=64, val_samples=16, width=8, layers=(1, 1, 1, 1), device="cpu", num_workers=0)
    base.update(kw)
    return TrainConfig(**base)


def child_ok(rank, out_dir):
    with open(os.path.join(out_dir, f"ok{rank}"), "w") as f:
        f.write(str(rank))


def child_raises(rank):
    if rank == 1:
        raise ValueError("boom from rank 1")
    time.sleep(30)


def child_killed(rank):
    if rank == 0:
        os.kill(os.getpid(), 9)
    time.sleep(30)


def child_suspends(rank):
    from pytorch_distributed_b200 import suspend
    if rank == 0:
        suspend.go_suspend()
    time.sleep(30)


def avg_hook(state, bucket):
    """A user comm hook in the delegate's calling convention: averages the bucket with an async all-reduce."""
    import torch.distributed as dist
    state["calls"] += 1
    t = bucket.buffer()
    # the view a hook gets: parameters of this bucket, their gradients aliasing the bucket's address range
    ps, gs = bucket.parameters(), bucket.gradients()
    assert len(ps) == len(gs) > 0 and sum(g.numel() for g in gs) <= t.numel()
    lo, hi = t.data_ptr(), t.data_ptr() + 4 * t.numel()
    assert all(lo <= g.data_ptr() < hi for g in gs)
    assert isinstance(bucket.index(), int) and isinstance(bucket.is_last(), bool)
    t.div_(dist.get_world_size())
    return dist.all_reduce(t, async_op=True).get_future().then(lambda f: f.value()[0])


def ddp_worker(rank, world, port, out_dir, use_ours, steps, sync_bn=False, hook=None):
    import torch.distributed as dist
    from pytorch_distributed_b200.models import ResNet
    from pytorch_distributed_b200.optim import SGD
    from pytorch_distributed_b200.parallel import DistributedDataParallel, convert_sync_batchnorm
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", world_size=world, rank=rank)
    torch.manual_seed(1234 + rank)          # different init per rank: the ctor broadcast must fix it
    model = ResNet((1, 1, 1, 1), num_classes=10, width=8)
    if sync_bn:
        model = convert_sync_batchnorm(model) if use_ours else torch.nn.SyncBatchNorm.convert_sync_batchnorm(model)
    if use_ours:
        ddp = DistributedDataParallel(model)
        hook_state = {"calls": 0}
        if hook == "py":
            ddp.register_comm_hook(hook_state, avg_hook)
        opt = SGD(ddp.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    else:
        ddp = torch.nn.parallel.DistributedDataParallel(model)
        opt = torch.optim.SGD(ddp.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(99)
    xs = torch.randn(steps, world, 4, 3, 32, 32, generator=g)
    ys = torch.randint(0, 10, (steps, world, 4), generator=g)
    losses = []
    for s in range(steps):
        out = ddp(xs[s, rank])
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(out, ys[s, rank])
        loss.backward()
        opt.step()
        losses.append(loss.item())
    flat = torch.cat([p.detach().reshape(-1) for p in ddp.module.parameters()])
    bufs = torch.cat([b.detach().float().reshape(-1) for b in ddp.module.buffers()])
    if use_ours and hook == "py":
        assert hook_state["calls"] == steps * len(ddp.bucket_ranges), hook_state
    tag = ("ours" if use_ours else "torch") + (f"_{hook}" if hook else "")
    torch.save({"losses": losses, "params": flat.clone(), "bufs": bufs.clone()}, os.path.join(out_dir, f"{tag}_{rank}.pt"))
    dist.destroy_process_group()


def harness_worker(rank, cfg):
    from pytorch_distributed_b200 import harness
    torch.set_num_threads(1)
    harness.main(cfg, rank)


def reference_style_main(local_rank, port, out_dir):
    """The body of the reference's DDP ``main(local_rank)`` (ref:restnet_ddp.py:75-150) written against the hfai shim
    only — torchvision transforms, hfai datasets / DDP / distributed / client — on CPU + gloo, tiny sizes."""
    import hf_env
    hf_env.set_env("202111")
    import hfai
    import hfai.nccl.distributed as dist
    from hfai.nn.parallel import DistributedDataParallel
    from torch import nn
    from torchvision import transforms as tv
    from pytorch_distributed_b200.models import ResNet
    torch.set_num_threads(1)
    world_size, rank = 2, local_rank
    dist.init_process_group(backend="gloo", init_method=f"tcp://127.0.0.1:{port}", world_size=world_size, rank=rank)
    model = DistributedDataParallel(ResNet((1, 1, 1, 1), num_classes=10, width=8), device_ids=None)
    tr = tv.Compose([tv.RandomResizedCrop(16), tv.RandomHorizontalFlip(), tv.ToTensor(),
                     tv.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    train_dataset = hfai.datasets.ImageNet("train", transform=tr, num_samples=32, image_size=16, num_classes=10)
    sampler = torch.utils.data.distributed.DistributedSampler(train_dataset, num_replicas=world_size, rank=rank)
    loader = train_dataset.loader(batch_size=4, sampler=sampler, num_workers=2, pin_memory=True)
    criterion = nn.CrossEntropyLoss()
    optimizer = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    scheduler = torch.optim.lr_scheduler.StepLR(optimizer, step_size=30, gamma=0.1)
    sampler.set_epoch(0)
    model.train()
    steps = 0
    for step, (samples, labels) in enumerate(loader):
        outputs = model(samples)
        optimizer.zero_grad()
        loss = criterion(outputs, labels)
        loss.backward()
        optimizer.step()
        steps += 1
        assert not hfai.client.receive_suspend_command()
    scheduler.step()
    acc = torch.tensor([float(loss), 1.0, 2.0, float(steps)])
    dist.reduce(acc, 0)
    if rank == 0 and local_rank == 0:
        torch.save({"model": model.module.state_dict(), "optimizer": optimizer.state_dict(),
                    "scheduler": scheduler.state_dict(), "acc": acc.tolist()}, os.path.join(out_dir, "latest.pt"))
    dist.destroy_process_group()
