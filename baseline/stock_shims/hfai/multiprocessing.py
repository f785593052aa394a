import torch.multiprocessing as _mp


def spawn(fn, args=(), nprocs=1, bind_numa=False, **kw):
    return _mp.spawn(fn, args=args, nprocs=nprocs, **kw)
