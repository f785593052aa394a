import torch
from torch.utils.data.distributed import DistributedSampler as TorchSampler

from pytorch_distributed_b200.data import DistributedSampler, SyntheticImageNet, transforms


class _DS:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


def test_sampler_matches_torch():
    for n, world in [(37, 4), (64, 8), (5, 8), (1281167, 8)]:
        for epoch in (0, 3):
            for rank in (0, world - 1):
                a = DistributedSampler(_DS(n), world, rank)
                b = TorchSampler(_DS(n), world, rank)
                a.set_epoch(epoch), b.set_epoch(epoch)
                assert len(a) == len(b)
                if n < 1000:
                    assert list(a) == list(b)
                else:
                    assert a.indices()[:100].tolist() == list(b)[:100]


def test_imagenet_step_count():
    s = DistributedSampler(_DS(1281167), 8, 0)
    assert len(s) == 160146 and -(-len(s) // 400) == 401


def test_loader_shapes_and_determinism():
    ds = SyntheticImageNet("train", num_samples=37, image_size=16, num_classes=10, pool=8)
    s = DistributedSampler(ds, 2, 1, shuffle=False)
    batches = list(ds.loader(8, sampler=s, num_workers=2))
    assert len(batches) == 3 and batches[0][0].shape == (8, 3, 16, 16) and batches[-1][0].shape[0] == 3
    assert batches[0][1].dtype == torch.int64 and batches[0][0].dtype == torch.float32
    x0, y0 = ds[int(s.indices()[0])]
    assert torch.equal(batches[0][0][0], x0) and int(batches[0][1][0]) == y0
    ring = list(ds.loader(8, mode="ring", ring=2))
    assert len(ring) == 5 and ring[0][0] is ring[2][0]


def test_transforms_pipeline():
    tr, va = transforms.train_transform(16), transforms.val_transform(16)
    ds = SyntheticImageNet("train", transform=tr, num_samples=4, image_size=16, pool=4)
    x, _ = ds[1]
    assert x.shape == (3, 16, 16) and x.dtype == torch.float32 and torch.equal(x, ds[1][0])
    dv = SyntheticImageNet("val", transform=va, num_samples=4, image_size=16, pool=4)
    assert dv[0][0].shape == (3, 16, 16)
    n = transforms.Normalize()(torch.full((3, 2, 2), 0.5))
    assert torch.allclose(n[:, 0, 0], (0.5 - torch.tensor(transforms.IMAGENET_MEAN)) / torch.tensor(transforms.IMAGENET_STD))
