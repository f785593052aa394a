import pytest
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


def _make_image_tree(root, classes=3, per_class=5, size=(40, 52)):
    import numpy as np
    from PIL import Image
    rng = np.random.RandomState(0)
    for split in ("train", "val"):
        for c in range(classes):
            d = root / split / f"class_{c}"
            d.mkdir(parents=True)
            for i in range(per_class):
                arr = rng.randint(0, 256, size=(size[0], size[1], 3), dtype=np.uint8)
                arr[..., 0] = c * 80                                  # the class is visible in the red channel
                Image.fromarray(arr).save(d / f"img_{i}.png")


def test_image_folder_dataset_and_loader(tmp_path):
    """Real-data path: <root>/{train,val}/<class>/<image> decoded by worker threads, reference transforms, the
    hfai-style ``.loader(...)`` call, sharded by our DistributedSampler."""
    from pytorch_distributed_b200.data import DistributedSampler, ImageNet
    _make_image_tree(tmp_path)
    tr = ImageNet("train", transform=transforms.train_transform(16), synthetic=False, root=str(tmp_path))
    va = ImageNet("val", transform=transforms.val_transform(16), synthetic=False, root=str(tmp_path))
    assert len(tr) == 15 and tr.num_classes == 3 and tr.classes == ["class_0", "class_1", "class_2"]
    x, y = va[7]
    assert x.shape == (3, 16, 16) and x.dtype == torch.float32 and y == 1
    # 8-bit pixels end up normalised: red channel of class 1 is 80/255 before Normalize
    want = (80 / 255 - transforms.IMAGENET_MEAN[0]) / transforms.IMAGENET_STD[0]
    assert abs(float(x[0].mean()) - want) < 0.02
    # same epoch -> same augmentation; next epoch -> different crop
    a0, _ = tr[3]
    assert torch.equal(a0, tr[3][0])
    tr.set_epoch(1)
    assert not torch.equal(a0, tr[3][0])
    # loader: all samples once, shapes, worker threads, sampler sharding
    batches = list(va.loader(batch_size=4, num_workers=3, pin_memory=False))
    assert [b[0].shape[0] for b in batches] == [4, 4, 4, 3] and batches[0][1].dtype == torch.int64
    assert sorted(int(v) for b in batches for v in b[1]) == sorted([0] * 5 + [1] * 5 + [2] * 5)
    shards = [DistributedSampler(tr, num_replicas=2, rank=r, shuffle=True, seed=0) for r in range(2)]
    seen = [int(v) for s in shards for b in tr.loader(batch_size=4, sampler=s, num_workers=2, pin_memory=False) for v in b[1]]
    assert len(seen) == 16                                              # ceil(15/2) per rank, padded like torch
    p1 = [int(v) for b in tr.loader(batch_size=15, shuffle=True, pin_memory=False) for v in b[1]]
    assert p1 != sorted(p1) and sorted(p1) == sorted([0] * 5 + [1] * 5 + [2] * 5)


def test_scripts_train_on_an_image_folder(tmp_path):
    """--data-root switches the harness from synthetic tensors to the folder reader end to end."""
    import subprocess, sys, os
    root = tmp_path / "data"
    _make_image_tree(root, classes=2, per_class=6)
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(here, "resnet_single_gpu.py"), "--device", "cpu", "--epochs", "1",
                        "--batch-size", "4", "--image-size", "32", "--num-classes", "2", "--width", "8", "--workers", "2",
                        "--data-root", str(root), "--save-path", str(tmp_path / "out")],
                       capture_output=True, text=True, timeout=600, cwd=here, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Acc1:" in r.stdout and (tmp_path / "out" / "best.pt").exists()


def test_image_folder_accepts_torchvision_transforms(tmp_path):
    """A reference-style script hands hfai.datasets.ImageNet a torchvision (PIL) pipeline (ref:restnet_ddp.py:100-107)."""
    tv = pytest.importorskip("torchvision.transforms")
    from pytorch_distributed_b200.data import ImageNet
    _make_image_tree(tmp_path, classes=2, per_class=2)
    t = tv.Compose([tv.Resize(20), tv.CenterCrop(16), tv.ToTensor(),
                    tv.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    ds = ImageNet("val", transform=t, synthetic=False, root=str(tmp_path))
    x, y = ds[3]
    assert x.shape == (3, 16, 16) and x.dtype == torch.float32 and y == 1
    xb, yb = next(iter(ds.loader(batch_size=4, num_workers=2, pin_memory=False)))
    assert xb.shape == (4, 3, 16, 16) and yb.tolist() == [0, 0, 1, 1]


def test_hfai_shim_reads_real_data_when_a_root_is_configured(tmp_path, monkeypatch):
    import hfai
    _make_image_tree(tmp_path, classes=2, per_class=2)
    assert type(hfai.datasets.ImageNet("val", num_samples=4, image_size=8)).__name__ == "SyntheticImageNet"
    monkeypatch.setenv("PDB200_DATA_ROOT", str(tmp_path))
    ds = hfai.datasets.ImageNet("val", transform=transforms.val_transform(8))
    assert type(ds).__name__ == "ImageFolder" and len(ds) == 4


def test_folder_loader_worker_processes_match_worker_threads(tmp_path):
    """num_workers decode PROCESSES (the default, like the reference's DataLoader workers) write into shared batch
    buffers; they must deliver exactly what the in-process thread pool delivers, epoch after epoch."""
    from pytorch_distributed_b200.data import ImageNet
    _make_image_tree(tmp_path)
    tr = ImageNet("train", transform=transforms.train_transform(16), synthetic=False, root=str(tmp_path))
    lp = tr.loader(batch_size=4, num_workers=2, pin_memory=False, shuffle=True)
    lt = tr.loader(batch_size=4, num_workers=2, pin_memory=False, shuffle=True, workers="thread")
    assert lp.workers == "process" and lt.workers == "thread"
    try:
        for epoch in range(2):
            tr.set_epoch(epoch)
            got = [(x.clone(), y.clone()) for x, y in lp]            # batches are views of a ring: copy before the next one
            want = list(lt)
            assert len(got) == len(want) == 4
            for (x, y), (xr, yr) in zip(got, want):
                assert torch.equal(y, yr) and torch.equal(x, xr)
        assert lp._pool is not None and all(p.is_alive() for p in lp._pool.procs)      # persistent across epochs
    finally:
        lp.close()
    assert lp._pool is None


def test_synthetic_loader_recycles_sample_buffers_but_not_labels():
    """Default mode for transform-free synthetic data: after the first ring of batches the pinned sample buffers are
    re-used as they are (no 240 MB host gather per step), while every batch still carries the labels of ITS indices."""
    from pytorch_distributed_b200.data.dataset import _label_of
    ds = SyntheticImageNet("train", num_samples=8 * 12, image_size=8, num_classes=10, pool=8)
    loader = ds.loader(8, num_workers=2, prefetch=1)          # ring = prefetch + 3 = 4 slots
    assert loader.mode == "recycle"
    seen = []
    for b, (x, y) in enumerate(loader):
        idx = torch.arange(b * 8, (b + 1) * 8)
        assert torch.equal(y, _label_of(idx, 10) % 10), b
        if b < 4:
            assert torch.equal(x[0], ds[b * 8][0])            # the first ring is assembled exactly
        seen.append(x.data_ptr())
    assert len(seen) == 12 and seen[4] == seen[0] and seen[9] == seen[1]
    exact = list(ds.loader(8, num_workers=2, mode="gather"))
    assert len(exact) == 12
