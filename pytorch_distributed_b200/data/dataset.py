"""ImageNet-shaped dataset and batch loader.

The reference reads ImageNet through ``hfai.datasets.ImageNet(split, transform)
.loader(batch_size, sampler, num_workers, pin_memory)`` (ref:restnet_ddp.py:107-119), a
closed ffrecord reader that is not on this box and needs a cluster file system.  This
module keeps that call shape and yields the same thing — ``(samples[B,3,224,224] fp32,
labels[B] int64)`` in pinned host memory — from a synthetic source.

Two assembly strategies:

* ``gather``: every sample is a pure function of its dataset index (pool image
  ``idx % pool`` + hashed label), so sharding / shuffling / resume tests are exact.
* ``ring``: a ring of fully formed pinned batches that is cycled; the per-step cost
  is exactly the host→device copy, which is what the throughput benchmark wants.
"""
from __future__ import annotations

import math
import os
import queue
import threading
from typing import Iterator, Optional, Tuple

import torch

from .sampler import DistributedSampler  # noqa: F401  (re-export convenience)


def _label_of(idx: torch.Tensor, num_classes: int) -> torch.Tensor:
    # Knuth multiplicative hash: deterministic, well spread, cheap.
    return ((idx.to(torch.int64) * 2654435761) % 4294967296) % num_classes


class SyntheticImageNet:
    """ImageNet-1k-shaped synthetic dataset (1,281,167 train / 50,000 val by default)."""

    def __init__(self, split: str = "train", transform=None, num_samples: Optional[int] = None,
                 image_size: int = 224, num_classes: int = 1000, seed: int = 0, pool: int = 32):
        if split not in ("train", "val"):
            raise ValueError(f"unknown split {split!r}")
        self.split = split
        self.transform = transform
        self.image_size = image_size
        self.num_classes = num_classes
        self.seed = seed + (0 if split == "train" else 1)
        if num_samples is None:
            num_samples = 1_281_167 if split == "train" else 50_000
        self.num_samples = int(num_samples)
        self.pool = int(min(pool, self.num_samples))
        self._images: Optional[torch.Tensor] = None

    # The pool is built lazily so constructing a dataset stays free (len() is all the
    # sampler needs) and so spawned workers do not pickle hundreds of MB.
    def _pool_images(self) -> torch.Tensor:
        if self._images is None:
            g = torch.Generator().manual_seed(self.seed)
            if self.transform is None:
                self._images = torch.randn(self.pool, 3, self.image_size, self.image_size, generator=g)
            else:
                raw = int(self.image_size * 256 / 224)
                self._images = torch.randint(0, 256, (self.pool, 3, raw, raw), generator=g,
                                             dtype=torch.uint8)
        return self._images

    def __len__(self) -> int:
        return self.num_samples

    def __getitem__(self, idx: int) -> Tuple[torch.Tensor, int]:
        if not 0 <= idx < self.num_samples:
            raise IndexError(idx)
        img = self._pool_images()[idx % self.pool]
        if self.transform is not None:
            from . import transforms as T
            if isinstance(self.transform, T.Compose):
                g = torch.Generator().manual_seed(self.seed * 1_000_003 + idx)
                img = self.transform(img, g)
            else:
                # a torchvision-style (PIL) pipeline handed in by a reference-style script (ref:restnet_ddp.py:100-107)
                from PIL import Image
                img = self.transform(Image.fromarray(img.permute(1, 2, 0).numpy()))
        label = int(_label_of(torch.tensor(idx), self.num_classes))
        return img, label

    def loader(self, batch_size: int, sampler=None, num_workers: int = 4, pin_memory: bool = True,
               mode: str = "auto", drop_last: bool = False, ring: int = 4, prefetch: int = 2):
        """mode: "gather" assembles every batch from its indices (exact per-index pixels, ~240 MB of host copies per
        400-image batch); "recycle" does that for the first few batches only and then re-uses those pinned sample buffers
        with index-exact LABELS — synthetic pixels carry no information, and eight ranks gathering 1.9 GB per step
        between them starve a host with few cores (measured: 38-94 ms steps on 8 GPUs instead of 28.5); "ring" is a fixed
        ring of whole batches.  "auto" = recycle for transform-free synthetic data, gather otherwise."""
        if mode == "auto":
            mode = "recycle" if self.transform is None else "gather"
        return BatchLoader(self, batch_size, sampler=sampler, num_workers=num_workers,
                           pin_memory=pin_memory, mode=mode, drop_last=drop_last, ring=ring,
                           prefetch=prefetch)


def ImageNet(split: str = "train", transform=None, synthetic: bool = True, root: Optional[str] = None, **kw):
    """``hfai.datasets.ImageNet`` call shape (ref:restnet_ddp.py:107,117).  ``synthetic=False`` reads an image-folder
    tree ``root/{train,val}/<class>/<image>`` (data/folder.py); the reference's own ffrecord files need the
    High-Flyer cluster file system and are not readable here."""
    if not synthetic:
        root = root or os.environ.get("PDB200_DATA_ROOT")
        if not root:
            raise RuntimeError("real data needs --data-root (or PDB200_DATA_ROOT) pointing at <root>/{train,val}/<class>/"
                               "<image>; the reference's ffrecord reader (hfai.datasets) is closed")
        from .folder import ImageFolder
        return ImageFolder(root, split=split, transform=transform, seed=kw.get("seed", 0))
    return SyntheticImageNet(split=split, transform=transform, **kw)


class BatchLoader:
    """Iterable of pinned ``(samples, labels)`` batches assembled off the training thread."""

    def __init__(self, dataset: SyntheticImageNet, batch_size: int, sampler=None, num_workers: int = 4,
                 pin_memory: bool = True, mode: str = "gather", drop_last: bool = False,
                 ring: int = 4, prefetch: int = 2):
        if mode not in ("gather", "ring", "recycle"):
            raise ValueError(mode)
        self.dataset, self.batch_size, self.sampler = dataset, int(batch_size), sampler
        self.num_workers, self.mode, self.drop_last = num_workers, mode, drop_last
        self.pin_memory = bool(pin_memory) and torch.cuda.is_available()
        self.ring_size, self.prefetch = max(1, ring), max(1, prefetch)
        self._ring = None

    def _num_indices(self) -> int:
        return len(self.sampler) if self.sampler is not None else len(self.dataset)

    def __len__(self) -> int:
        n = self._num_indices()
        return n // self.batch_size if self.drop_last else math.ceil(n / self.batch_size)

    def _index_tensor(self) -> torch.Tensor:
        if self.sampler is None:
            return torch.arange(len(self.dataset))
        if hasattr(self.sampler, "indices"):
            return self.sampler.indices()
        return torch.tensor(list(iter(self.sampler)), dtype=torch.int64)

    def _alloc(self, b: int):
        s = self.dataset.image_size
        x = torch.empty(b, 3, s, s, dtype=torch.float32, pin_memory=self.pin_memory)
        y = torch.empty(b, dtype=torch.int64, pin_memory=self.pin_memory)
        return x, y

    def _assemble(self, idx: torch.Tensor, out=None, pool=None):
        ds = self.dataset
        x, y = out if out is not None else self._alloc(idx.numel())
        x, y = x[: idx.numel()], y[: idx.numel()]
        if ds.transform is None:
            src, sel = ds._pool_images(), idx % ds.pool
            if pool is None or idx.numel() < 64:
                torch.index_select(src, 0, sel, out=x)
            else:
                # a 400-image batch is 240 MB: one thread copies it in ~25 ms, the engine consumes one every 28 ms —
                # split the gather over the worker threads (index_select releases the GIL)
                n, w = idx.numel(), self.num_workers
                per = -(-n // w)
                list(pool.map(lambda a: torch.index_select(src, 0, sel[a:a + per], out=x[a:a + per]), range(0, n, per)))
        else:
            for k, i in enumerate(idx.tolist()):
                x[k] = ds[i][0]
        torch.remainder(_label_of(idx, ds.num_classes), ds.num_classes, out=y)
        return x, y

    def _iter_ring(self) -> Iterator:
        if getattr(self, "_fixed_ring", None) is None:
            base = torch.arange(self.batch_size)
            self._fixed_ring = [self._assemble(base + k * self.batch_size) for k in range(self.ring_size)]
        for k in range(len(self)):
            yield self._fixed_ring[k % self.ring_size]

    def _iter_gather(self) -> Iterator:
        index = self._index_tensor()
        nb = len(self)
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        stop = threading.Event()

        def work():
            # Pinned batch buffers are allocated ONCE and recycled: a cudaHostAlloc of 240 MB per step costs more than
            # the step.  The queue holds <= prefetch batches, the consumer one, one more may still be in its H2D copy,
            # and the producer fills one: prefetch + 3 slots.
            from concurrent.futures import ThreadPoolExecutor
            try:
                if self._ring is None or len(self._ring) < self.prefetch + 3 or self._ring[0][0].shape[0] != self.batch_size:
                    self._ring = [self._alloc(self.batch_size) for _ in range(self.prefetch + 3)]
                    self._ring_filled = 0
                with ThreadPoolExecutor(max(1, self.num_workers), thread_name_prefix="pdb200-gather") as pool:
                    for b in range(nb):
                        if stop.is_set():
                            return
                        idx = index[b * self.batch_size: (b + 1) * self.batch_size]
                        slot = self._ring[b % len(self._ring)]
                        if self.mode == "recycle" and b >= len(self._ring) and idx.numel() == self.batch_size \
                                and getattr(self, "_ring_filled", 0) >= len(self._ring):
                            # sample buffer of this slot is already filled (and full-sized): labels only
                            y = slot[1][: idx.numel()]
                            torch.remainder(_label_of(idx, self.dataset.num_classes), self.dataset.num_classes, out=y)
                            q.put((slot[0][: idx.numel()], y))
                            continue
                        q.put(self._assemble(idx, out=slot, pool=pool))
                        if idx.numel() == self.batch_size:
                            self._ring_filled = getattr(self, "_ring_filled", 0) + 1
                q.put(None)
            except BaseException as e:  # surface worker errors in the consumer
                q.put(e)

        if self.num_workers <= 0:
            for b in range(nb):     # (no worker thread: plain per-batch assembly)
                yield self._assemble(index[b * self.batch_size: (b + 1) * self.batch_size])
            return
        t = threading.Thread(target=work, daemon=True, name="pdb200-loader")
        t.start()
        try:
            while True:
                item = q.get()
                if item is None:
                    return
                if isinstance(item, BaseException):
                    raise item
                yield item
        finally:
            stop.set()
            while not q.empty():
                try:
                    q.get_nowait()
                except queue.Empty:
                    break

    def __iter__(self) -> Iterator:
        return self._iter_ring() if self.mode == "ring" else self._iter_gather()
