"""ImageNet from an image-folder tree: ``root/{train,val}/<class>/<image>``.

The reference reads ImageNet through High-Flyer's closed ffrecord reader
(``hfai.datasets.ImageNet(split, transform).loader(...)``, ref:restnet_ddp.py:107-119); outside that
cluster the same data normally lives as a directory per class.  This module gives that layout the same
call shape: ``ImageNet(split, transform, synthetic=False, root=...)`` returns an :class:`ImageFolder`
whose ``loader(batch_size, sampler, num_workers, pin_memory)`` yields ``(samples[B,3,S,S] fp32,
labels[B] int64)`` in pinned host memory, decoded by ``num_workers`` threads (PIL releases the GIL while
decoding) and transformed by the tensor transforms of ``data/transforms.py``.
"""
from __future__ import annotations

import os
import queue
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import Iterator, List, Optional, Sequence, Tuple

import torch

EXTENSIONS = (".jpg", ".jpeg", ".png", ".bmp", ".webp", ".ppm")


def _decode(path: str) -> torch.Tensor:
    """File -> uint8 CHW RGB tensor."""
    from PIL import Image
    import numpy as np
    with Image.open(path) as im:
        arr = np.asarray(im.convert("RGB"))
    return torch.from_numpy(arr.copy()).permute(2, 0, 1).contiguous()


def _takes_generator(transform) -> bool:
    from . import transforms as T
    return isinstance(transform, T.Compose)


class ImageFolder:
    def __init__(self, root: str, split: str = "train", transform=None, seed: int = 0,
                 extensions: Sequence[str] = EXTENSIONS, **_ignored):
        base = os.path.join(root, split)
        if not os.path.isdir(base):
            raise FileNotFoundError(f"{base}: expected <root>/{split}/<class>/<image files>")
        self.classes: List[str] = sorted(d for d in os.listdir(base) if os.path.isdir(os.path.join(base, d)))
        if not self.classes:
            raise FileNotFoundError(f"{base} has no class directories")
        self.class_to_idx = {c: i for i, c in enumerate(self.classes)}
        self.samples: List[Tuple[str, int]] = []
        for c in self.classes:
            d = os.path.join(base, c)
            for f in sorted(os.listdir(d)):
                if f.lower().endswith(tuple(extensions)):
                    self.samples.append((os.path.join(d, f), self.class_to_idx[c]))
        if not self.samples:
            raise FileNotFoundError(f"no image files under {base}")
        self.split, self.transform = split, transform
        self.seed = seed + (0 if split == "train" else 1)
        self.epoch = 0
        self.num_classes = len(self.classes)

    def set_epoch(self, epoch: int) -> None:
        """Different augmentation draws every epoch (the sampler's set_epoch reshuffles; this reseeds the crops)."""
        self.epoch = int(epoch)

    def __len__(self) -> int:
        return len(self.samples)

    def __getitem__(self, idx: int) -> Tuple[torch.Tensor, int]:
        path, label = self.samples[idx]
        if self.transform is not None and not _takes_generator(self.transform):
            # a torchvision-style pipeline (what a reference script passes in): it wants a PIL image
            from PIL import Image
            with Image.open(path) as im:
                out = self.transform(im.convert("RGB"))
            return (out if torch.is_tensor(out) else torch.as_tensor(out)), label
        img = _decode(path)
        if self.transform is not None:
            g = torch.Generator().manual_seed((self.seed * 1_000_003 + self.epoch) * 1_000_033 + idx)
            img = self.transform(img, g)
        return img, label

    def loader(self, batch_size: int, sampler=None, num_workers: int = 4, pin_memory: bool = True,
               drop_last: bool = False, prefetch: int = 2, shuffle: bool = False, **_ignored):
        return FolderLoader(self, batch_size, sampler, num_workers, pin_memory, drop_last, prefetch, shuffle)


class FolderLoader:
    """Batches assembled by a pool of decode threads, ``prefetch`` batches ahead of the consumer."""

    def __init__(self, dataset: ImageFolder, batch_size: int, sampler=None, num_workers: int = 4,
                 pin_memory: bool = True, drop_last: bool = False, prefetch: int = 2, shuffle: bool = False):
        self.dataset, self.batch_size, self.sampler = dataset, int(batch_size), sampler
        self.shuffle = bool(shuffle) and sampler is None
        self.num_workers, self.drop_last, self.prefetch = max(1, int(num_workers)), drop_last, max(1, int(prefetch))
        self.pin_memory = bool(pin_memory) and torch.cuda.is_available()
        self._ring = None              # queue holds <= prefetch batches, the consumer one (plus one whose H2D copy may be in flight), the producer one

    def _indices(self) -> List[int]:
        if self.sampler is not None:
            return list(iter(self.sampler))
        if self.shuffle:                         # a new permutation per epoch (dataset.set_epoch)
            g = torch.Generator().manual_seed(self.dataset.seed * 7_919 + self.dataset.epoch)
            return torch.randperm(len(self.dataset), generator=g).tolist()
        return list(range(len(self.dataset)))

    def __len__(self) -> int:
        n = len(self.sampler) if self.sampler is not None else len(self.dataset)
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)

    def _assemble(self, pool: ThreadPoolExecutor, idx: List[int], slot: int):
        items = list(pool.map(self.dataset.__getitem__, idx))
        if not self.pin_memory:
            return (torch.stack([it[0] for it in items]).to(torch.float32),
                    torch.tensor([it[1] for it in items], dtype=torch.int64))
        # a ring of pinned batch buffers, allocated once: pinning per batch costs a cudaHostAlloc every step
        shape = tuple(items[0][0].shape)
        if self._ring is None or self._ring[0][0].shape[1:] != shape:
            self._ring = [(torch.empty((self.batch_size, *shape), dtype=torch.float32).pin_memory(),
                           torch.empty(self.batch_size, dtype=torch.int64).pin_memory())
                          for _ in range(self.prefetch + 3)]
        x, y = self._ring[slot % len(self._ring)]
        x, y = x[:len(items)], y[:len(items)]
        for i, (img, label) in enumerate(items):
            x[i].copy_(img)
            y[i] = label
        return x, y

    def __iter__(self) -> Iterator:
        idx = self._indices()
        chunks = [idx[i:i + self.batch_size] for i in range(0, len(idx), self.batch_size)]
        if self.drop_last and chunks and len(chunks[-1]) < self.batch_size:
            chunks.pop()
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        stop = threading.Event()

        def produce():
            try:
                with ThreadPoolExecutor(self.num_workers, thread_name_prefix="pdb200-decode") as pool:
                    for slot, c in enumerate(chunks):
                        if stop.is_set():
                            return
                        q.put(self._assemble(pool, c, slot))
                q.put(None)
            except BaseException as e:          # surface decode errors in the consumer
                q.put(e)

        t = threading.Thread(target=produce, name="pdb200-folder-loader", daemon=True)
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
            while t.is_alive():                 # unblock a producer waiting on a full queue
                try:
                    q.get_nowait()
                except queue.Empty:
                    t.join(timeout=0.05)
