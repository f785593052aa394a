"""ImageNet from an image-folder tree: ``root/{train,val}/<class>/<image>``.

The reference reads ImageNet through High-Flyer's closed ffrecord reader
(``hfai.datasets.ImageNet(split, transform).loader(...)``, ref:restnet_ddp.py:107-119); outside that
cluster the same data normally lives as a directory per class.  This module gives that layout the same
call shape: ``ImageNet(split, transform, synthetic=False, root=...)`` returns an :class:`ImageFolder`
whose ``loader(batch_size, sampler, num_workers, pin_memory)`` yields ``(samples[B,3,S,S] fp32,
labels[B] int64)`` in pinned host memory.  Like the reference's loader call (``num_workers=4,
pin_memory=True``, ref:restnet_ddp.py:109) the decoding runs in ``num_workers`` worker PROCESSES: JPEG
decode releases the GIL but resize / crop / normalise on tensors does not, and at the engine's ~14 k
images/s per GPU a thread pool inside the training process cannot keep up.  The workers are persistent
(spawned once, reused every epoch) and write their samples straight into a ring of shared-memory batch
buffers that the parent has page-locked (cudaHostRegister), so a finished batch is handed to the H2D
copy without another pass over it.  ``workers="thread"`` (or PDB200_LOADER_THREADS=1) keeps the
in-process thread pool for debugging and for transforms that cannot be pickled.
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
               drop_last: bool = False, prefetch: int = 2, shuffle: bool = False, workers: Optional[str] = None,
               **_ignored):
        return FolderLoader(self, batch_size, sampler, num_workers, pin_memory, drop_last, prefetch, shuffle, workers)


def _worker_main(dataset: "ImageFolder", ring, tasks, results) -> None:
    """Body of a decode process: (slot, row0, indices, epoch) -> samples written into ring[slot] rows row0.."""
    import traceback
    torch.set_num_threads(1)
    while True:
        task = tasks.get()
        if task is None:
            return
        slot, row0, indices, epoch = task
        try:
            dataset.epoch = epoch
            x, y = ring[slot]
            for j, idx in enumerate(indices):
                img, label = dataset[idx]
                x[row0 + j].copy_(img)
                y[row0 + j] = label
            results.put((slot, len(indices), None))
        except BaseException:
            results.put((slot, 0, traceback.format_exc()))


class _ProcessPool:
    """Persistent decode processes + the shared (and, on a CUDA machine, page-locked) ring of batch buffers."""

    def __init__(self, dataset: "ImageFolder", num_workers: int, batch_size: int, sample_shape, slots: int, pin: bool):
        import torch.multiprocessing as mp
        ctx = mp.get_context("spawn")            # never fork a process that may hold a CUDA context
        self.ring = [(torch.empty((batch_size, *sample_shape), dtype=torch.float32).share_memory_(),
                      torch.empty(batch_size, dtype=torch.int64).share_memory_()) for _ in range(slots)]
        self.pinned = False
        if pin:
            try:
                rt = torch.cuda.cudart()
                for x, y in self.ring:
                    for t in (x, y):
                        err = rt.cudaHostRegister(t.data_ptr(), t.numel() * t.element_size(), 0)
                        if int(err) != 0:
                            raise RuntimeError(f"cudaHostRegister failed with {err}")
                self.pinned = True
            except Exception as e:               # still correct, only the H2D copy is staged by the driver
                import warnings
                warnings.warn(f"pdb200: could not page-lock the loader's shared batch buffers: {e}")
        self.tasks, self.results = ctx.Queue(), ctx.Queue()
        self.procs = [ctx.Process(target=_worker_main, args=(dataset, self.ring, self.tasks, self.results), daemon=True)
                      for _ in range(num_workers)]
        for p in self.procs:
            p.start()

    def close(self) -> None:
        for _ in self.procs:
            try:
                self.tasks.put(None)
            except Exception:
                pass
        for p in self.procs:
            p.join(timeout=2)
            if p.is_alive():
                p.terminate()
        if self.pinned:
            try:
                rt = torch.cuda.cudart()
                for x, y in self.ring:
                    rt.cudaHostUnregister(x.data_ptr()), rt.cudaHostUnregister(y.data_ptr())
            except Exception:
                pass
        self.procs, self.pinned = [], False


class FolderLoader:
    """Batches assembled by a pool of decode threads, ``prefetch`` batches ahead of the consumer."""

    def __init__(self, dataset: ImageFolder, batch_size: int, sampler=None, num_workers: int = 4,
                 pin_memory: bool = True, drop_last: bool = False, prefetch: int = 2, shuffle: bool = False,
                 workers: Optional[str] = None):
        if workers is None:
            workers = "thread" if (os.environ.get("PDB200_LOADER_THREADS") == "1" or int(num_workers) < 1) else "process"
        if workers not in ("process", "thread"):
            raise ValueError("workers must be 'process' or 'thread'")
        self.workers = workers
        self._pool: Optional[_ProcessPool] = None
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

    # -- worker processes ---------------------------------------------------------------------------------
    def _ensure_pool(self, first_index: int) -> _ProcessPool:
        if self._pool is None:
            shape = tuple(self.dataset[first_index][0].shape)          # one decode in the parent: the sample geometry
            self._pool = _ProcessPool(self.dataset, self.num_workers, self.batch_size, shape, self.prefetch + 3,
                                      self.pin_memory)
            import atexit
            atexit.register(self.close)
        return self._pool

    def close(self) -> None:
        if self._pool is not None:
            self._pool.close()
            self._pool = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _iter_processes(self, chunks) -> Iterator:
        """Up to `prefetch` batches are in the workers' hands while the consumer holds one: batch b lives in ring slot
        b % slots, each batch is cut into one piece per worker."""
        pool = self._ensure_pool(chunks[0][0])
        slots, epoch = len(pool.ring), self.dataset.epoch
        pending = {}                         # slot -> samples still missing
        sent = 0

        def send(b):
            c = chunks[b]
            per = -(-len(c) // self.num_workers)
            pending[b % slots] = len(c)
            for w in range(0, len(c), per):
                pool.tasks.put((b % slots, w, c[w:w + per], epoch))

        try:
            for b in range(len(chunks)):
                while sent < len(chunks) and sent <= b + self.prefetch:
                    send(sent)
                    sent += 1
                slot = b % slots
                while pending[slot] > 0:
                    try:
                        s, n, err = pool.results.get(timeout=120)
                    except queue.Empty:
                        dead = [p.pid for p in pool.procs if not p.is_alive()]
                        raise RuntimeError(f"image-folder loader: no sample from the decode workers for 120 s (dead workers: {dead})")
                    if err is not None:
                        raise RuntimeError(f"image-folder decode worker failed:\n{err}")
                    pending[s] -= n
                x, y = pool.ring[slot]
                yield x[:len(chunks[b])], y[:len(chunks[b])]
        except BaseException:
            self.close()                     # tasks of an abandoned epoch must not write into the next one's buffers
            raise
        else:
            if sent != len(chunks):
                self.close()

    def __iter__(self) -> Iterator:
        idx = self._indices()
        chunks = [idx[i:i + self.batch_size] for i in range(0, len(idx), self.batch_size)]
        if self.drop_last and chunks and len(chunks[-1]) < self.batch_size:
            chunks.pop()
        if self.workers == "process" and chunks:
            yield from self._iter_processes(chunks)
            return
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
