"""Batch loaders.

The reference wraps ``DataLoader`` in ``prefetch_generator.BackgroundGenerator`` and uses
``pin_memory=True, persistent_workers=True, drop_last=True`` + ``.to(device, non_blocking=True)``
(``resnet50_test.py:41-43, 321-352``).  Equivalents here:

* :class:`BackgroundLoader` — a producer thread + bounded queue around any iterable (the
  ``DataLoaderX`` replacement; works with ``torch.utils.data.DataLoader`` and worker processes).
* :class:`TensorBatchLoader` — for datasets that fit in host memory as tensors (CIFAR-10 is
  150 MB of uint8): a batch is one ``index_select`` into a **pinned** staging buffer, so no worker
  processes, no per-sample Python, and the H2D copy is a single async DMA.  Augmentation and
  normalisation happen on the GPU (``csrc/input_head.cu``), so the loader ships uint8.
* :class:`DevicePrefetcher` — double-buffers the H2D copies on a side stream so step N+1's inputs
  land while step N computes.
"""
from __future__ import annotations

import queue
import threading
from typing import Iterable, Iterator, List, Optional, Sequence, Tuple

import torch

from ..parallel.sampler import ShardedSampler


class BackgroundLoader:
    def __init__(self, iterable: Iterable, max_prefetch: int = 2):
        self.iterable, self.max_prefetch = iterable, max_prefetch

    def __len__(self):
        return len(self.iterable)

    def __iter__(self) -> Iterator:
        q: "queue.Queue" = queue.Queue(self.max_prefetch)
        sentinel = object()

        def produce():
            try:
                for item in self.iterable:
                    q.put(item)
            except BaseException as e:  # surface worker errors in the consumer
                q.put(e)
            q.put(sentinel)

        t = threading.Thread(target=produce, daemon=True)
        t.start()
        while True:
            item = q.get()
            if item is sentinel:
                break
            if isinstance(item, BaseException):
                raise item
            yield item
        t.join()


class _StagedBatch(tuple):
    """A batch living in a pinned staging slot; :meth:`mark_copied` (called by whoever enqueued the asynchronous
    host→device copy) lets the loader know when the slot may be refilled."""

    def __new__(cls, tensors, loader, slot):
        self = super().__new__(cls, tensors)
        self._loader, self._slot = loader, slot
        return self

    def mark_copied(self, stream=None):
        if torch.cuda.is_available():
            ev = torch.cuda.Event()
            ev.record(stream if stream is not None else torch.cuda.current_stream())
            self._loader._slot_events[self._slot] = ev


class TensorBatchLoader:
    def __init__(self, tensors: Sequence[torch.Tensor], batch_size: int, sampler: Optional[ShardedSampler] = None,
                 shuffle: bool = True, drop_last: bool = True, pin: Optional[bool] = None, ring: int = 8, seed: int = 0):
        self.tensors = [t.contiguous() for t in tensors]
        n = self.tensors[0].shape[0]
        self.sampler = sampler or ShardedSampler(n, 1, 0, shuffle=shuffle, seed=seed)
        self.batch_size, self.drop_last = batch_size, drop_last
        self.pin = torch.cuda.is_available() if pin is None else pin
        # a staging buffer is rewritten `ring` batches later; BackgroundLoader (2 queued + 1 in hand) and
        # DevicePrefetcher (1 ahead + 1 current) keep at most 5 batches alive, so 8 is safe
        self.ring = max(ring, 8)
        self._stage: List[List[torch.Tensor]] = []
        # one CUDA event per staging slot: DevicePrefetcher records it after the slot's host→device copy was enqueued,
        # and the slot is not refilled before that copy has finished (ordering, not timing)
        self._slot_events: List[Optional["torch.cuda.Event"]] = [None] * self.ring
        for _ in range(self.ring):
            bufs = []
            for t in self.tensors:
                b = torch.empty((batch_size,) + tuple(t.shape[1:]), dtype=t.dtype)
                bufs.append(b.pin_memory() if self.pin else b)
            self._stage.append(bufs)

    def set_epoch(self, epoch: int):
        self.sampler.set_epoch(epoch)

    def __len__(self):
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        idx = self.sampler.indices()
        nb = len(self)
        for b in range(nb):
            sel = idx[b * self.batch_size:(b + 1) * self.batch_size]
            slot = b % self.ring
            bufs = self._stage[slot]
            ev = self._slot_events[slot]
            if ev is not None:
                ev.synchronize()              # the previous H2D copy out of this pinned slot is complete
                self._slot_events[slot] = None
            out = []
            for t, buf in zip(self.tensors, bufs):
                if sel.numel() == self.batch_size:
                    torch.index_select(t, 0, sel, out=buf)
                    out.append(buf)
                else:
                    out.append(t.index_select(0, sel))
            yield _StagedBatch(out, self, slot) if self.pin else tuple(out)

    def bytes_per_batch(self) -> int:
        return sum(b.numel() * b.element_size() for b in self._stage[0])


class DevicePrefetcher:
    """Iterates ``loader`` and returns batches already on ``device`` (copied on a side stream)."""

    def __init__(self, loader: Iterable, device, transform=None):
        self.loader, self.device, self.transform = loader, torch.device(device), transform
        self.stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None

    def __len__(self):
        return len(self.loader)

    def _to_device(self, batch):
        if self.stream is None:
            out = tuple(b.to(self.device) for b in batch)
            return self.transform(out) if self.transform else out
        with torch.cuda.stream(self.stream):
            out = tuple(b.to(self.device, non_blocking=True) for b in batch)
            if hasattr(batch, "mark_copied"):
                batch.mark_copied(self.stream)
            if self.transform:
                out = self.transform(out)
        return out

    def __iter__(self):
        it = iter(self.loader)
        try:
            nxt = self._to_device(next(it))
        except StopIteration:
            return
        while nxt is not None:
            if self.stream is not None:
                torch.cuda.current_stream(self.device).wait_stream(self.stream)
                for t in nxt:
                    if torch.is_tensor(t):
                        t.record_stream(torch.cuda.current_stream(self.device))
            cur = nxt
            try:
                nxt = self._to_device(next(it))
            except StopIteration:
                nxt = None
            yield cur
