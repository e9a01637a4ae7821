"""Batch loader over an LMDB image database -- the data side of the hot path as one iterator.

Replaces what ``LoaderConfig.make`` assembles in the reference (``torchbooster/config.py:348-379``:
``data_sampler`` + ``DataLoader(batch_size, sampler, num_workers, pin_memory, drop_last)`` whose
workers run the per-sample transform): here a background thread gathers the raw records of the
next batches straight out of the memory map into pinned buffers (C++, GIL released), and the
consumer thread hands each pinned batch to ``BatchTransform``, whose kernel reads it in place.

    reader  = LMDBReader(path)
    sampler = data_sampler(range(len(reader)), shuffle=True, distributed=True)   # reference's sharding
    loader  = BatchLoader(reader, (256, 256, 3), batch_size=512, sampler=sampler,
                          transform=BatchTransform(spec, seed=42), labels=reader.read_labels())
    for x, y in loader: ...
"""
from __future__ import annotations

import queue
import threading
from typing import Iterable, Iterator, Optional, Sequence, Tuple

import numpy as np
import torch

from .lmdb import LMDBReader


class BatchLoader:
    """Yields ``(batch, labels)`` per step.

    ``batch`` is the transform's output on the GPU (or the pinned ``uint8 [B,H,W,C]`` batch itself
    when ``transform`` is None); ``labels`` is an int64 tensor on the transform's device (host when
    there is no transform), or ``None`` without a label array.  ``sampler`` is any iterable of
    record indices (default: ``range(len(reader))``); ``drop_last`` as in ``DataLoader``.
    A yielded batch stays valid until ``prefetch`` further batches have been requested.
    """

    def __init__(self, reader: LMDBReader, shape: Tuple[int, int, int], batch_size: int,
                 sampler: Optional[Iterable[int]] = None, transform=None, labels=None, drop_last: bool = False,
                 prefetch: int = 2, threads: int = 8, in_place: bool = False, ragged: bool = False) -> None:
        if batch_size < 1 or prefetch < 1:
            raise ValueError("batch_size and prefetch must be >= 1")
        self.reader, self.shape, self.batch_size = reader, tuple(int(s) for s in shape), int(batch_size)
        self.sampler, self.transform, self.drop_last = sampler, transform, drop_last
        self.prefetch, self.threads = int(prefetch), int(threads)
        self.record_bytes = int(np.prod(self.shape))
        self.labels = None if labels is None else torch.as_tensor(np.asarray(labels), dtype=torch.int64)
        self._pin = torch.cuda.is_available()
        # ragged: records of different sizes (write_ragged_image_lmdb); `shape` is ignored, the transform
        # must be a resize (+ center_crop) or random_resized_crop pipeline
        self.ragged = bool(ragged)
        if self.ragged and (transform is None or in_place):
            raise ValueError("ragged needs a transform and excludes in_place")
        # in_place: the database map is registered with CUDA and the kernel reads the records where
        # they lie -- no gather at all; the background thread only looks the record offsets up.
        self._base = None
        if in_place:
            if transform is None:
                raise ValueError("in_place needs a transform")
            self._base = reader.cuda_register()

    def _indices(self) -> Sequence[int]:
        return list(self.sampler) if self.sampler is not None else list(range(len(self.reader)))

    def __len__(self) -> int:
        n = len(self.sampler) if self.sampler is not None else len(self.reader)
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, Optional[torch.Tensor]]]:
        indices = np.asarray(self._indices(), dtype=np.int64)
        n = indices.shape[0]
        nb = n // self.batch_size if self.drop_last else -(-n // self.batch_size)
        nslots = self.prefetch + 2
        slots = [] if self._base is not None else [
            torch.empty((self.batch_size, self.record_bytes), dtype=torch.uint8, pin_memory=self._pin)
            for _ in range(nslots)]
        if self.ragged:
            slots = []
        ragged_slots = [None] * nslots
        released = [None] * nslots          # CUDA event after the last kernel that read the slot
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        stop = threading.Event()

        def produce():
            try:
                for k in range(nb):
                    if stop.is_set():
                        return
                    s = k % nslots
                    idx = indices[k * self.batch_size:(k + 1) * self.batch_size]
                    if self._base is not None:
                        q.put((self.reader.record_offsets(idx, self.record_bytes), idx))
                        continue
                    ev = released[s]
                    if ev is not None:
                        ev.synchronize()
                    if self.ragged:
                        buf, off, _ = self.reader.get_ragged_batch(idx, out=ragged_slots[s], threads=self.threads)
                        ragged_slots[s] = buf            # grown when a batch did not fit
                        q.put(((s, off), idx))
                        continue
                    out = slots[s][:idx.shape[0]]
                    self.reader.get_batch(idx, self.record_bytes, out=out, threads=self.threads)
                    q.put((s, idx))
                q.put(None)
            except BaseException as e:      # surfaces in the consumer
                q.put(e)

        worker = threading.Thread(target=produce, daemon=True)
        worker.start()
        try:
            while True:
                item = q.get()
                if item is None:
                    return
                if isinstance(item, BaseException):
                    raise item
                s, idx = item
                if self._base is not None:
                    x = self.transform.from_records(self._base, s, self.shape)
                    y = None
                    if self.labels is not None:
                        y = self.labels[torch.from_numpy(idx)].pin_memory().to(self.transform.device, non_blocking=True)
                    yield x, y
                    continue
                if self.ragged:
                    slot, off = s
                    x = self.transform.ragged_records(ragged_slots[slot], off)
                    y = None
                    if self.labels is not None:
                        y = self.labels[torch.from_numpy(idx)].pin_memory().to(self.transform.device, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(torch.cuda.current_stream(self.transform.device))
                    released[slot] = ev
                    yield x, y
                    continue
                raw = slots[s][:idx.shape[0]].view(idx.shape[0], *self.shape)
                y = None
                if self.transform is None:
                    x = raw
                    if self.labels is not None:
                        y = self.labels[torch.from_numpy(idx)]
                else:
                    x = self.transform(raw)
                    if self.labels is not None:
                        yh = self.labels[torch.from_numpy(idx)]
                        y = (yh.pin_memory() if self._pin else yh).to(self.transform.device, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(torch.cuda.current_stream(self.transform.device))
                    released[s] = ev
                yield x, y
        finally:
            stop.set()
            try:                              # unblock a producer waiting on a full queue
                while True:
                    q.get_nowait()
            except queue.Empty:
                pass
