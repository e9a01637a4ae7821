"""Host -> device staging for the uint8 batches (replaces DataLoader-pin + blocking ``.to("cuda")``).

The reference collates fp32 NCHW on the host and copies it with a blocking ``value.to("cuda")``
(``torchbooster/config.py:371-379,174-175``): 4 bytes per pixel-channel over PCIe and no overlap.
``DeviceStager`` ships the raw uint8 batch instead (4x fewer bytes; 8x for bf16 training) on a side
stream, one batch ahead of the compute stream, into a small ring of device buffers; CUDA events
hand each buffer over, so the transform of batch i runs while batch i+1 is still in flight.
"""
from __future__ import annotations

from typing import Iterable, Iterator, Sequence, Tuple

import torch


class DeviceStager:
    def __init__(self, device, depth: int = 2) -> None:
        if depth < 2:
            raise ValueError("depth must be >= 2 (one buffer in use, one in flight)")
        from ._native import cuda_device
        self.device = cuda_device(device)
        self.depth = depth
        self.copy_stream = torch.cuda.Stream(self.device)
        self._buffers = [dict() for _ in range(depth)]
        self._ready = [torch.cuda.Event() for _ in range(depth)]
        self._free = [None] * depth

    def _buffer(self, slot: int, pos: int, host: torch.Tensor) -> torch.Tensor:
        key = (pos, tuple(host.shape), host.dtype)
        buf = self._buffers[slot].get(key)
        if buf is None:
            buf = torch.empty(host.shape, dtype=host.dtype, device=self.device)
            self._buffers[slot][key] = buf
        return buf

    def _issue(self, slot: int, hosts: Sequence[torch.Tensor]):
        with torch.cuda.stream(self.copy_stream):
            if self._free[slot] is not None:
                self.copy_stream.wait_event(self._free[slot])
            devs = []
            for pos, h in enumerate(hosts):
                buf = self._buffer(slot, pos, h)
                buf.copy_(h, non_blocking=True)
                devs.append(buf)
            self._ready[slot].record(self.copy_stream)
        return tuple(devs)

    def run(self, batches: Iterable[Tuple[Sequence[torch.Tensor], object]]) -> Iterator[Tuple[tuple, object]]:
        """``batches`` yields ``(host_tensors, extra)``; this yields ``(device_tensors, extra)`` in
        order.  Host tensors should be pinned.  The device tensors of one step stay valid until the
        step after the next is requested."""
        it = iter(batches)
        compute = torch.cuda.current_stream(self.device)
        queue = []
        i = 0
        try:
            hosts, extra = next(it)
        except StopIteration:
            return
        queue.append((self._issue(0, hosts), extra, 0))
        while queue:
            nxt = None
            try:
                nxt = next(it)
            except StopIteration:
                pass
            if nxt is not None:
                slot = (i + 1) % self.depth
                queue.append((self._issue(slot, nxt[0]), nxt[1], slot))
            devs, extra, slot = queue.pop(0)
            compute.wait_event(self._ready[slot])
            yield devs, extra
            ev = torch.cuda.Event()
            ev.record(compute)
            self._free[slot] = ev
            i += 1
