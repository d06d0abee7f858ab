"""Host -> device input pipeline for the step loop.

The reference moves each batch with `images.to(device, non_blocking=True)` on the compute stream
(/root/reference/train_advanced_v4.py:211-213, DataLoader(pin_memory=True) :608-616) and reads the
loss back with `loss.item()` every batch (:268), which serialises the copy with the kernels and
drains the launch queue once per batch.  Here the copy of batch i+1 runs on a second stream while
batch i computes, and scalar results travel through pinned memory without stalling the host.
"""
from __future__ import annotations

import torch


_copy_streams: dict = {}      # one copy stream per device, shared by every prefetcher


def _copy_stream(device: torch.device) -> "torch.cuda.Stream":
    """The caching allocator keeps one block pool per stream: a new stream per epoch would
    cudaMalloc its staging buffers again every time (measured: 5 cudaMallocs per prefetcher)."""
    key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
    if key not in _copy_streams:
        _copy_streams[key] = torch.cuda.Stream(device)
    return _copy_streams[key]


class DevicePrefetcher:
    """Iterate `(images, labels, ...)` batches of pinned host tensors as device tensors, one batch
    ahead: `for images, labels in DevicePrefetcher(loader, device): ...`"""

    def __init__(self, loader, device):
        self.loader, self.device = loader, torch.device(device)
        self.stream = _copy_stream(self.device)

    def _load(self, it):
        try:
            batch = next(it)
        except StopIteration:
            return None
        with torch.cuda.stream(self.stream):
            dev = [t.to(self.device, non_blocking=True) if torch.is_tensor(t) else t for t in batch]
        ev = torch.cuda.Event()
        ev.record(self.stream)
        return dev, ev

    def __iter__(self):
        it = iter(self.loader)
        nxt = self._load(it)
        while nxt is not None:
            dev, ev = nxt
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            for t in dev:
                if torch.is_tensor(t):
                    t.record_stream(cur)           # allocated on the copy stream, used on this one
            nxt = self._load(it)                   # next copy is in flight while the caller computes
            yield tuple(dev)


class AsyncScalarReader:
    """Device scalar -> host without a stall: `push(t)` starts a copy into pinned memory, `pop()`
    returns the oldest value whose copy has finished being waited for (one step of latency)."""

    def __init__(self, device, depth: int = 2):
        self.device = torch.device(device)
        self.slots = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(depth)]
        self.events = [None] * depth
        self.head = 0          # next slot to write
        self.pending = []      # slots in flight, oldest first

    def push(self, t: torch.Tensor):
        vals = []
        if len(self.pending) == len(self.slots):      # ring full: the oldest copy finished long ago
            vals.append(self.pop())
        i = self.head
        self.slots[i].copy_(t.detach().reshape(1).float(), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.events[i] = ev
        self.pending.append(i)
        self.head = (i + 1) % len(self.slots)
        return vals

    def pop(self) -> float:
        i = self.pending.pop(0)
        self.events[i].synchronize()
        return float(self.slots[i][0])

    def drain(self):
        return [self.pop() for _ in range(len(self.pending))]
