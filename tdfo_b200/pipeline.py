"""Input pipeline overlap: keep the next batches on their way to the GPU while the
current one trains.  The reference does this with `DataLoader(pin_memory=True)` + `.to(device)`
(torchrec/data.py:29-34, torchrec/train.py:97) and, in the jax variant, with
`flax.jax_utils.prefetch_to_device(it, size=2)` (jax-flax/train_dp.py:211), whose depth of two is the
default here.

    for batch in DevicePrefetcher(loader, device, depth=2):   # batch: tensors / KeyedJaggedTensor / dict / tuple
        loss = model(batch) ...

Host->device copies run on a dedicated copy stream from pinned memory; the consumer's stream waits on
the batch's event only (never on the host), and the batch's storage is tied to the consumer's stream
with `record_stream` so the allocator cannot hand it out again early.  On a CPU device the wrapper
only moves the batches (nothing to overlap), which is what the CPU tests exercise.
"""
from __future__ import annotations

from collections import deque
from typing import Any, Iterable, Iterator

import torch

from .sparse import JaggedTensor, KeyedJaggedTensor


def _map(obj: Any, fn) -> Any:
    if isinstance(obj, (torch.Tensor, KeyedJaggedTensor, JaggedTensor)):
        return fn(obj)
    if isinstance(obj, dict):
        return {k: _map(v, fn) for k, v in obj.items()}
    if isinstance(obj, tuple) and hasattr(obj, "_fields"):          # namedtuple
        return type(obj)(*(_map(v, fn) for v in obj))
    if isinstance(obj, (list, tuple)):
        return type(obj)(_map(v, fn) for v in obj)
    return obj


def _pin(x):
    if isinstance(x, torch.Tensor):
        return x if (x.is_cuda or x.is_pinned()) else x.pin_memory()
    if isinstance(x, KeyedJaggedTensor):
        return x if x.values().is_cuda or x.values().is_pinned() else x.pin_memory()
    return x


class DevicePrefetcher:
    """Iterate `iterable` with `depth` batches already copied (or being copied) to `device`."""

    def __init__(self, iterable: Iterable, device, depth: int = 2, pin: bool = True):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self._src = iterable
        self._device = torch.device(device)
        self._depth = depth
        self._pin = pin
        self._cuda = self._device.type == "cuda"
        self._copy_stream = torch.cuda.Stream(device=self._device) if self._cuda else None

    def _stage(self, batch):
        if not self._cuda:
            return _map(batch, lambda t: t.to(self._device)), None
        if self._pin:
            batch = _map(batch, _pin)
        with torch.cuda.stream(self._copy_stream):
            moved = _map(batch, lambda t: t.to(self._device, non_blocking=True))
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        return moved, ev

    def __iter__(self) -> Iterator:
        it = iter(self._src)
        queue: deque = deque()
        done = False
        while True:
            while not done and len(queue) < self._depth:
                try:
                    queue.append(self._stage(next(it)))
                except StopIteration:
                    done = True
            if not queue:
                return
            batch, ev = queue.popleft()
            if ev is not None:
                cur = torch.cuda.current_stream(self._device)
                cur.wait_event(ev)                      # device-side wait; the host runs on
                _map(batch, lambda t: t.record_stream(cur))
            yield batch
