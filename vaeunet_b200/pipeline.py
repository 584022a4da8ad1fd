"""Host <-> device staging around a device step (CUDA streams + events, no threads).

The reference moves every batch to the GPU synchronously and pulls results back with `.cpu()` (train.py:382-383,
visualize_vae.py:330,361): transfers and compute serialise. Here a step's inputs are copied from pinned host memory on a
copy-in stream while the previous step computes, and its results go back on a copy-out stream while the next step computes;
the compute stream only ever waits on events. At 512x512 a batch of 16 patches is 50 MB in and 50 MB out (three maps):
~2 ms of PCIe time per 14 ms step that disappears from the critical path.

    inp = StagedInput(device, depth=2)
    out = StagedOutput(device, depth=2)
    for host_batch in batches:                      # pinned host tensors
        x, = inp.push(host_batch)                   # device tensor, valid on the current stream
        inp.prefetch(next_host_batch)               # optional: its copy overlaps this step
        res = model.mc_predict(x, ...)
        inp.release()                               # the slot may be overwritten once the work enqueued so far is done
        out.push({k: res[k] for k in ("mean", "std", "entropy")}, on_ready=consume)   # consume(dict of pinned host tensors)
    out.flush()
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence

import torch


class StagedInput:
    """`depth` rotating device buffers filled on a side stream. push(*pinned_host_tensors) -> device tensors that are ready
    on the CURRENT stream (it waits on the copy's event); a buffer is only overwritten after the compute work that read it
    (everything enqueued on the current stream up to the next push of the same slot) has finished."""

    def __init__(self, device, depth: int = 2):
        self.dev = torch.device(device)
        self.depth = depth
        self.stream = torch.cuda.Stream(device=self.dev)
        self.bufs: List[Optional[List[torch.Tensor]]] = [None] * depth
        self.free: List[Optional[torch.cuda.Event]] = [None] * depth       # compute done with slot s
        self.i = 0
        self.cur = 0
        self._pending = None

    def _copy(self, host: Sequence[torch.Tensor]):
        s = self.i % self.depth
        self.i += 1
        if self.bufs[s] is None or any(b.shape != h.shape or b.dtype != h.dtype for b, h in zip(self.bufs[s], host)):
            self.bufs[s] = [torch.empty(h.shape, dtype=h.dtype, device=self.dev) for h in host]
        with torch.cuda.stream(self.stream):
            if self.free[s] is not None:
                self.stream.wait_event(self.free[s])
            for b, h in zip(self.bufs[s], host):
                b.copy_(h, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return s, ev

    def prefetch(self, *host: torch.Tensor) -> None:
        """Start the copy of the NEXT batch now (call before running the current step)."""
        assert self._pending is None
        self._pending = self._copy(host)

    def push(self, *host: torch.Tensor) -> List[torch.Tensor]:
        """Device tensors of this batch (uses the prefetched copy if there is one)."""
        s, ev = self._pending if self._pending is not None else self._copy(host)
        self._pending = None
        self.cur = s
        torch.cuda.current_stream(self.dev).wait_event(ev)
        return self.bufs[s]

    def release(self) -> None:
        """Mark the most recently pushed slot as consumed by everything enqueued on the current stream so far."""
        s = self.cur
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.dev))
        self.free[s] = ev


class StagedOutput:
    """Results go to pinned host buffers on a copy-out stream; `on_ready(host_dict)` runs when a slot is reused or flushed."""

    def __init__(self, device, depth: int = 2):
        self.dev = torch.device(device)
        self.depth = depth
        self.stream = torch.cuda.Stream(device=self.dev)
        self.host: List[Optional[Dict[str, torch.Tensor]]] = [None] * depth
        self.done: List[Optional[torch.cuda.Event]] = [None] * depth
        self.cb: List[Optional[Callable]] = [None] * depth
        self.i = 0

    def _retire(self, s: int) -> None:
        if self.done[s] is not None:
            self.done[s].synchronize()
            if self.cb[s] is not None:
                self.cb[s](self.host[s])
            self.done[s], self.cb[s] = None, None

    def push(self, tensors: Dict[str, torch.Tensor], on_ready: Optional[Callable] = None) -> None:
        s = self.i % self.depth
        self.i += 1
        self._retire(s)
        if self.host[s] is None or any(k not in self.host[s] or self.host[s][k].shape != t.shape for k, t in tensors.items()):
            self.host[s] = {k: torch.empty(t.shape, dtype=t.dtype).pin_memory() for k, t in tensors.items()}
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ready)
            for k, t in tensors.items():
                self.host[s][k].copy_(t, non_blocking=True)
                t.record_stream(self.stream)            # the caching allocator must not recycle t before the copy ran
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.done[s], self.cb[s] = ev, on_ready

    def flush(self) -> None:
        for k in range(self.depth):
            self._retire((self.i + k) % self.depth)
