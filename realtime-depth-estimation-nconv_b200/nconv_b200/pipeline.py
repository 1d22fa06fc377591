"""Host<->device pipelining for batched inference through the module call.

The reference's loops copy a batch to the GPU, run the model and read the result back serially
(/root/reference/train_step1.py:53-54,61; utils.py:28-36).  At >1000 frames/s per GPU the PCIe copies
(4.9 MB of float RGB+depth in, 1.2 MB out per frame) are 10-15 % of a step if left in line, so this
helper double-buffers them on two copy streams: while batch i computes, batch i+1 uploads from pinned
host memory and batch i-1's result downloads.  Every batch still crosses PCIe in both directions —
nothing is cached or skipped — the copies just overlap the kernels.
"""
from __future__ import annotations

from typing import Callable, Iterable, List, Sequence

import torch


class StreamedInference:
    """run(batches) -> list of host outputs.  `model(*device_inputs)` must return a tensor or a tuple
    of tensors; `batches` yields sequences of pinned host tensors of identical shapes."""

    def __init__(self, model: Callable, device: torch.device, depth: int = 2):
        self.model, self.dev, self.depth = model, device, depth
        self.h2d = torch.cuda.Stream(device)
        self.d2h = torch.cuda.Stream(device)
        self._in = None       # depth x list of device input buffers
        self._out = None      # depth x list of device outputs kept alive until downloaded
        self._host = None     # depth x list of pinned host outputs

    def _alloc_inputs(self, batch: Sequence[torch.Tensor]):
        self._in = [[torch.empty(t.shape, dtype=t.dtype, device=self.dev) for t in batch] for _ in range(self.depth)]

    def run(self, batches: Iterable[Sequence[torch.Tensor]], out_host: List[List[torch.Tensor]] = None):
        compute = torch.cuda.current_stream(self.dev)
        it = iter(batches)
        slots_ready = [None] * self.depth          # event: upload of slot s finished
        slots_free = [None] * self.depth           # event: compute that read slot s finished
        down_done = [None] * self.depth            # event: download of slot s finished
        results = []

        def upload(s, batch):
            if self._in is None:
                self._alloc_inputs(batch)
            with torch.cuda.stream(self.h2d):
                if slots_free[s] is not None:
                    self.h2d.wait_event(slots_free[s])
                for dst, src in zip(self._in[s], batch):
                    dst.copy_(src, non_blocking=True)
                ev = torch.cuda.Event(); ev.record(self.h2d)
                slots_ready[s] = ev

        nxt = next(it, None)
        if nxt is None:
            return results
        upload(0, nxt)
        i = 0
        with torch.no_grad():
            while nxt is not None:
                s = i % self.depth
                nxt = next(it, None)
                if nxt is not None:
                    upload((i + 1) % self.depth, nxt)          # overlaps the compute of batch i
                compute.wait_event(slots_ready[s])
                out = self.model(*self._in[s])
                outs = list(out) if isinstance(out, (tuple, list)) else [out]
                done = torch.cuda.Event(); done.record(compute)
                slots_free[s] = done
                if self._host is None:
                    self._host = [[torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in outs] for _ in range(self.depth)]
                    self._out = [None] * self.depth
                with torch.cuda.stream(self.d2h):
                    self.d2h.wait_event(done)
                    if down_done[s] is not None:
                        down_done[s].synchronize()           # the host buffer of this slot was consumed last round
                    self._out[s] = outs
                    for h, o in zip(self._host[s], outs):
                        h.copy_(o, non_blocking=True)
                        o.record_stream(self.d2h)
                    ev = torch.cuda.Event(); ev.record(self.d2h)
                    down_done[s] = ev
                results.append((s, down_done[s]))
                i += 1
        for ev in down_done:
            if ev is not None:
                ev.synchronize()
        return [self._host[s] for s, _ in results[-self.depth:]]
