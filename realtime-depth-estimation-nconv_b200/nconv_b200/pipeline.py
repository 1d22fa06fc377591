"""Host<->device pipelining for batched inference through the module call.

The reference's loops copy a batch to the GPU, run the model and read the result back serially
(/root/reference/train_step1.py:53-54,61; utils.py:28-36).  At >1000 frames/s per GPU the PCIe copies
(4.9 MB of float RGB+depth in, 1.2 MB out per frame) are 10-15 % of a step if left in line, so this
helper double-buffers them on two copy streams: while batch i computes, batch i+1 uploads from pinned
host memory and batch i-1's result downloads.  Every batch still crosses PCIe in both directions —
nothing is cached or skipped — the copies just overlap the kernels.
"""
from __future__ import annotations

from typing import Callable, Iterable, Optional, Sequence

import torch


class StreamedInference:
    """run(batches, consume=None) -> list with one entry per batch, in order.

    `model(*device_inputs)` must return a tensor or a tuple of tensors; `batches` yields sequences of pinned host
    tensors of identical shapes.  Every batch's outputs are downloaded into one of `depth` pinned host slots; as
    soon as a slot's download has finished — and before the slot is reused — its tensors are handed to
    `consume(index, host_outputs)` (a view of the pinned slot, valid only during the call: copy or reduce there) and
    the value `consume` returns is what `run` collects.  The default `consume` clones the slot, so `run(batches)`
    returns every batch's outputs as ordinary host tensors."""

    def __init__(self, model: Callable, device: torch.device, depth: int = 2):
        if depth < 2:
            raise ValueError("StreamedInference: depth >= 2 (double buffering)")
        self.model, self.dev, self.depth = model, device, depth
        self.h2d = torch.cuda.Stream(device)
        self.d2h = torch.cuda.Stream(device)
        self._in = None       # depth x list of device input buffers
        self._out = None      # depth x list of device outputs kept alive until downloaded
        self._host = None     # depth x list of pinned host outputs

    def _alloc_inputs(self, batch: Sequence[torch.Tensor]):
        self._in = [[torch.empty(t.shape, dtype=t.dtype, device=self.dev) for t in batch] for _ in range(self.depth)]

    def run(self, batches: Iterable[Sequence[torch.Tensor]], consume: Optional[Callable] = None):
        if consume is None:
            consume = lambda i, outs: [o.clone() for o in outs]      # noqa: E731
        compute = torch.cuda.current_stream(self.dev)
        it = iter(batches)
        slots_ready = [None] * self.depth          # event: upload of slot s finished
        slots_free = [None] * self.depth           # event: compute that read slot s finished
        down_done = [None] * self.depth            # (batch index, event: download of slot s finished), not yet consumed
        results = []

        def retire(s):
            """Hand slot s's finished download to the consumer; after this the pinned slot may be overwritten."""
            if down_done[s] is not None:
                idx, ev = down_done[s]
                ev.synchronize()
                results.append((idx, consume(idx, self._host[s])))
                down_done[s] = None

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
            return []
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
                retire(s)                                      # batch i - depth leaves this slot before it is refilled
                with torch.cuda.stream(self.d2h):
                    self.d2h.wait_event(done)
                    self._out[s] = outs
                    for h, o in zip(self._host[s], outs):
                        h.copy_(o, non_blocking=True)
                        o.record_stream(self.d2h)
                    ev = torch.cuda.Event(); ev.record(self.d2h)
                    down_done[s] = (i, ev)
                i += 1
        for k in range(self.depth):                            # drain in batch order
            retire((i + k) % self.depth)
        results.sort(key=lambda r: r[0])
        return [r for _, r in results]


class GraphedInference:
    """The module's inference forward captured ONCE in a CUDA graph and replayed per call: one `cudaGraphLaunch` instead of
    ~50 kernel launches plus the Python between them (the step is launch-latency-free, and on an 8-GPU box the eight ranks
    no longer compete for host cores to enqueue).  Shapes are fixed at capture; `__call__` copies the caller's device
    tensors into the graph's static input buffers (device-to-device, ~0.05 ms for a 32-frame batch) and returns the
    graph's static outputs — valid until the next call on this object (copy them if they must outlive it).

    Everything the launchers do lazily (kernel attributes, driver entry points, packed weights, P8 / N8 workspaces, scratch
    buffers of the capture stream) happens in the warm-up passes, which run on the capture stream before capturing.
    The captured graph bakes in the parameter pointers AND the packed weights: re-capture (`GraphedInference(...)` again)
    after the weights change."""

    def __init__(self, model: Callable, example_inputs: Sequence[torch.Tensor], warmup: int = 3):
        self.model = model
        self.static_in = [t.detach().clone() for t in example_inputs]
        dev = self.static_in[0].device
        self.graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.no_grad(), torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                out = model(*self.static_in)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        with torch.no_grad(), torch.cuda.graph(self.graph, stream=side):
            out = model(*self.static_in)
        self.static_out = list(out) if isinstance(out, (tuple, list)) else [out]
        self._tuple = isinstance(out, (tuple, list))

    def __call__(self, *inputs: torch.Tensor):
        if len(inputs) != len(self.static_in):
            raise ValueError("GraphedInference: expected %d inputs" % len(self.static_in))
        for dst, src in zip(self.static_in, inputs):
            if src.shape != dst.shape or src.dtype != dst.dtype:
                raise ValueError("GraphedInference: input shape/dtype differs from the captured one")
            if src.data_ptr() != dst.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return tuple(self.static_out) if self._tuple else self.static_out[0]
