"""Multi-GPU host logic for the NConv path: one process per GPU (torchrun), NCCL for the plumbing.

The reference's only parallelism is `nn.DataParallel` (/root/reference/train_step1.py:153,
train_step2.py:135): a single process scatters the batch, replicates the parameters and reduces
the gradients onto GPU 0 every iteration.  The B200 form of the same semantics:

* inference — frames are independent (no cross-sample op, eval-mode BatchNorm), so the batch axis
  is partitioned across ranks with NO data-path collective (`shard_range`, `shard_batch`);
* training — batch-parallel replicas; BatchNorm statistics stay per GPU (as under DataParallel, the
  reference has no SyncBN); the only exchange is ONE NCCL all-reduce per step over a flat fp32
  bucket of every gradient (`GradBucket`): SETP2 has 978 336 trainable parameters = 3.9 MB, SETP1
  10 129 = 40 KB, so the collective is latency-bound and launching it once matters more than
  overlapping it.  Parameters that never receive a gradient (frozen step 1, NConv2d's unused
  `bnorm`, SETP2_BP_TRAIN's leftover `rgb_encoder4`) are left out of the bucket.

Nothing here touches CUDA directly: it works on whatever device the tensors live on, which is what
lets the world-size-2 `gloo` tests run on CPU.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced partition of range(n): the first n % world ranks get one extra item."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("shard_range: bad rank/world %d/%d" % (rank, world))
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(tensors: Sequence[torch.Tensor], rank: int, world: int) -> List[torch.Tensor]:
    """This rank's slice of every tensor along the batch axis (all must share dim 0)."""
    n = tensors[0].shape[0]
    if any(t.shape[0] != n for t in tensors):
        raise ValueError("shard_batch: tensors disagree on the batch size")
    lo, hi = shard_range(n, rank, world)
    return [t[lo:hi] for t in tensors]


class GradBucket:
    """Flat fp32 gradient bucket: one all-reduce per training step.

    The parameter list is fixed at construction (every parameter with requires_grad); a parameter
    whose .grad is still None at reduce time (never reached by backward, e.g. the unused BatchNorm
    of an NConv2d) contributes zeros on every rank and keeps .grad None afterwards, exactly as it
    would on one GPU."""

    def __init__(self, params: Iterable[torch.nn.Parameter]):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("GradBucket: no trainable parameters")
        dev = self.params[0].device
        if any(p.device != dev or p.dtype != torch.float32 for p in self.params):
            raise ValueError("GradBucket: parameters must be float32 on one device")
        self.offsets = []
        total = 0
        for p in self.params:
            self.offsets.append(total)
            total += p.numel()
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)

    @property
    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def allreduce_mean(self, group=None, async_op: bool = False):
        """grad <- mean over ranks of grad, for every parameter, with one collective."""
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        views = []
        self.flat.zero_()
        for p, off in zip(self.params, self.offsets):
            v = self.flat[off:off + p.numel()]
            if p.grad is not None:
                v.copy_(p.grad.reshape(-1))
            views.append(v)
        work = None
        if world > 1:
            work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if async_op and work is not None:
            return _Pending(self, work, world, views)
        self._scatter(world, views)
        return None

    def _scatter(self, world, views):
        if world > 1:
            self.flat.mul_(1.0 / world)
        for p, v in zip(self.params, views):
            if p.grad is not None:
                p.grad.copy_(v.view_as(p.grad))


class _Pending:
    def __init__(self, bucket, work, world, views):
        self.bucket, self.work, self.world, self.views = bucket, work, world, views

    def wait(self):
        self.work.wait()
        self.bucket._scatter(self.world, self.views)


def broadcast_parameters(module: torch.nn.Module, src: int = 0, group=None) -> None:
    """Make every rank start from rank `src`'s parameters and buffers (DataParallel's `replicate`)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=src, group=group)
