"""World-size-2 `gloo` tests (CPU) of the multi-GPU host logic in nconv_b200/distributed.py:
batch sharding (inference: no collective) and the flat-bucket gradient all-reduce (training).
The reference's counterpart is nn.DataParallel (/root/reference/train_step1.py:153)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _net():
    torch.manual_seed(3)
    m = torch.nn.Sequential(torch.nn.Conv2d(1, 4, 3, padding=1), torch.nn.ReLU(), torch.nn.Conv2d(4, 1, 3, padding=1))
    m.add_module("unused", torch.nn.BatchNorm2d(4))       # never reached by backward, like NConv2d.bnorm
    frozen = torch.nn.Conv2d(1, 1, 1)
    for p in frozen.parameters():
        p.requires_grad = False
    m.add_module("frozen", frozen)
    return m


def _fwd(m, x):
    return m[2](m[1](m[0](x)))


def _worker(rank, world, port, tmp):
    for p in (ROOT, PKG):
        sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from nconv_b200.distributed import GradBucket, broadcast_parameters, shard_batch, shard_range
        g = torch.Generator().manual_seed(11)
        x = torch.rand(6, 1, 8, 8, generator=g); y = torch.rand(6, 1, 8, 8, generator=g)
        m = _net()
        if rank == 1:                                   # start from different weights: broadcast must fix it
            with torch.no_grad():
                for p in m.parameters():
                    p.add_(1.0)
        broadcast_parameters(m)
        xs, ys = shard_batch([x, y], rank, world)
        lo, hi = shard_range(6, rank, world)
        assert xs.shape[0] == hi - lo
        # sum-of-squares loss normalised by the GLOBAL batch => mean over ranks of (local grads * world)
        loss = ((_fwd(m, xs) - ys) ** 2).sum() / 6 * world
        loss.backward()
        bucket = GradBucket(m.parameters())
        assert bucket.nbytes == 4 * sum(p.numel() for p in m.parameters() if p.requires_grad)
        pend = bucket.allreduce_mean(async_op=True)
        pend.wait()
        assert m.unused.weight.grad is None
        torch.save({k: (None if p.grad is None else p.grad.clone()) for k, p in m.named_parameters()},
                   os.path.join(tmp, "g%d.pt" % rank))
        # sharded inference: each rank computes its slice; gathered result equals the full-batch result
        with torch.no_grad():
            out = _fwd(m, xs)
        outs = [torch.empty(shard_range(6, r, world)[1] - shard_range(6, r, world)[0], 1, 8, 8) for r in range(world)]
        dist.all_gather(outs, out)
        torch.save(torch.cat(outs), os.path.join(tmp, "o%d.pt" % rank))
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions():
    from nconv_b200.distributed import shard_range
    for n in (0, 1, 5, 32, 33, 64):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


@pytest.mark.timeout(120)
def test_gloo_world2_bucket_allreduce_and_sharded_inference(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    # single-process ground truth on the whole batch
    g = torch.Generator().manual_seed(11)
    x = torch.rand(6, 1, 8, 8, generator=g); y = torch.rand(6, 1, 8, 8, generator=g)
    m = _net()
    (((_fwd(m, x) - y) ** 2).sum() / 6).backward()
    g0 = torch.load(os.path.join(tmp_path, "g0.pt")); g1 = torch.load(os.path.join(tmp_path, "g1.pt"))
    for k, p in m.named_parameters():
        if p.grad is None:
            assert g0[k] is None and g1[k] is None
            continue
        assert torch.equal(g0[k], g1[k]), k                               # ranks agree bit for bit
        assert torch.allclose(g0[k], p.grad, rtol=1e-5, atol=1e-7), k      # and match the 1-process gradient
    with torch.no_grad():
        full = _fwd(m, x)
    o0 = torch.load(os.path.join(tmp_path, "o0.pt"))
    assert torch.equal(o0, torch.load(os.path.join(tmp_path, "o1.pt")))
    assert torch.allclose(o0, full, rtol=1e-6, atol=1e-7)
