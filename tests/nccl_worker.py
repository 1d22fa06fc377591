"""Worker of tests/test_gpu_multi.py, launched once per GPU by torch.distributed.run (NCCL).

Checks on hardware what tests/test_distributed_cpu.py proves on gloo with a toy net (SURVEY §4 tier 4, §8e):
  1. `FusedAdamW.allreduce_mean` (the path tools/train_bench.py and bench.py's train_step leg use: ONE NCCL
     all-reduce on the flat gradient buffer) — the reduced SETP2_BP_TRAIN gradients of `world` shards equal the
     mean of the single-GPU gradients of each shard (BatchNorm batch statistics per shard, as nn.DataParallel
     computes them: the reference has no SyncBN), within 1e-5 of each tensor's max magnitude;
  2. after one optimizer step every rank holds bit-identical parameters;
  3. batch-sharded inference is bit-identical to the full-batch result (no data-path collective).
Rank 0 writes the figures as JSON to argv[1]."""
import copy
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def grads_of(model, batch, gt, U):
    model.train()
    a, b = model(*batch)
    depths = [torch.cat((x, y), 0) for x, y in zip(a, b)]
    loss = U.calculate_loss_multi_resolution(depths, gt, False, all_samples=True)
    loss.backward()
    return loss


def main():
    out_path = sys.argv[1]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from nconv_b200 import ops
    from nconv_b200 import train_utils as U
    from nconv_b200.distributed import shard_batch
    from nconv_b200.synthetic import setp2_inputs
    import train_bench

    H, W, per = 96, 128, 2                       # frames per stream per rank
    _, model0, _, _ = train_bench.make_trainer(dev, 2 * per, H, W, rank=0)      # same seeds on every rank + broadcast
    # the global batch, generated identically on every rank, sharded on the batch axis
    full = [t.to(dev) for t in setp2_inputs(per * world, seed=321, H=H, W=W)]
    g = torch.Generator().manual_seed(99)
    gt_s = [(torch.rand(per * world, 1, H, W, generator=g) * 9.5 + 0.5) * (torch.rand(per * world, 1, H, W, generator=g) > 0.1)
            for _ in range(2)]                   # ground truth per stream

    def shard(r):
        b = shard_batch(full, r, world)
        gts = shard_batch(gt_s, r, world)
        return b, torch.cat(gts, 0).to(dev)

    # ---- 1. distributed: this rank's shard, then ONE all-reduce on the flat buffer ------------------------------
    m = copy.deepcopy(model0)
    opt = ops.FusedAdamW([p for p in m.parameters() if p.requires_grad], lr=1e-3, weight_decay=1e-7)
    batch, gt = shard(rank)
    grads_of(m, batch, gt, U)
    opt.allreduce_mean()
    torch.cuda.synchronize()
    names = [n for n, p in m.named_parameters() if p.grad is not None]
    g_dist = {n: p.grad.detach().clone() for n, p in m.named_parameters() if p.grad is not None}
    bucket_bytes = opt.flat_grad.numel() * 4
    opt.step()
    # ---- 2. parameters identical on every rank after the step -----------------------------------------------------
    flat = opt.flat_param.detach().clone()
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    params_identical = all(torch.equal(gathered[0], t) for t in gathered[1:])
    # ---- single-GPU reference on rank 0: every shard separately from the same initial state, gradients averaged -----
    result = None
    if rank == 0:
        acc = None
        for r in range(world):
            mr = copy.deepcopy(model0)
            b, gtr = shard(r)
            grads_of(mr, b, gtr, U)
            gr = {n: p.grad.detach().double() for n, p in mr.named_parameters() if p.grad is not None}
            assert sorted(gr) == sorted(names)
            acc = gr if acc is None else {n: acc[n] + gr[n] for n in acc}
        worst, worst_name = 0.0, None
        for n in names:
            ref = acc[n] / world
            scale = float(ref.abs().max())
            if scale < 1e-12:
                continue
            err = float((g_dist[n].double() - ref).abs().max()) / scale
            if err > worst:
                worst, worst_name = err, n
        result = {"world": world, "tensors": len(names), "grad_bucket_bytes": bucket_bytes,
                  "allreduce_grad_max_rel_err": worst, "worst_tensor": worst_name, "params_identical_after_step": params_identical}
    # ---- 3. sharded inference == full-batch inference, bit for bit -------------------------------------------------
    me = copy.deepcopy(model0).eval()
    with torch.no_grad():
        a, b = me(*batch)
        mine = torch.cat((a[3], b[3]), 0).contiguous()
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        if rank == 0:
            fa, fb = me(*full)
            ok = True
            for r in range(world):
                lo, hi = r * per, (r + 1) * per
                ok = ok and torch.equal(parts[r][:per], fa[3][lo:hi]) and torch.equal(parts[r][per:], fb[3][lo:hi])
            result["sharded_inference_bit_identical"] = bool(ok)
    if rank == 0:
        with open(out_path, "w") as f:
            json.dump(result, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
