#!/usr/bin/env python
"""BASELINE.json config 5: SETP1+SETP2 training step at 480x640, batch-parallel over N GPUs.

  python tools/train_bench.py --frames 16 --steps 5                       # one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
         --master-port P tools/train_bench.py --frames 16 --steps 5       # N GPUs, NCCL

One step = what /root/reference/train_step2.py:59-65 does per iteration: model.train(); zero_grad;
forward of SETP2_BP_TRAIN (frozen step 1 with the EnforcePos hook firing, BatchNorm batch statistics
per GPU — the reference's nn.DataParallel has no SyncBN); multi-resolution loss; backward; ONE
all-reduce of the flat gradient bucket over NCCL (nconv_b200.distributed.GradBucket); AdamW step
(lr 1e-4, wd 1e-7: train_step2.py:17-18, utils.py:53-61).

`--frames` is the per-GPU batch (frames per stream = frames/2; config 5 = 16 per GPU x 8 GPUs = 128).
The loss is utils.py:63-71's multi-resolution masked MSE with a batch mean instead of the reference's
sample-0-only indexing, and the optimizer utils.py:53-55's AdamW; both run on the
package's own kernels (csrc/loss_optim.cu: one launch per scale per direction, one launch for the
update over the flat buffer whose gradient half is also the all-reduce bucket).  `--torch-loss-optim`
keeps the previous caller-side PyTorch loss + torch.optim.AdamW + GradBucket for comparison.
Prints one JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402


def multires_loss(depths, gt):
    """utils.py:63-71 + :138-151 with use_gradient_loss=False, batch mean, in eager PyTorch."""
    total = 0.0
    for d in depths:
        up = F.interpolate(d, size=gt.shape[2:], mode="bilinear", align_corners=False)
        total = total + F.mse_loss(up.masked_fill(gt == 0, 0), gt)
    return total / len(depths)


def make_trainer(dev, frames: int, H: int, W: int, rank: int = 0, fused: bool = True, seed_offset: int = 0):
    """Build SETP2_BP_TRAIN (frozen step 1 from a freshly written checkpoint, train_step2.py:134), its optimizer and
    one synthetic per-GPU batch; return (step, model, opt, bucket).  `step()` = train_step2.py:59-65 plus the
    data-parallel gradient all-reduce."""
    from models.step1 import SETP1_NCONV
    from models.step2 import SETP2_BP_TRAIN
    from nconv_b200.distributed import GradBucket, broadcast_parameters
    from nconv_b200.synthetic import setp2_inputs
    from nconv_b200 import train_utils as U
    per = frames // 2
    td = tempfile.mkdtemp()
    os.makedirs(os.path.join(td, "checkpoints"))
    torch.manual_seed(3)
    torch.save({"state_dict": SETP1_NCONV().state_dict()}, os.path.join(td, "checkpoints", "s1.pth.tar"))
    cwd = os.getcwd(); os.chdir(td)
    try:
        torch.manual_seed(5)
        model = SETP2_BP_TRAIN("s1").to(dev).train()
    finally:
        os.chdir(cwd)
    broadcast_parameters(model)
    params = [p for p in model.parameters() if p.requires_grad]
    if fused:
        opt = U.get_optimizer(model, "adam", 1e-4, 1e-7)
        bucket = None
    else:
        opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=1e-7)
        bucket = GradBucket(params)
    rgb0, d0, rgb1, d1 = (t.to(dev) for t in setp2_inputs(per, seed=100 + rank + seed_offset, H=H, W=W))
    g = torch.Generator().manual_seed(200 + rank + seed_offset)
    gt = (torch.rand(2 * per, 1, H, W, generator=g) * 9.5 + 0.5) * (torch.rand(2 * per, 1, H, W, generator=g) > 0.1)
    gt = gt.to(dev)

    def step(sync: bool = True):
        model.train()
        opt.zero_grad(set_to_none=False)
        a, b = model(rgb0, d0, rgb1, d1)
        depths = [torch.cat((x, y), 0) for x, y in zip(a, b)]
        if fused:
            loss = U.calculate_loss_multi_resolution(depths, gt, False, all_samples=True)
            loss.backward()
            if sync:
                opt.allreduce_mean()
        else:
            loss = multires_loss(depths, gt)
            loss.backward()
            if sync:
                bucket.allreduce_mean()
        opt.step()
        return loss

    return step, model, opt, bucket


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--height", type=int, default=480)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--torch-loss-optim", action="store_true")
    args = ap.parse_args()
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from nconv_b200 import _lib, ops

    fused = not args.torch_loss_optim
    step, model, opt, bucket = make_trainer(dev, args.frames, args.height, args.width, rank, fused)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    _lib.profile_enable(True)
    n0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    records = _lib.profile_records(); _lib.profile_enable(False)
    if rank == 0:
        by = {}
        for label, kms, fl, by_ in records:
            key = label.split(" @")[0]
            d = by.setdefault(key, [0.0, 0, 0.0]); d[0] += kms; d[1] += 1; d[2] += fl
        top = sorted(by.items(), key=lambda kv: -kv[1][0])[:56]
        print(json.dumps({
            "metric": "setp1+setp2_training_frames_per_s_480x640", "value": args.frames * world * args.steps / (ms / 1e3),
            "unit": "frames/s", "n_gpus": world, "steps": args.steps, "ms_per_step": ms / args.steps,
            "frames_per_gpu": args.frames, "global_batch": args.frames * world, "loss": float(loss),
            "loss_optim": "fused" if fused else "torch",
            "grad_bucket_bytes": opt.flat_grad.numel() * 4 if fused else bucket.nbytes, "gpu_launches_per_step": (ops.launch_count() - n0) / args.steps,
            "peak_mem_GB": torch.cuda.max_memory_allocated(dev) / 2 ** 30,
            "kernels": [{"kernel": k, "ms_per_step": round(v[0] / args.steps, 3), "launches_per_step": v[1] / args.steps,
                         "tflops": round(v[2] / v[0] / 1e9, 2) if v[0] > 0 else 0.0} for k, v in top]}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
