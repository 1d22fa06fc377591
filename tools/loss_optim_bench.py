#!/usr/bin/env python
"""Times the training-loop callers (csrc/loss_optim.cu) against the eager PyTorch sequence the reference runs
(utils.py:63-71, :138-151, :53-55) on the same GPU: 4-scale masked MSE forward+backward at 480x640, the
0.8*RMSE+0.2*Sobel-L1 loss of train_step1.py, and AdamW over the SETP2 parameter count.

  python tools/loss_optim_bench.py [--frames 16] [--iters 20]

Inputs are larger than nothing in particular (a 16-frame batch is 19.7 MB per full-resolution map), so L2 is
flushed (clean: write + read back) before every timed iteration.  One JSON line per case.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    from nconv_b200 import train_utils as U
    from nconv_b200 import _lib, ops
    from nconv_b200.synthetic import loss_case

    def kernel_breakdown(fn):
        _lib.profile_enable(True)
        fn()
        torch.cuda.synchronize()
        rec = [(label.split(" @")[0], round(ms, 4)) for label, ms, _, _ in _lib.profile_records()]
        _lib.profile_enable(False)
        return rec

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn):
        for _ in range(3):
            fn()
        tot = 0.0
        n0 = ops.launch_count()
        for _ in range(args.iters):
            flush.fill_(1); flush.sum()
            torch.cuda._sleep(4000000)          # ~2 ms: the host enqueues the whole region before the device starts it
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / args.iters, (ops.launch_count() - n0) / args.iters

    B, H, W = args.frames, 480, 640
    depths, gt = loss_case(B, 5)
    gt = gt.to(dev)
    depths = [d.to(dev).requires_grad_(True) for d in depths]
    pix = B * H * W
    lo = sum(d.numel() for d in depths)

    def torch_mr(sobel):
        def f():
            for d in depths:
                d.grad = None
            total = 0.0
            for d in depths:
                up = F.interpolate(d, size=(H, W), mode="bilinear", align_corners=False)
                rec = up.masked_fill(gt == 0, 0)
                if sobel:
                    kx = torch.tensor([[[[1., 0, -1], [2, 0, -2], [1, 0, -1]]]], device=dev)
                    ky = torch.tensor([[[[1., 2, 1], [0, 0, 0], [-1, -2, -1]]]], device=dev)
                    diff = gt - rec
                    total = total + torch.sqrt(F.mse_loss(rec, gt)) * 0.8 + \
                        (F.conv2d(diff, kx, padding=1).abs().mean() + F.conv2d(diff, ky, padding=1).abs().mean()) * 0.2
                else:
                    total = total + F.mse_loss(rec, gt)
            (total / len(depths)).backward()
        return f

    def fused_mr(sobel):
        def f():
            for d in depths:
                d.grad = None
            U.calculate_loss_multi_resolution(depths, gt, sobel, all_samples=True).backward()
        return f

    for sobel in (False, True):
        t_ref, _ = timed(torch_mr(sobel))
        t_new, launches = timed(fused_mr(sobel))
        # algorithmic traffic: forward reads gt + every scale, backward reads them again and writes the gradients
        bytes_ = 4.0 * (len(depths) * 2 * pix + 3 * lo)
        print(json.dumps({"case": "multires_loss_fwd_bwd" + ("_sobel" if sobel else ""), "frames": B, "scales": len(depths),
                          "torch_eager_ms": round(t_ref, 4), "nconv_b200_ms": round(t_new, 4),
                          "speedup": round(t_ref / t_new, 2), "launches": launches,
                          "algorithmic_GB": round(bytes_ / 1e9, 4), "GBps": round(bytes_ / t_new / 1e6, 1),
                          "kernels_ms": kernel_breakdown(fused_mr(sobel))}))

    # AdamW over SETP2_BP_TRAIN's trainable parameter count
    n = 1_700_000
    torch.manual_seed(1)
    shapes = [(64, 64, 3, 3)] * 40 + [(64,)] * 80 + [(32, 32, 3, 3)] * 20 + [(n - 40 * 36864 - 80 * 64 - 20 * 9216,)]
    pa = [torch.nn.Parameter(torch.randn(s, device=dev)) for s in shapes]
    pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
    for p in pa + pb:
        p.grad = torch.randn_like(p)
    oa = torch.optim.AdamW(pa, lr=1e-4, weight_decay=1e-7)
    ob = ops.FusedAdamW(pb, lr=1e-4, weight_decay=1e-7)
    t_ref, _ = timed(oa.step)
    t_new, launches = timed(ob.step)
    nel = sum(p.numel() for p in pa)
    print(json.dumps({"case": "adamw_step", "parameters": nel, "tensors": len(pa), "torch_ms": round(t_ref, 4),
                      "nconv_b200_ms": round(t_new, 4), "speedup": round(t_ref / t_new, 2), "launches": launches,
                      "GBps": round(28.0 * nel / t_new / 1e6, 1)}))


if __name__ == "__main__":
    main()
