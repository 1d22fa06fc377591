#!/usr/bin/env python
"""BASELINE.json config 3: SETP1_NCONV batched inference on KITTI-shape 352x1216 sparse depth (~5 % density),
batch 64 sharded over N GPUs (8 frames/GPU at N=8; --frames sets the per-GPU batch), and config 1's shape
(--nyu: 480x640, 500 points).  One process per GPU, no collective.  Reports frames/s, ms/step and the
layer-wise algorithmic bytes/flops of SURVEY §8d (241.89 MB, 6.569 GFLOP per KITTI frame; 173.61 MB, 4.715 GFLOP
per NYU frame) as GB/s and TFLOP/s."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--nyu", action="store_true")
    ap.add_argument("--kernels", action="store_true")
    args = ap.parse_args()
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local); dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from models.step1 import SETP1_NCONV
    from nconv_b200 import ops
    from nconv_b200.synthetic import sparse_depth_kitti, sparse_depth_nyu
    torch.manual_seed(0)
    m = SETP1_NCONV().to(dev)
    m.train()
    with torch.no_grad():
        m(torch.zeros(1, 1, 16, 16, device=dev))     # one training-mode forward: the hook applies softplus
    m.eval()
    S = (sparse_depth_nyu(args.frames, seed=1 + rank) if args.nyu else sparse_depth_kitti(args.frames, seed=2000 + rank)).to(dev)
    with torch.no_grad():
        for _ in range(3):
            y = m(S)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        from nconv_b200 import _lib
        if args.kernels: _lib.profile_enable(True)
        n0 = ops.launch_count()
        e0.record()
        for _ in range(args.steps):
            y = m(S)
        e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    if args.kernels and rank == 0:
        by = {}
        for label, kms, fl, byt in _lib.profile_records():
            d = by.setdefault(label, [0.0, 0, 0.0]); d[0] += kms; d[1] += 1; d[2] += fl
        for k, v in by.items():
            print('   %-28s n=%3d  %.4f ms  %.1f TFLOP/s' % (k, v[1], v[0] / v[1], v[2] / v[0] / 1e9))
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
    mb, gf = (173.61, 4.715) if args.nyu else (241.89, 6.569)
    if rank == 0:
        print(json.dumps({"metric": "setp1_inference_frames_per_s_%s" % ("480x640" if args.nyu else "352x1216"),
                          "value": args.frames * world / (ms / 1e3), "unit": "frames/s", "n_gpus": world, "ms_per_step": ms,
                          "frames_per_gpu": args.frames, "launches_per_step": (ops.launch_count() - n0) / args.steps,
                          "layerwise_TBs": mb * args.frames / ms / 1e3, "layerwise_fp32_TFLOPs": gf * args.frames / ms,
                          "out_shape": list(y.shape)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
