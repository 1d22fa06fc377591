#!/usr/bin/env python
"""Long-run hang / corruption check (run under `timeout`): many back-to-back steps of the three measured workloads,
results compared with the first step's (the kernels are deterministic: bit-identical or it is a defect).
  config 4 (SETP1+SETP2 inference, 32 frames) x N4, config 3 (SETP1, 64 KITTI frames) x N3, config 5 (training) x N5"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200"), os.path.join(ROOT, "tools")):
    sys.path.insert(0, p)
import torch
import bench
from nconv_b200.synthetic import setp2_inputs, sparse_depth_kitti
N4, N3, N5 = (int(v) for v in (sys.argv[1:4] + ["1000", "600", "40"][len(sys.argv) - 1:]))
dev = torch.device("cuda:0")
out = {}
model = bench.build_model(dev)
res = [t.to(dev) for t in setp2_inputs(bench.FRAMES_PER_STREAM, seed=7)]
t0 = time.time()
with torch.no_grad():
    first = [o.clone() for o in model(*res)]
    bad = 0
    for i in range(N4):
        o = model(*res)
        if i % 50 == 49:
            bad += int(not all(torch.equal(a, b) for a, b in zip(o, first)))
torch.cuda.synchronize()
out["config4_steps"] = N4; out["config4_mismatches"] = bad; out["config4_s"] = round(time.time() - t0, 1)
from models.step1 import SETP1_NCONV
torch.manual_seed(0)
m1 = SETP1_NCONV().to(dev).eval()
S = sparse_depth_kitti(64, seed=2000).to(dev)
t0 = time.time()
with torch.no_grad():
    first = m1(S).clone()
    bad = 0
    for i in range(N3):
        o = m1(S)
        if i % 50 == 49:
            bad += int(not torch.equal(o, first))
torch.cuda.synchronize()
out["config3_steps"] = N3; out["config3_mismatches"] = bad; out["config3_s"] = round(time.time() - t0, 1)
del m1, S, model, res
import train_bench
step, _, _, _ = train_bench.make_trainer(dev, 16, 480, 640, rank=0)
t0 = time.time()
losses = [float(step()) for _ in range(N5)]
torch.cuda.synchronize()
out["config5_steps"] = N5; out["config5_finite"] = all(l == l and abs(l) < 1e30 for l in losses); out["config5_s"] = round(time.time() - t0, 1)
out["config5_loss_first_last"] = [losses[0], losses[-1]]
print(json.dumps(out))
