#!/usr/bin/env python
"""Run one NConv2d forward shape a few times (for an ncu capture of a single kernel).
Usage: python tools/prof_nconv.py CI CO K [B] [--bwd]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
from nconv_b200 import ops  # noqa
ci, co, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
B = int(sys.argv[4]) if len(sys.argv) > 4 and sys.argv[4].isdigit() else 16
bwd = "--bwd" in sys.argv
dev = torch.device("cuda:0")
torch.manual_seed(0)
x = torch.rand(B, ci, 480, 640, device=dev) * 10
c = torch.rand(B, ci, 480, 640, device=dev) * (torch.rand(B, ci, 480, 640, device=dev) < 0.5)
w = torch.nn.functional.softplus(torch.rand(co, ci, k, k, device=dev), beta=10)
b = torch.full((co,), 0.01, device=dev)
K = torch.ops.nconv_b200
for _ in range(4):
    y, cc = K.nconv2d(x, c, w, b, k // 2, k // 2, 1e-7)
    if bwd:
        K.nconv2d_backward(torch.ones_like(y), torch.ones_like(cc), x, c, w, b, y, cc, k // 2, k // 2, 1e-7, True, True, True)
torch.cuda.synchronize()
print("ok", y.shape)
