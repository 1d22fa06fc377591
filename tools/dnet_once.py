#!/usr/bin/env python
"""One SETP1 forward at a given batch/size (debug aid: run under CUDA_LAUNCH_BLOCKING=1 or compute-sanitizer)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
from models.step1 import SETP1_NCONV
from nconv_b200.synthetic import sparse_depth_kitti
B, H, W = (int(v) for v in sys.argv[1:4])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 1
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = SETP1_NCONV().to(dev).eval()
S = sparse_depth_kitti(B, seed=2000, H=H, W=W).to(dev)
from nconv_b200 import _lib
dbg = torch.zeros(16, dtype=torch.int32).pin_memory()
_lib.load().nconv_b200_debug_buffer(dbg.data_ptr())
import atexit
atexit.register(lambda: print("dbg", dbg.tolist()[:5], flush=True))
with torch.no_grad():
    for i in range(reps):
        y = m(S)
        torch.cuda.synchronize()
print("ok", B, H, W, float(y.mean()))
