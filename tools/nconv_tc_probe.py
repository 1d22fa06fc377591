#!/usr/bin/env python
"""Debug aid: one tc NConv layer at a given batch/size with selected outputs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
from nconv_b200.ops import N8, nconv_tc
B, H, W = (int(v) for v in sys.argv[1:4])
flags = sys.argv[4] if len(sys.argv) > 4 else "n8,pool"
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(1)
x = torch.rand(1, 8, H, W, generator=g).expand(B, 8, H, W).contiguous().to(dev) * 10
c = (torch.rand(1, 8, H, W, generator=g) < 0.5).float().expand(B, 8, H, W).contiguous().to(dev)
w = torch.rand(8, 8, 5, 5, generator=g).to(dev) + 0.01
b = torch.zeros(8, device=dev)
a = N8.pack(x, c)
del x, c
torch.cuda.synchronize()
print("packed", flush=True)
out = nconv_tc(a, None, w, b, want_n8="n8" in flags, want_pool="pool" in flags, want_up="up" in flags, want_f32="f32" in flags)
torch.cuda.synchronize()
print("ok", B, H, W, flags, {k: (v.buf.numel() if hasattr(v, "buf") else v.numel()) for k, v in out.items()})
