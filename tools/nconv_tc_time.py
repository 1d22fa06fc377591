#!/usr/bin/env python
"""Time one tc NConv layer (k5 with n8+pool outputs, or k3 tail) at B x H x W; used with NCONV_B200_NT_* switches."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
from nconv_b200.ops import N8, nconv_tc
B, H, W = (int(v) for v in sys.argv[1:4])
mode = sys.argv[4] if len(sys.argv) > 4 else "k5"
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(1)
x = (torch.rand(1, 8, H, W, generator=g) * 10).expand(B, 8, H, W).contiguous().to(dev)
c = (torch.rand(1, 8, H, W, generator=g) < 0.5).float().expand(B, 8, H, W).contiguous().to(dev)
a = N8.pack(x, c)
b8 = torch.zeros(8, device=dev)
if mode == "k5":
    w = torch.rand(8, 8, 5, 5, generator=g).to(dev) + 0.01
    fn = lambda: nconv_tc(a, None, w, b8, want_n8=True, want_pool=True)
else:
    w = torch.rand(8, 16, 3, 3, generator=g).to(dev) + 0.01
    w7 = torch.rand(1, 8, 1, 1, generator=g).to(dev); b7 = torch.zeros(1, device=dev)
    fn = lambda: nconv_tc(a, a, w, b8, a_first=False, tail=(w7, b7))
for _ in range(3):
    fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
from nconv_b200 import _lib
_lib.profile_enable(True)
for _ in range(10):
    fn()
torch.cuda.synchronize()
ms = [r[1] for r in _lib.profile_records() if r[0].startswith("tc_nconv 8")]
print("%s dbg=%s issuers=%s esets=%s: %.4f ms" % (mode, os.environ.get("NCONV_B200_NT_DEBUG"), os.environ.get("NCONV_B200_NT_ISSUERS"),
                                                    os.environ.get("NCONV_B200_NT_ESETS"), sorted(ms)[len(ms) // 2]))
