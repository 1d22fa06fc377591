#!/usr/bin/env python
"""GB/s of the fp32 NCHW <-> P8 conversions and the fused ReLU-backward pack (training path), per shape."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
from nconv_b200 import ops
from nconv_b200.ops import P8
dev = torch.device("cuda:0")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for shape in [(16, 32, 480, 640), (16, 64, 240, 320), (16, 128, 240, 320), (16, 64, 60, 80)]:
    x = torch.randn(*shape, device=dev); y = torch.relu(torch.randn(*shape, device=dev))
    n = x.numel()
    p = P8.pack(x)
    out = {"shape": shape}
    t = timed(lambda: P8.pack(x)); out["pack_ms"] = round(t, 4); out["pack_gbs"] = round(8 * n / t / 1e6)
    t = timed(lambda: p.unpack()); out["unpack_ms"] = round(t, 4); out["unpack_gbs"] = round(8 * n / t / 1e6)
    t = timed(lambda: ops._relu_bias_bwd_p8(x, y, True)); out["relu_bwd_p8_ms"] = round(t, 4); out["relu_bwd_p8_gbs"] = round(12 * n / t / 1e6)
    t = timed(lambda: ops._relu_bias_bwd(x, y, True, True, True)); out["relu_bwd_ms"] = round(t, 4); out["relu_bwd_gbs"] = round(12 * n / t / 1e6)
    print(json.dumps(out))
