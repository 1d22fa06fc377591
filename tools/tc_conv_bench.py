#!/usr/bin/env python
"""Per-layer timing of the tcgen05 conv3x3 on the SETP2 layer shapes (batch 32, 480x640 pyramid).
Reports ms, algorithmic TFLOP/s (2*Ci*Co*9*px; the kernel issues 3x that in bf16 MMAs) and GB/s."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
from nconv_b200.ops import P8, tc_pack_weights, conv_tc, TC_S1

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
ONLY = int(sys.argv[2]) if len(sys.argv) > 2 else None     # single shape (for ncu captures)
SHAPES = [(64, 0, 32, 480, 640), (32, 0, 32, 480, 640), (32, 32, 32, 480, 640), (64, 64, 64, 240, 320), (64, 0, 64, 240, 320),
          (64, 0, 32, 240, 320), (32, 0, 32, 240, 320), (64, 64, 64, 120, 160), (64, 0, 64, 120, 160),
          (64, 64, 64, 60, 80), (64, 0, 64, 60, 80)]
for CA, CB, CO, H, W in (SHAPES if ONLY is None else SHAPES[ONLY:ONLY + 1]):
    a = P8.pack(torch.randn(B, CA, H, W, device=dev))
    b = P8.pack(torch.randn(B, CB, H, W, device=dev)) if CB else None
    wp = tc_pack_weights(TC_S1, torch.randn(CO, CA + CB, 3, 3, device=dev) * 0.05, CA, CB)
    shift = torch.zeros(CO, device=dev)
    y = P8.empty(B, CO, H, W, dev)
    for _ in range(3):
        conv_tc(TC_S1, a, b, wp, CO, None, shift, True, out=y)
    torch.cuda.synchronize()
    n = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        conv_tc(TC_S1, a, b, wp, CO, None, shift, True, out=y)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    px = B * H * W
    fl = 2.0 * (CA + CB) * CO * 9 * px
    by = 4.0 * (CA + CB + CO) * px
    print(json.dumps({"ci": CA + CB, "co": CO, "H": H, "W": W, "B": B, "ms": round(ms, 4), "alg_TFLOPs": round(fl / ms / 1e9, 1),
                      "mma_TFLOPs": round(3 * fl / ms / 1e9, 1), "GBs": round(by / ms / 1e6, 1)}), flush=True)
    del a, b, y
