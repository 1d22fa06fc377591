#!/usr/bin/env python
"""Calibration: what a plain device copy reaches at the config-2 launch sizes (cold L2, one launch
timed at a time, same protocol as tools/nconv2d_sweep.py).  read+write bytes / time."""
import json, sys, torch
MODE = sys.argv[1] if len(sys.argv) > 1 else "write"   # "read": read the flush buffer back -> clean L2
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = []
for total_mb in (118, 157, 315, 1258, 4096):
    n = total_mb * (1 << 20) // 2 // 4
    a = torch.rand(n, device=dev); b = torch.empty_like(a)
    ts = []
    for it in range(25):
        flush.fill_(it & 1)
        if MODE == "read":
            flush.view(torch.float32).sum()
        torch.cuda._sleep(400000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
        if it >= 5: ts.append(e0.elapsed_time(e1))
    ts.sort(); ms = ts[len(ts) // 2]
    out.append({"flush": MODE, "traffic_MB": total_mb, "ms": round(ms, 4), "GBs": round(2 * n * 4 / ms / 1e6, 1)})
    print(json.dumps(out[-1]))
