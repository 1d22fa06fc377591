#!/usr/bin/env python
"""BASELINE.json config 2: single NConv2d layer microbench sweep on one B200.

k in {3,5} (pad k//2), (Ci,Co) in {(1,2),(2,2),(4,4),(16,16)} plus DNET's own layer shapes, forward
and backward timed separately at 480x640, batch 16, through the C ABI
(torch.ops.nconv_b200.nconv2d / nconv2d_backward).  Working set per launch >= 118 MB and the 126 MB
L2 is flushed between timed launches by writing a 256 MB buffer, so every launch starts cold.

Algorithmic bytes: fwd 4*B*(2*Ci*H*W + 2*Co*Ho*Wo); bwd (consumes saved y, cout
instead of recomputing D, N) 4*B*(4*Co*Ho*Wo [gy,gcout,y,cout] + 2*Ci*H*W [x,c] + 2*Ci*H*W
[gx,gc]).  Prints one JSON object per case and a final summary line.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

SWEEP = [(1, 2), (2, 2), (4, 4), (16, 16)]
DNET_SHAPES = [(1, 8, 5, 2), (8, 8, 5, 2), (16, 8, 3, 1)]


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--height", type=int, default=480)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-bwd", action="store_true")
    ap.add_argument("--flush", choices=["write", "read"], default="write")
    ap.add_argument("--sleep-cycles", type=int, default=400000)
    ap.add_argument("--breakdown", action="store_true", help="per-kernel times of one cold backward")
    ap.add_argument("--only", type=str, default="", help="comma list of ci-co-k cases")
    args = ap.parse_args()
    from nconv_b200 import ops  # noqa: F401
    K = torch.ops.nconv_b200
    dev = torch.device("cuda:0")
    hbm, src = peaks()
    B, H, W = args.batch, args.height, args.width
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    flush_f = flush.view(torch.float32)

    def do_flush():
        # "write": L2 is left full of DIRTY lines, whose write-back the timed kernel then pays for;
        # "read": the written buffer is read back, so the timed kernel starts on a cold, CLEAN L2.
        flush.fill_(1)
        if args.flush == "read":
            flush_f.sum()
    cases = [(ci, co, k, k // 2) for k in (3, 5) for ci, co in SWEEP] + DNET_SHAPES
    rows = []
    only = set(args.only.split(",")) if args.only else None
    from nconv_b200 import _lib
    for idx, (ci, co, k, pad) in enumerate(cases):
        if only and "%d-%d-%d" % (ci, co, k) not in only:
            continue
        g = torch.Generator(device="cpu"); g.manual_seed(1000 + idx)
        x = (torch.rand(B, ci, H, W, generator=g) * 10).to(dev)
        c = (torch.rand(B, ci, H, W, generator=g) * (torch.rand(B, ci, H, W, generator=g) < 0.5)).to(dev)
        w = torch.nn.functional.softplus(torch.rand(co, ci, k, k, generator=g), beta=10).to(dev)
        b = torch.full((co,), 0.01, device=dev)
        HO, WO = H + 2 * pad - k + 1, W + 2 * pad - k + 1
        gy = torch.randn(B, co, HO, WO, generator=g).to(dev)
        gc = torch.randn(B, co, HO, WO, generator=g).to(dev)

        def timed(fn):
            for _ in range(3):
                fn()
            ts = []
            for _ in range(args.iters):
                do_flush()
                # a device-side spin lets the host finish enqueueing fn()'s launches before the
                # timed region starts on the GPU (several launches + allocator calls per backward)
                torch.cuda._sleep(args.sleep_cycles)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ts.sort()
            return ts[len(ts) // 2]

        y, co_ = K.nconv2d(x, c, w, b, pad, pad, 1e-7)
        ms_f = timed(lambda: K.nconv2d(x, c, w, b, pad, pad, 1e-7))
        bytes_f = 4.0 * B * (2 * ci * H * W + 2 * co * HO * WO)
        flops_f = 4.0 * ci * co * k * k * B * HO * WO
        row = {"k": k, "ci": ci, "co": co, "B": B, "H": H, "W": W,
               "fwd_ms": ms_f, "fwd_GBs": bytes_f / ms_f / 1e6, "fwd_hbm_frac": bytes_f / ms_f / 1e6 / hbm,
               "fwd_TFLOPs": flops_f / ms_f / 1e9}
        if not args.no_bwd:
            ms_b = timed(lambda: K.nconv2d_backward(gy, gc, x, c, w, b, y, co_, pad, pad, 1e-7, True, True, True))
            bytes_b = 4.0 * B * (4 * co * HO * WO + 4 * ci * H * W)
            row.update({"bwd_ms": ms_b, "bwd_GBs": bytes_b / ms_b / 1e6, "bwd_hbm_frac": bytes_b / ms_b / 1e6 / hbm,
                        "bwd_TFLOPs": 2 * flops_f / ms_b / 1e9})
            if args.breakdown:
                do_flush(); torch.cuda.synchronize()
                _lib.profile_enable(True)
                K.nconv2d_backward(gy, gc, x, c, w, b, y, co_, pad, pad, 1e-7, True, True, True)
                torch.cuda.synchronize()
                row["bwd_kernels"] = [(l, round(ms, 4)) for l, ms, _, _ in _lib.profile_records()]
                _lib.profile_enable(False)
        rows.append(row)
        print(json.dumps({k_: (round(v, 4) if isinstance(v, float) else v) for k_, v in row.items()}), flush=True)
        del x, c, gy, gc, y, co_
    print(json.dumps({"summary": "nconv2d config-2 sweep", "hbm_peak_GBs": hbm, "peak_source": src,
                      "best_fwd_frac": max(r["fwd_hbm_frac"] for r in rows),
                      "l2": "256 MB flush (%s) between timed launches" % args.flush}))


if __name__ == "__main__":
    main()
