#!/usr/bin/env python
"""Where does the end-to-end step time go?  Times, on one GPU: raw pinned H2D / D2H bandwidth of one config-4 batch,
the resident step, and StreamedInference with depth 2 / 3 and with / without a consumer."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
import bench
from nconv_b200.pipeline import StreamedInference
from nconv_b200.synthetic import setp2_inputs

dev = torch.device("cuda:0")
model = bench.build_model(dev)
host = [t.pin_memory() for t in setp2_inputs(bench.FRAMES_PER_STREAM, seed=7)]
res = [t.to(dev) for t in host]
out = {}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(fn, n=10):
    fn(); torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


nb = sum(t.numel() * 4 for t in host)
out["h2d_ms"] = timed(lambda: [d.copy_(h, non_blocking=True) for d, h in zip(res, host)])
out["h2d_gbs"] = nb / out["h2d_ms"] / 1e6
with torch.no_grad():
    o = model(*res)
ho = [torch.empty(t.shape).pin_memory() for t in o]
out["d2h_ms"] = timed(lambda: [h.copy_(d, non_blocking=True) for d, h in zip(o, ho)])
with torch.no_grad():
    out["resident_ms"] = timed(lambda: model(*res))
for depth in (2, 3):
    for name, consume in (("clone_sample", lambda i, outs: [x[..., ::16, ::16].clone() for x in outs]), ("none", lambda i, outs: None)):
        s = StreamedInference(model, dev, depth=depth)
        s.run(host for _ in range(3))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        s.run((host for _ in range(10)), consume=consume)
        torch.cuda.synchronize()
        out["streamed_depth%d_%s_ms" % (depth, name)] = (time.perf_counter() - t0) * 100
# H2D of one batch WHILE the resident step runs on the compute stream (copy engine vs kernels)
side = torch.cuda.Stream(dev)
c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.no_grad():
    for _ in range(3):
        model(*res)
    with torch.cuda.stream(side):
        c0.record(side)
        for d, h in zip([torch.empty_like(r) for r in res], host):
            d.copy_(h, non_blocking=True)
        c1.record(side)
    for _ in range(3):
        model(*res)
torch.cuda.synchronize()
out["h2d_ms_under_compute"] = c0.elapsed_time(c1)
# host time to ENQUEUE one step (queue drained first; no synchronisation inside)
torch.cuda.synchronize()
with torch.no_grad():
    t0 = time.perf_counter()
    for _ in range(5):
        model(*res)
    out["host_enqueue_ms_per_step"] = (time.perf_counter() - t0) * 200
torch.cuda.synchronize()
out["cpus"] = len(os.sched_getaffinity(0))
try:
    import subprocess
    out["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout[-700:]
except Exception as e:       # noqa: BLE001
    out["topo"] = str(e)
print(json.dumps(out))
