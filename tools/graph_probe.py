#!/usr/bin/env python
"""Step time of the config-4 batch: eager, eager with per-launch profiling events, and CUDA-graph replay."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
import bench
from nconv_b200 import _lib
from nconv_b200.pipeline import GraphedInference
from nconv_b200.synthetic import setp2_inputs
dev = torch.device("cuda:0")
model = bench.build_model(dev)
res = [t.to(dev) for t in setp2_inputs(bench.FRAMES_PER_STREAM, seed=7)]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
out = {}
with torch.no_grad():
    out["eager_ms"] = timed(lambda: model(*res))
    _lib.profile_enable(True)
    out["eager_profiled_ms"] = timed(lambda: model(*res))
    _lib.profile_enable(False)
g = GraphedInference(model, res)
out["graph_ms"] = timed(lambda: g(*res))
print(json.dumps(out))
