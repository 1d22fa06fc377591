#!/usr/bin/env python
"""Step time of the config-4 batch: eager, eager with per-launch profiling events, and CUDA-graph replay
(max over ranks when launched by torch.distributed.run: all ranks step at the same time)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch
import bench
from nconv_b200 import _lib
from nconv_b200.pipeline import GraphedInference
from nconv_b200.synthetic import setp2_inputs
import torch.distributed as dist
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
model = bench.build_model(dev)
res = [t.to(dev) for t in setp2_inputs(bench.FRAMES_PER_STREAM, seed=7)]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
out = {}
with torch.no_grad():
    out["eager_ms"] = timed(lambda: model(*res))
    _lib.profile_enable(True)
    out["eager_profiled_ms"] = timed(lambda: model(*res))
    _lib.profile_enable(False)
g = GraphedInference(model, res)
out["graph_ms"] = timed(lambda: g(*res))
out["world"] = world
if int(os.environ.get("RANK", "0")) == 0:
    print(json.dumps(out))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
