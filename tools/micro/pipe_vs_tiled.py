#!/usr/bin/env python
"""SETP1 forward with nconv_pipe (default) against nconv_tiled (NCONV_B200_NO_PIPE=1 in a child process): the two
kernels share arithmetic and accumulation order, so the outputs must be bit-identical."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import torch


def run(path):
    from models.step1 import SETP1_NCONV
    from nconv_b200.synthetic import sparse_depth_kitti, sparse_depth_nyu
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = SETP1_NCONV().to(dev)
    m.train()
    with torch.no_grad():
        m(torch.zeros(1, 1, 16, 16, device=dev))
    m.eval()
    outs = []
    with torch.no_grad():
        for S in (sparse_depth_nyu(3, seed=1), sparse_depth_kitti(2, seed=2), sparse_depth_nyu(2, seed=3, H=72, W=136, points=300)):
            y = m(S.to(dev))
            outs += [t.cpu() for t in (y if isinstance(y, (tuple, list)) else [y])]
    torch.save(outs, path)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        run(sys.argv[1])
    else:
        a, b = "/tmp/pipe_a.pt", "/tmp/pipe_b.pt"
        subprocess.check_call([sys.executable, __file__, a])
        env = dict(os.environ, NCONV_B200_NO_PIPE="1")
        subprocess.check_call([sys.executable, __file__, b], env=env)
        A, B = torch.load(a), torch.load(b)
        same = all(torch.equal(x, y) for x, y in zip(A, B))
        print("tensors", len(A), "bit-identical", same, "max diff", max(float((x - y).abs().max()) for x, y in zip(A, B)))
        sys.exit(0 if same else 1)
