"""Diagnostic: per-tensor gradient distances of the CUDA SETP2 training step vs the reference golden."""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
from models.step1 import SETP1_NCONV
from models.step2 import SETP2_BP_TRAIN
from nconv_b200.synthetic import setp2_inputs
g = np.load(os.path.join(ROOT, "tests/golden/setp2_train.npz"))
H, W = (int(v) for v in g["hw"])
dev = torch.device("cuda:0")
td = tempfile.mkdtemp(); os.makedirs(td + "/checkpoints")
torch.manual_seed(3); s1 = SETP1_NCONV()
torch.save({"state_dict": {"module." + k: v for k, v in s1.state_dict().items()}}, td + "/checkpoints/golden_step1.pth.tar")
os.chdir(td)
torch.manual_seed(5)
m = SETP2_BP_TRAIN("golden_step1").to(dev).train()
rgb0, d0, rgb1, d1 = (t.to(dev) for t in setp2_inputs(1, seed=9, H=H, W=W))
gen = torch.Generator().manual_seed(77)
Gs = [torch.randn(2, 1, H // f, W // f, generator=gen).to(dev) for f in (8, 4, 2, 1)]
a, b = m(rgb0, d0, rgb1, d1)
depths = [torch.cat((x, y), 0) for x, y in zip(a, b)]
sum((d * G).sum() for d, G in zip(depths, Gs)).backward()
sub = lambda t: (lambda f: f if f.numel() <= 20000 else f[::13])(t.detach().cpu().reshape(-1))
for k in [str(n) for n in g["grad_names"]]:
    g64 = torch.from_numpy(g["g64_" + k]).double(); g32 = torch.from_numpy(g["g_" + k]).double()
    sc = g64.abs().max().clamp_min(1e-30)
    o = sub(dict(m.named_parameters())[k].grad).double()
    print("%-44s ours-64 %.2e  ref32-64 %.2e  ours-ref32 %.2e  |g| %.2e" % (k, (o - g64).abs().max() / sc, (g32 - g64).abs().max() / sc, (o - g32).abs().max() / sc, sc))
