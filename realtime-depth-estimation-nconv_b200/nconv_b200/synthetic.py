"""Seeded synthetic inputs for the BASELINE.json configs (there is no dataset on the box).

Value conventions follow the reference's loaders: depth in metres with 0 = missing
(dataset/nyuloader.py:83-124, dataset/kittiloader.py:68), RGB as raw 0..255 float BGR
(dataset/nyuloader.py:69).  Everything is generated on the CPU from an explicit
torch.Generator so the same tensors can be rebuilt anywhere from (config, seed).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

# (Ci, Co, K, pad, B, H, W) — per-layer parity cases: DNET's shapes, BASELINE config 2's sweep,
# ragged sizes (W % 4 != 0 -> scalar load path) and shapes served by the generic kernels.
NCONV_LAYER_CASES = [
    (1, 8, 5, 2, 2, 20, 36), (8, 8, 5, 2, 2, 20, 36), (16, 8, 3, 1, 2, 20, 36), (16, 8, 3, 0, 2, 20, 36),
    (8, 1, 1, 2, 2, 20, 36),
    (1, 2, 3, 1, 2, 24, 72), (2, 2, 3, 1, 2, 24, 72), (4, 4, 3, 1, 2, 24, 72), (16, 16, 3, 1, 2, 24, 72),
    (1, 2, 5, 2, 2, 24, 72), (2, 2, 5, 2, 2, 24, 72), (4, 4, 5, 2, 2, 24, 72), (16, 16, 5, 2, 2, 24, 72),
    (8, 8, 5, 2, 1, 17, 67), (16, 8, 3, 1, 1, 33, 130), (1, 8, 5, 2, 1, 5, 3),
    (3, 5, 3, 1, 1, 12, 20), (2, 3, 7, 3, 1, 12, 20), (4, 4, 3, 2, 1, 9, 11),
]


def _gen(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def nconv_layer_inputs(ci, co, k, pad, B, H, W, seed):
    """BASELINE config 2 distributions: data~U[0,10), conf~U[0,1)*Bernoulli(0.5) (exact zeros
    included), upstream grads ~N(0,1)."""
    g = _gen(seed)
    x = torch.rand(B, ci, H, W, generator=g) * 10
    c = torch.rand(B, ci, H, W, generator=g) * (torch.rand(B, ci, H, W, generator=g) < 0.5)
    HO, WO = H + 2 * pad - k + 1, W + 2 * pad - k + 1
    gy = torch.randn(B, co, HO, WO, generator=g)
    gc = torch.randn(B, co, HO, WO, generator=g)
    return x, c, gy, gc


def nconv_layer_case(idx: int):
    ci, co, k, pad, B, H, W = NCONV_LAYER_CASES[idx]
    return nconv_layer_inputs(ci, co, k, pad, B, H, W, 1000 + idx)


def sparse_depth_nyu(B: int, seed: int = 0, H: int = 480, W: int = 640, points: int = 500):
    """BASELINE config 1: `points` valid pixels per frame at random positions, U[1,10) metres."""
    g = _gen(seed)
    S = torch.zeros(B, 1, H, W)
    for b in range(B):
        idx = torch.randperm(H * W, generator=g)[:points]
        S[b].view(-1)[idx] = torch.rand(points, generator=g)[:idx.numel()] * 9 + 1    # a frame of fewer than `points` pixels: all valid
    return S


def sparse_depth_kitti(B: int, seed: int = 2000, H: int = 352, W: int = 1216, density: float = 0.05):
    """BASELINE config 3: Bernoulli(density) mask x U[1,80) metres."""
    g = _gen(seed)
    mask = torch.rand(B, 1, H, W, generator=g) < density
    return mask * (torch.rand(B, 1, H, W, generator=g) * 79 + 1)


def setp2_inputs(B: int, seed: int = 7, H: int = 480, W: int = 640):
    """BASELINE config 4: two streams of B frames: rgb~U[0,255) (raw BGR floats), NYU-style depth."""
    g = _gen(seed)
    rgb0 = torch.rand(B, 3, H, W, generator=g) * 255
    rgb1 = torch.rand(B, 3, H, W, generator=g) * 255
    d0 = sparse_depth_nyu(B, seed * 1000 + 1, H, W)
    d1 = sparse_depth_nyu(B, seed * 1000 + 2, H, W)
    return rgb0, d0, rgb1, d1


def loss_case(B: int, seed: int, H: int = 480, W: int = 640, scales=(1, 2, 4, 8), valid: float = 0.7):
    """Ground truth with holes (exact zeros, as in the NYU/KITTI gt maps) and one prediction per scale
    (train_step2.py:62-64: SETP2_BP_TRAIN returns 4 depth maps, full to 1/8 resolution)."""
    g = _gen(seed)
    gt = (torch.rand(B, 1, H, W, generator=g) * 9 + 1) * (torch.rand(B, 1, H, W, generator=g) < valid)
    depths = [torch.rand(B, 1, H // s, W // s, generator=g) * 10 for s in scales]
    return depths, gt


def randomise_bn_stats(model: torch.nn.Module, seed: int = 55) -> None:
    """Give every BatchNorm2d of the RGB-guided stack non-trivial running statistics and affine
    parameters so that BN folding is exercised (the unused NConv2d.bnorm modules are skipped)."""
    g = _gen(seed)
    for name, mod in model.named_modules():
        if isinstance(mod, torch.nn.BatchNorm2d) and not name.endswith("bnorm"):
            n = mod.num_features
            with torch.no_grad():
                mod.running_mean.copy_(torch.randn(n, generator=g))
                mod.running_var.copy_(torch.rand(n, generator=g) * 1.5 + 0.5)
                mod.weight.copy_(torch.rand(n, generator=g) + 0.5)
                mod.bias.copy_(torch.rand(n, generator=g) - 0.5)


def synthetic_state_dict(seed: int = 5):
    """A full SETP2_BP_EXPORT-shaped state_dict (step2.py:86-102 for the shapes) with
    random-init weights of the reference's distributions: NConv weights = Poisson-bump + U[0,1)
    (step1.py:159-169) passed once through softplus (what one training forward leaves stored,
    step1.py:190-207), nconv7 positive, conv weights Kaiming-uniform-like, BatchNorm with
    non-trivial running statistics.  Not RNG-identical to the reference constructors — it only
    has to be the same tensors for the CUDA arm and the CPU arm."""
    import numpy as np
    from scipy.stats import poisson
    g = torch.Generator().manual_seed(int(seed))
    sd = {}

    def nconv(name, ci, co, k):
        pmf = poisson(k / 2).pmf(np.arange(0, k))
        bump = torch.tensor(np.outer(pmf, pmf), dtype=torch.float32)
        w = bump.expand(co, ci, k, k) + torch.rand(co, ci, k, k, generator=g)
        sd["step1.d_net.%s.weight" % name] = F.softplus(w, beta=10).contiguous()
        sd["step1.d_net.%s.bias" % name] = torch.full((co,), 0.01)

    nconv("nconv1", 1, 8, 5)
    for n in ("nconv2", "nconv_down1", "nconv_down2", "nconv_down3"):
        nconv(n, 8, 8, 5)
    for n in ("nconv4", "nconv5", "nconv6"):
        nconv(n, 16, 8, 3)
    nconv("nconv7", 8, 1, 1)

    def conv(name, co, ci, k, bias):
        bound = (1.0 / (ci * k * k)) ** 0.5
        sd[name + ".weight"] = (torch.rand(co, ci, k, k, generator=g) * 2 - 1) * bound
        if bias:
            sd[name + ".bias"] = (torch.rand(co, generator=g) * 2 - 1) * bound

    def bn(name, c):
        sd[name + ".weight"] = torch.rand(c, generator=g) + 0.5
        sd[name + ".bias"] = torch.rand(c, generator=g) - 0.5
        sd[name + ".running_mean"] = torch.randn(c, generator=g) * 0.1
        sd[name + ".running_var"] = torch.rand(c, generator=g) * 1.5 + 0.5
        sd[name + ".num_batches_tracked"] = torch.tensor(1)

    for i, (ci, co) in enumerate(((3, 32), (32, 64), (64, 64), (64, 64))):
        p = "rgb_encoder%d" % i
        conv(p + ".encoder.0", co, ci, 3, True)
        bn(p + ".encoder.1", co)
        conv(p + ".downsample.0", co, ci, 1, False)

    def fusion(p, c, cout):
        conv(p + ".rgb_conv.conv", c, c, 3, True)
        conv(p + ".depth_conv.conv", c, 1, 3, True)
        conv(p + ".fuse_conv1.conv", c, 2 * c, 3, True)
        conv(p + ".fuse_conv2.conv", cout, c, 3, True)
        conv(p + ".fuse_conv3.conv", cout, cout, 3, True)

    fusion("fuse0.fuse", 64, 64)
    conv("fuse0.conv", 1, 64, 3, False)
    for name, c, cout in (("fuse1", 64, 64), ("fuse2", 64, 32), ("fuse3", 32, 32)):
        bound = (1.0 / ((c + 1) * 16)) ** 0.5
        sd[name + ".upcat.upf.conv.weight"] = (torch.rand(c + 1, c, 4, 4, generator=g) * 2 - 1) * bound
        bn(name + ".upcat.upf.bn", c)
        conv(name + ".upcat.conv.conv.conv", c, 2 * c, 3, False)
        bn(name + ".upcat.conv.conv.bn", c)
        fusion(name + ".fuse", c, cout)
        conv(name + ".conv", 1, cout, 3, False)
    return sd
