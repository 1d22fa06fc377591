"""tcgen05 NConv2d layers (csrc/nconv_tc.cu) and the N8 layout against the CPU oracle (`O.nconv2d`, step1.py:116-149,
max-pools :62-74, nearest upsample :78-89, nconv7 + crop tail :92-94).

Tolerance: y / cout max-rel <= 1e-4 with |ref| floored at 1e-3 (BASELINE north_star); the split-bf16 contraction
keeps 16 mantissa bits per operand (SURVEY 7.3 measured 1.6e-5), so the achieved figures are reported too.  Pools are
compared with F.max_pool2d of the ORACLE's y / cout at the same tolerance: the pooling itself is an exact selection
(bit-exact on identical pool inputs — checked separately by feeding the kernel's own fp32 outputs to F.max_pool2d)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import REL_TOL, report
from oracle import nconv_oracle as O

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


def _weight(co, ci, k, gen):
    from scipy.stats import poisson
    pmf = poisson(k / 2).pmf(np.arange(0, k))
    bump = torch.tensor(np.outer(pmf, pmf), dtype=torch.float32)
    return O.softplus_beta10(bump.expand(co, ci, k, k) + torch.rand(co, ci, k, k, generator=gen)).contiguous()


def _pair(B, H, W, gen, sparse=0.5):
    x = torch.rand(B, 8, H, W, generator=gen) * 10
    c = torch.rand(B, 8, H, W, generator=gen) * (torch.rand(B, 8, H, W, generator=gen) < sparse)
    return x, c


@pytest.mark.parametrize("shape", [(2, 5, 7), (1, 33, 70)])
def test_n8_roundtrip_and_ring(shape):
    from nconv_b200.ops import N8
    B, H, W = shape
    g = torch.Generator().manual_seed(H)
    x, c = _pair(B, H, W, g)
    t = N8.pack(x.to(dev()), c.to(dev()))
    xc, cc = t.unpack()
    assert O.max_rel_err(xc, x * c, floor=1e-6) <= 2.0 ** -16
    assert O.max_rel_err(cc, c, floor=1e-6) <= 2.0 ** -16
    raw = t.buf[: B * 4 * (H + 4) * (W + 4) * 16].view(torch.bfloat16).view(B, 4, H + 4, W + 4, 8).float()
    assert float(raw[:, :, :2].abs().max()) == 0 and float(raw[:, :, -2:].abs().max()) == 0
    assert float(raw[:, :, :, :2].abs().max()) == 0 and float(raw[:, :, :, -2:].abs().max()) == 0


@pytest.mark.parametrize("shape", [(2, 20, 36), (1, 45, 70), (2, 64, 96), (1, 120, 160)])
def test_nconv1_n8_vs_oracle(shape):
    from nconv_b200.ops import nconv1_n8
    B, H, W = shape
    g = torch.Generator().manual_seed(B + H)
    S = torch.rand(B, 1, H, W, generator=g) * 9 * (torch.rand(B, 1, H, W, generator=g) < 0.05)
    S[0, 0, 0, 0] = 0.005                                   # below the threshold: data kept, confidence 0
    w = _weight(8, 1, 5, g); b = torch.full((8,), 0.01)
    y, cout = O.nconv2d(S, (S > 0.01).float(), w, b, (2, 2))
    xc, cc = nconv1_n8(S.to(dev()), w.to(dev()), b.to(dev())).unpack()
    assert O.max_rel_err(cc, cout) <= REL_TOL
    assert O.max_rel_err(xc, y * cout) <= REL_TOL


@pytest.mark.parametrize("shape", [(2, 20, 36), (1, 45, 70), (2, 64, 96), (1, 121, 163), (3, 8, 28)])
def test_nconv_tc_k5_vs_oracle(shape):
    """8->8 5x5 (nconv2 / nconv_down1..3): every output form."""
    from nconv_b200.ops import N8, nconv_tc
    B, H, W = shape
    g = torch.Generator().manual_seed(H * 7 + W)
    x, c = _pair(B, H, W, g)
    w = _weight(8, 8, 5, g); b = torch.rand(8, generator=g) * 0.1
    y, cout = O.nconv2d(x, c, w, b, (2, 2))
    out = nconv_tc(N8.pack(x.to(dev()), c.to(dev())), None, w.to(dev()), b.to(dev()), want_n8=True, want_pool=True,
                   want_up=True, want_f32=True)
    torch.cuda.synchronize()
    e_y, e_c = O.max_rel_err(out["y"], y), O.max_rel_err(out["cout"], cout)
    report("nconv_tc_k5_%dx%dx%d" % shape, y_max_rel=e_y, cout_max_rel=e_c)
    assert e_y <= REL_TOL and e_c <= REL_TOL
    xc, cc = out["n8"].unpack()
    assert O.max_rel_err(cc, cout) <= REL_TOL and O.max_rel_err(xc, y * cout) <= REL_TOL
    # pools: exact selection on the kernel's own outputs, and within tolerance of the oracle's pools
    yp, cp = F.max_pool2d(out["y"], 2, 2), F.max_pool2d(out["cout"], 2, 2)
    pxc, pc = out["pool"].unpack()
    assert O.max_rel_err(pc, cp, floor=1e-6) <= 2.0 ** -16 and O.max_rel_err(pxc, yp * cp, floor=1e-6) <= 2.0 ** -16
    assert O.max_rel_err(pc, F.max_pool2d(cout, 2, 2)) <= REL_TOL
    # upsampled copy == nearest 2x of the same-size N8
    uxc, uc = out["up"].unpack()
    assert torch.equal(uxc, F.interpolate(xc, scale_factor=2, mode="nearest")) and torch.equal(uc, F.interpolate(cc, scale_factor=2, mode="nearest"))


@pytest.mark.parametrize("a_first", [True, False])
@pytest.mark.parametrize("shape", [(2, 20, 36), (1, 45, 70), (2, 64, 96), (1, 120, 160)])
def test_nconv_tc_k3_two_sources_vs_oracle(shape, a_first):
    """[8|8]->8 3x3 (nconv4 / nconv5: skip first; nconv6: upsampled first, step1.py:80,85,90)."""
    from nconv_b200.ops import N8, nconv_tc
    B, H, W = shape
    g = torch.Generator().manual_seed(H * 3 + W + int(a_first))
    xa, ca = _pair(B, H, W, g)
    xb, cb = _pair(B, H, W, g, sparse=0.8)
    w = _weight(8, 16, 3, g); b = torch.rand(8, generator=g) * 0.1
    cat = (lambda p, q: torch.cat((p, q), 1)) if a_first else (lambda p, q: torch.cat((q, p), 1))
    y, cout = O.nconv2d(cat(xa, xb), cat(ca, cb), w, b, (1, 1))
    out = nconv_tc(N8.pack(xa.to(dev()), ca.to(dev())), N8.pack(xb.to(dev()), cb.to(dev())), w.to(dev()), b.to(dev()),
                   a_first=a_first, want_f32=True, want_up=True)
    e_y, e_c = O.max_rel_err(out["y"], y), O.max_rel_err(out["cout"], cout)
    report("nconv_tc_k3_%dx%dx%d_afirst%d" % (shape + (int(a_first),)), y_max_rel=e_y, cout_max_rel=e_c)
    assert e_y <= REL_TOL and e_c <= REL_TOL
    uxc, uc = out["up"].unpack()
    assert O.max_rel_err(uc, F.interpolate(cout, scale_factor=2, mode="nearest")) <= REL_TOL


@pytest.mark.parametrize("shape", [(2, 20, 36), (1, 64, 96)])
def test_nconv_tc_tail_vs_oracle(shape):
    """nconv6 (valid 3x3) -> nconv7 (1x1, pad 2) -> crop [1:H+1, 1:W+1] (step1.py:90-94): the border ring is exactly
    nconv7's bias."""
    from nconv_b200.ops import N8, nconv_tc
    B, H, W = shape
    g = torch.Generator().manual_seed(H + W)
    xa, ca = _pair(B, H, W, g)
    xb, cb = _pair(B, H, W, g, sparse=0.9)
    w6 = _weight(8, 16, 3, g); b6 = torch.full((8,), 0.01)
    w7 = O.softplus_beta10(torch.rand(1, 8, 1, 1, generator=g)); b7 = torch.tensor([0.01])
    x6, c6 = O.nconv2d(torch.cat((xb, xa), 1), torch.cat((cb, ca), 1), w6, b6, (0, 0))
    x7, _ = O.nconv2d(x6, c6, w7, b7, (2, 2))
    ref = x7[:, :, 1:H + 1, 1:W + 1]
    out = nconv_tc(N8.pack(xa.to(dev()), ca.to(dev())), N8.pack(xb.to(dev()), cb.to(dev())), w6.to(dev()), b6.to(dev()),
                   a_first=False, tail=(w7.to(dev()), b7.to(dev())))["tail"]
    assert O.max_rel_err(out, ref) <= REL_TOL
    o = out.cpu()
    for ring in (o[:, 0, 0], o[:, 0, -1], o[:, 0, :, 0], o[:, 0, :, -1]):
        assert torch.all(ring == 0.01)
