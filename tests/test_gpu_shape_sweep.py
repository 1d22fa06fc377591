"""Seeded random shape sweep of the tcgen05 kernels.  Tile geometry (rows per tile, tile width, M-tiles per TMEM buffer,
pipeline depth, which kernel form runs a layer) is planned per launch from (B, H, W, Ci, Co), so the fixed shapes of
test_gpu_parity.py leave most plans unvisited: here ~150 drawn shapes per run, all small enough for a double-precision
torch convolution on the same GPU as the checker (tests only; cuDNN never runs on the product path).

Tolerance as in test_gpu_parity.py: max|a-b| <= 2e-5 * max|ref| for the split-bf16 kernels against fp64, zero ring of
every P8 output intact."""
import random

import pytest
import torch
import torch.nn.functional as F

from conftest import report

pytestmark = pytest.mark.gpu

TC_TOL = 2e-5


def dev():
    return torch.device("cuda:0")


def _err(out, ref):
    return float((out.double() - ref).abs().max() / ref.abs().max().clamp_min(1e-30))


def _ring_ok(y):
    B, C, H, W = y.B, y.C, y.H, y.W
    raw = y.buf[: 2 * ((C + 7) // 8) * B * (H + 2) * (W + 2) * 16].view(torch.bfloat16).view(B, (C + 7) // 8, 2, H + 2, W + 2, 8)
    return (bool((raw[:, :, :, 0] == 0).all()) and bool((raw[:, :, :, -1] == 0).all()) and
            bool((raw[:, :, :, :, 0] == 0).all()) and bool((raw[:, :, :, :, -1] == 0).all()))


def _draw_hw(rng, lo=3, hi=150):
    # mix of tiny, ragged and wide frames
    kind = rng.random()
    if kind < 0.25:
        return rng.randint(lo, 12), rng.randint(lo, hi)
    if kind < 0.5:
        return rng.randint(lo, hi), rng.randint(lo, 12)
    return rng.randint(8, 70), rng.randint(8, hi)


@pytest.mark.parametrize("seed", range(12))
def test_conv3x3_s1_random_shapes(seed):
    """3x3 / stride 1: pixels-on-M (Co = 64, Co = 32 with 32 input channels, stacked accumulators), pixels-on-N
    (Co = 32 with >= 64 input channels), one and two sources, P8 and fp32 epilogues."""
    from nconv_b200 import ops
    from nconv_b200.ops import P8, TC_S1
    rng = random.Random(1000 + seed)
    torch.manual_seed(1000 + seed)
    worst = 0.0
    for k in range(10):
        CO = rng.choice((32, 64))
        CA = rng.choice((16, 32, 48, 64, 96, 128))
        CB = rng.choice((0, 0, 16, 32, 64))
        B = rng.randint(1, 3)
        H, W = _draw_hw(rng)
        relu, affine = rng.random() < 0.5, rng.random() < 0.5
        a = torch.randn(B, CA, H, W, device=dev()) * 3
        b = torch.randn(B, CB, H, W, device=dev()) * 3 if CB else None
        w = torch.randn(CO, CA + CB, 3, 3, device=dev()) / (3 * (CA + CB) ** 0.5)
        scale = torch.rand(CO, device=dev()) + 0.5 if affine else None
        shift = torch.randn(CO, device=dev())
        x = a if b is None else torch.cat((a, b), 1)
        ref = F.conv2d(x.double(), w.double(), padding=1)
        if scale is not None:
            ref = ref * scale.double().view(1, -1, 1, 1)
        ref = ref + shift.double().view(1, -1, 1, 1)
        if relu:
            ref = ref.relu()
        pa, pb = P8.pack(a), (P8.pack(b) if b is not None else None)
        wp = ops.tc_pack_weights(TC_S1, w, CA, CB)
        try:
            y = ops.conv_tc(TC_S1, pa, pb, wp, CO, scale, shift, relu)
        except RuntimeError as ex:
            raise AssertionError(("s1 launch", seed, k, (CA, CB, CO, B, H, W), str(ex)))
        e = _err(y.unpack(), ref)
        worst = max(worst, e)
        assert e <= TC_TOL, ("s1", seed, k, (CA, CB, CO, B, H, W), e)
        assert _ring_ok(y), ("s1 ring", seed, k, (CA, CB, CO, B, H, W))
        if CB == 0 and scale is None:
            out = torch.full((B, CO + 8, H, W), 9.0, device=dev())
            if ops.conv_tc_f32(pa, wp, CO, out, 8, shift=shift, relu=relu):
                assert _err(out[:, 8:], ref) <= TC_TOL, ("s1 f32", seed, k, (CA, CO, B, H, W))
                assert bool((out[:, :8] == 9).all())
    report("shape_sweep_conv3x3_s1_seed%d" % seed, worst_rel=worst, shapes=10)


@pytest.mark.parametrize("seed", range(6))
def test_conv3x3_s2_and_transposed_random_shapes(seed):
    from nconv_b200 import ops
    from nconv_b200.ops import P8, TC_S2, TC_T
    rng = random.Random(2000 + seed)
    torch.manual_seed(2000 + seed)
    worst = 0.0
    for k in range(8):
        # ---- 3x3 / stride 2 (+ 1x1 / stride 2 residual), even sizes
        CI, CO = rng.choice((32, 64)), 64
        B = rng.randint(1, 2)
        H, W = 2 * rng.randint(2, 40), 2 * rng.randint(2, 70)
        res = rng.random() < 0.6
        x = torch.randn(B, CI, H, W, device=dev()) * 3
        w = torch.randn(CO, CI, 3, 3, device=dev()) / (3 * CI ** 0.5)
        w1 = torch.randn(CO, CI, 1, 1, device=dev()) / CI ** 0.5 if res else None
        scale, shift = torch.rand(CO, device=dev()) + 0.5, torch.randn(CO, device=dev())
        ref = (F.conv2d(x.double(), w.double(), stride=2, padding=1) * scale.double().view(1, -1, 1, 1)
               + shift.double().view(1, -1, 1, 1)).relu()
        if res:
            ref = ref + F.conv2d(x.double(), w1.double(), stride=2)
        px = P8.pack(x, with_q=True)
        wp = ops.tc_pack_weights(TC_S2, w, CI, 0, w1=w1)
        y = ops.conv_tc(TC_S2, px, None, wp, CO, scale, shift, True, has_1x1=res)
        e = _err(y.unpack(), ref)
        worst = max(worst, e)
        assert e <= TC_TOL, ("s2", seed, k, (CI, CO, B, H, W, res), e)
        assert _ring_ok(y)
        # ---- ConvTranspose2d(4, 2, 1) on cat[depth(1), features] or on features alone
        lone = rng.random() < 0.5
        CF, CO = rng.choice(((64, 64), (32, 32), (64, 32), (32, 64)))
        B = rng.randint(1, 2)
        H, W = rng.randint(2, 40), rng.randint(2, 70)
        f = torch.randn(B, CF, H, W, device=dev()) * 3
        d = torch.randn(B, 1, H, W, device=dev()) * 5 if lone else None
        CI = CF + (1 if lone else 0)
        w = torch.randn(CI, CO, 4, 4, device=dev()) / (2 * CI ** 0.5)
        scale, shift = torch.rand(CO, device=dev()) + 0.5, torch.randn(CO, device=dev())
        xin = torch.cat((d, f), 1) if lone else f
        ref = (F.conv_transpose2d(xin.double(), w.double(), stride=2, padding=1) * scale.double().view(1, -1, 1, 1)
               + shift.double().view(1, -1, 1, 1)).relu()
        pf = P8.pack(f)
        if lone:
            y = ops.conv_tc(TC_T, P8.pack(d), pf, ops.tc_pack_weights(TC_T, w, 1, CF), CO, scale, shift, True)
        else:
            y = ops.conv_tc(TC_T, pf, None, ops.tc_pack_weights(TC_T, w, CF, 0), CO, scale, shift, True)
        e = _err(y.unpack(), ref)
        worst = max(worst, e)
        assert e <= TC_TOL, ("T", seed, k, (lone, CF, CO, B, H, W), e)
        assert _ring_ok(y)
    report("shape_sweep_conv_s2_T_seed%d" % seed, worst_rel=worst, shapes=16)


@pytest.mark.parametrize("seed", range(5))
def test_conv_wgrad_tc_random_shapes(seed):
    """3x3 weight gradient on tcgen05 (MN-major operands, taps split over CTAs, per-CTA partials) against fp64."""
    from nconv_b200 import ops
    from nconv_b200.ops import P8
    rng = random.Random(3000 + seed)
    torch.manual_seed(3000 + seed)
    worst = 0.0
    for k in range(8):
        CI = rng.choice((32, 64, 128))
        CO = rng.choice((32, 64))
        B = rng.randint(1, 3)
        H, W = _draw_hw(rng, lo=3, hi=110)
        x = torch.randn(B, CI, H, W, device=dev())
        g = torch.randn(B, CO, H, W, device=dev())
        xd = x.double().requires_grad_(False)
        wd = torch.zeros(CO, CI, 3, 3, device=dev(), dtype=torch.double, requires_grad=True)
        (F.conv2d(xd, wd, padding=1) * g.double()).sum().backward()
        gw = ops._conv_wgrad_tc(P8.pack(x), P8.pack(g))
        e = _err(gw, wd.grad)
        worst = max(worst, e)
        assert e <= TC_TOL, ("wgrad", seed, k, (CI, CO, B, H, W), e)
    report("shape_sweep_wgrad_tc_seed%d" % seed, worst_rel=worst, shapes=8)


def _nconv_ref(x, c, w, b, pad, eps=1e-7):
    """NConv2d.forward (step1.py:116-149) in fp64 on the GPU: the checker of this file."""
    x, c, w, b = x.double(), c.double(), w.double(), b.double()
    denom = F.conv2d(c, w, padding=pad)
    y = F.conv2d(x * c, w, padding=pad) / (denom + eps) + b.view(1, -1, 1, 1)
    return y, denom / w.sum((1, 2, 3)).view(1, -1, 1, 1)


def _rel_floor(a, ref, floor=1e-3):
    return float(((a.double() - ref).abs() / ref.abs().clamp_min(floor)).max())


@pytest.mark.parametrize("seed", range(5))
def test_nconv_tc_random_shapes(seed):
    """8->8 5x5 and [8|8]->8 3x3 NConv2d on tcgen05 at drawn sizes (tile height, ragged last M-tile, frames narrower than
    one staged row): y / cout within 1e-4 at a 1e-3 floor of the fp64 formula."""
    from nconv_b200.ops import N8, nconv_tc
    rng = random.Random(4000 + seed)
    torch.manual_seed(4000 + seed)
    worst = 0.0
    for k in range(6):
        B = rng.randint(1, 3)
        H, W = rng.randint(2, 90), rng.randint(2, 140)
        def pair(sparse):
            x = torch.rand(B, 8, H, W, device=dev()) * 10
            c = torch.rand(B, 8, H, W, device=dev()) * (torch.rand(B, 8, H, W, device=dev()) < sparse)
            return x, c
        xa, ca = pair(0.5)
        w5 = F.softplus(torch.rand(8, 8, 5, 5, device=dev()), beta=10); b = torch.rand(8, device=dev()) * 0.1
        y, cout = _nconv_ref(xa, ca, w5, b, 2)
        out = nconv_tc(N8.pack(xa, ca), None, w5, b, want_f32=True, want_n8=True)
        e = max(_rel_floor(out["y"], y), _rel_floor(out["cout"], cout))
        worst = max(worst, e)
        assert e <= 1e-4, ("k5", seed, k, (B, H, W), e)
        xb, cb = pair(0.8)
        a_first = rng.random() < 0.5
        w3 = F.softplus(torch.rand(8, 16, 3, 3, device=dev()), beta=10)
        cat = (lambda p, q: torch.cat((p, q), 1)) if a_first else (lambda p, q: torch.cat((q, p), 1))
        y, cout = _nconv_ref(cat(xa, xb), cat(ca, cb), w3, b, 1)
        out = nconv_tc(N8.pack(xa, ca), N8.pack(xb, cb), w3, b, a_first=a_first, want_f32=True)
        e = max(_rel_floor(out["y"], y), _rel_floor(out["cout"], cout))
        worst = max(worst, e)
        assert e <= 1e-4, ("k3", seed, k, (B, H, W, a_first), e)
    report("shape_sweep_nconv_tc_seed%d" % seed, worst_rel_floor1e3=worst, shapes=12)


@pytest.mark.parametrize("seed", range(3))
def test_setp1_pipeline_random_sizes_tc_vs_fp32(seed):
    """The whole SETP1 forward at drawn frame sizes that are eligible for the tcgen05 pipeline (multiples of 8, W >= 224):
    the tensor-core pipeline against the fp32 CUDA-core pipeline (itself held to the reference goldens) within 1e-4 at
    the 1e-3 floor."""
    from models.step1 import SETP1_NCONV
    rng = random.Random(5000 + seed)
    torch.manual_seed(5000 + seed)
    m = SETP1_NCONV().to(dev())
    m.train()
    with torch.no_grad():
        m(torch.zeros(1, 1, 16, 16, device=dev()))       # the hook applies softplus once (training-mode forward)
    m.eval()
    worst = 0.0
    for k in range(4):
        B = rng.randint(1, 4)
        H, W = 8 * rng.randint(4, 40), 8 * rng.randint(28, 90)
        S = (torch.rand(B, 1, H, W, device=dev()) * 9 + 1) * (torch.rand(B, 1, H, W, device=dev()) < rng.choice((0.002, 0.05, 0.3)))
        with torch.no_grad():
            tc = m(S)
            m.d_net.fp32_only = True
            ref = m(S)
            m.d_net.fp32_only = False
        assert tc.shape == ref.shape == (B, 1, H, W)
        e = _rel_floor(tc, ref.double())
        worst = max(worst, e)
        assert e <= 1e-4, ("setp1", seed, k, (B, H, W), e)
    report("shape_sweep_setp1_tc_vs_fp32_seed%d" % seed, worst_rel_floor1e3=worst, shapes=4)


def _rel_scale(a, ref):
    ref = ref.detach().double()
    return float((a.detach().double() - ref).abs().max() / ref.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("seed", range(5))
def test_training_kernels_random_shapes(seed):
    """ops.conv_train / convT_train / bn_train (forward, data / weight / bias gradients; fp32 CUDA-core kernels plus the
    tcgen05 data and weight gradients where the shape allows) through autograd at drawn shapes, against torch autograd in
    fp64 on the same GPU.  Tolerances of test_gpu_parity.py: forward 1e-5 (2e-5 on tcgen05), gradients 2e-5 (3e-5 for a
    tcgen05 data gradient) of each tensor's scale."""
    from nconv_b200 import ops
    rng = random.Random(6000 + seed)
    torch.manual_seed(6000 + seed)
    worst = 0.0
    for k in range(6):
        # ---- Conv2d K in {1, 3}, stride in {1, 2}, optional bias / ReLU
        CI = rng.choice((1, 3, 16, 32, 33, 64, 65, 128))
        CO = rng.choice((1, 32, 64)) if CI >= 16 else rng.choice((32, 64))
        K, stride = rng.choice(((3, 1), (3, 1), (3, 2), (1, 1), (1, 2)))
        if CO == 1:
            stride = 1           # the stack's single-channel heads are stride 1; stride 2 needs Co % 16 == 0 (UNSUPPORTED otherwise)
        relu, bias = rng.random() < 0.5, rng.random() < 0.6
        B = rng.randint(1, 3)
        H, W = rng.randint(4, 40), rng.randint(4, 90)
        x = torch.rand(B, CI, H, W, device=dev()) * 4 - 1
        w = (torch.rand(CO, CI, K, K, device=dev()) - 0.5) * 0.2
        bb = torch.rand(CO, device=dev()) - 0.5 if bias else None
        xr, wr = x.double().requires_grad_(True), w.double().requires_grad_(True)
        br = bb.double().requires_grad_(True) if bias else None
        yr = F.conv2d(xr, wr, br, stride=stride, padding=K // 2)
        if relu:
            yr = F.relu(yr)
        G = torch.randn(yr.shape, device=dev())
        (yr * G.double()).sum().backward()
        xd, wd = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
        bd = bb.clone().requires_grad_(True) if bias else None
        y = ops.conv_train(xd, wd, bd, None, stride, relu)
        (y * G).sum().backward()
        tag = ("conv", seed, k, (CI, CO, K, stride, relu, bias, B, H, W))
        errs = [_rel_scale(y, yr), _rel_scale(wd.grad, wr.grad), _rel_scale(xd.grad, xr.grad)]
        assert errs[0] <= 2e-5 and errs[1] <= 2e-5 and errs[2] <= 3e-5, (tag, errs)
        if bias:
            assert _rel_scale(bd.grad, br.grad) <= 2e-5, tag
        worst = max(worst, *errs)
        # ---- ConvTranspose2d(4, 2, 1)
        CI, CO = rng.choice(((65, 64), (33, 32), (64, 64), (9, 16), (32, 32)))
        B = rng.randint(1, 2)
        H, W = rng.randint(2, 24), rng.randint(2, 40)
        x = torch.rand(B, CI, H, W, device=dev()) * 2 - 0.5
        w = (torch.rand(CI, CO, 4, 4, device=dev()) - 0.5) * 0.2
        xr, wr = x.double().requires_grad_(True), w.double().requires_grad_(True)
        yr = F.conv_transpose2d(xr, wr, None, stride=2, padding=1)
        G = torch.randn(yr.shape, device=dev())
        (yr * G.double()).sum().backward()
        xd, wd = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
        y = ops.convT_train(xd, wd)
        (y * G).sum().backward()
        errs = [_rel_scale(y, yr), _rel_scale(wd.grad, wr.grad), _rel_scale(xd.grad, xr.grad)]
        assert errs[0] <= 1e-5 and errs[1] <= 2e-5 and errs[2] <= 2e-5, (("convT", seed, k, (CI, CO, B, H, W)), errs)
        worst = max(worst, *errs)
        # ---- BatchNorm2d, batch statistics (+ReLU), un-normalised inputs
        C = rng.choice((8, 32, 64))
        B = rng.randint(2, 4)
        H, W = rng.randint(3, 30), rng.randint(3, 50)
        relu = rng.random() < 0.7
        z = torch.randn(B, C, H, W, device=dev()) * 3 + 20
        bn = torch.nn.BatchNorm2d(C).to(dev()).train()
        with torch.no_grad():
            bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-1, 1)
        bn64 = torch.nn.BatchNorm2d(C).to(dev()).double().train()
        bn64.load_state_dict({k2: (v.double() if v.is_floating_point() else v) for k2, v in bn.state_dict().items()})
        zr = z.double().requires_grad_(True)
        yr = bn64(zr)
        if relu:
            yr = F.relu(yr)
        G = torch.randn(yr.shape, device=dev())
        (yr * G.double()).sum().backward()
        zd = z.clone().requires_grad_(True)
        y = ops.bn_train(zd, bn, None, relu)
        (y * G).sum().backward()
        errs = [_rel_scale(y, yr), _rel_scale(zd.grad, zr.grad), _rel_scale(bn.weight.grad, bn64.weight.grad),
                _rel_scale(bn.bias.grad, bn64.bias.grad), _rel_scale(bn.running_var, bn64.running_var)]
        assert max(errs) <= 2e-5, (("bn", seed, k, (B, C, H, W, relu)), errs)
        worst = max(worst, *errs)
    report("shape_sweep_training_kernels_seed%d" % seed, worst_rel=worst, shapes=18)


@pytest.mark.parametrize("seed", range(6))
def test_nconv2d_layer_random_shapes(seed):
    """torch.ops.nconv_b200.nconv2d forward + backward (the stream / tiled / pipelined / generic fp32 kernels behind one
    dispatcher, incl. the single-pass k3 and k5 backward) at drawn (Ci, Co, k, B, H, W): widths that are and are not
    multiples of 4 (vector and scalar paths), frames smaller than a tile, exact-zero confidences.  Checker: the layer's
    formula (step1.py:116-149) through torch autograd in fp64 on the same GPU.  y / cout max-rel 1e-4 at the 1e-3 floor,
    gradients 1e-4 of each tensor's scale (the bounds of test_gpu_parity.py)."""
    from nconv_b200 import ops  # noqa: F401  (registers torch.ops.nconv_b200.*)
    K = torch.ops.nconv_b200
    rng = random.Random(7000 + seed)
    torch.manual_seed(7000 + seed)
    worst_f = worst_g = 0.0
    for k_ in range(8):
        ci, co, k = rng.choice(((1, 2, 3), (2, 2, 3), (4, 4, 3), (16, 16, 3), (1, 2, 5), (2, 2, 5), (4, 4, 5), (16, 16, 5),
                                (1, 8, 5), (8, 8, 5), (16, 8, 3), (3, 5, 3), (8, 1, 1)))
        pad = rng.choice((k // 2, k // 2, 0)) if k > 1 else rng.choice((0, 2))
        B = rng.randint(1, 3)
        H, W = rng.randint(max(k, 3), 70), rng.choice((rng.randint(max(k, 3), 150), 4 * rng.randint(2, 40)))
        x = torch.rand(B, ci, H, W, device=dev()) * 10
        c = torch.rand(B, ci, H, W, device=dev()) * (torch.rand(B, ci, H, W, device=dev()) < 0.5)
        w = F.softplus(torch.rand(co, ci, k, k, device=dev()), beta=10)
        b = torch.rand(co, device=dev()) * 0.1
        xr, cr, wr, br = (t.double().requires_grad_(True) for t in (x, c, w, b))
        denom = F.conv2d(cr, wr, padding=pad)
        yr = F.conv2d(xr * cr, wr, padding=pad) / (denom + 1e-7) + br.view(1, -1, 1, 1)
        cor = denom / wr.sum((1, 2, 3)).view(1, -1, 1, 1)
        gy, gc = torch.randn(yr.shape, device=dev()), torch.randn(yr.shape, device=dev())
        ((yr * gy.double()).sum() + (cor * gc.double()).sum()).backward()
        xg, cg, wg, bg = (t.clone().requires_grad_(True) for t in (x, c, w, b))
        y, cout = K.nconv2d(xg, cg, wg, bg, pad, pad, 1e-7)
        ((y * gy).sum() + (cout * gc).sum()).backward()
        tag = (seed, k_, (ci, co, k, pad, B, H, W))
        ef = max(_rel_floor(y, yr.detach()), _rel_floor(cout, cor.detach()))
        assert ef <= 1e-4, ("fwd", tag, ef)
        eg = max(_rel_scale(xg.grad, xr.grad), _rel_scale(cg.grad, cr.grad), _rel_scale(wg.grad, wr.grad), _rel_scale(bg.grad, br.grad))
        assert eg <= 1e-4, ("bwd", tag, eg)
        worst_f, worst_g = max(worst_f, ef), max(worst_g, eg)
    report("shape_sweep_nconv2d_layer_seed%d" % seed, worst_fwd_rel_floor1e3=worst_f, worst_grad_rel=worst_g, shapes=8)
