"""GPU: the training-loop callers (masked multi-resolution MSE, fused AdamW) through the C-ABI against the
oracle restatement and the reference goldens (utils.py:63-71, :138-151, :53-55)."""
import numpy as np
import pytest
import torch

from conftest import T
from oracle import nconv_oracle as O
from nconv_b200.synthetic import loss_case

pytestmark = pytest.mark.gpu

LOSS_CASES = [(2, 11, (1, 2, 4, 8)), (1, 12, (1,)), (3, 13, (2, 4)), (2, 14, (1, 3, 5, 16))]   # = gen_golden.py
ADAMW_SHAPES = [(7,), (3, 5), (2, 3, 3, 3), (1,), (129,)]
LOSS_TOL = 2e-6       # scalar loss: double partial sums on the device, fp32 per-pixel residuals
GRAD_TOL = 1e-5       # gradients relative to max |ref|, against the reference's own fp32 run
GRAD_TOL64 = 1e-4     # against fp64: ATen (and the kernel) evaluate the source index (dst+0.5)*in/out-0.5 in fp32, so for
                      # a non-integer ratio the interpolation weight carries ~src*2^-24 (1e-5 at src=200) of rounding;
                      # 1e-4 is BASELINE.json north_star's bound


def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def U():
    from nconv_b200 import train_utils as U  # the drop-in mirror of the reference's utils.py (raises if the .so is missing)
    assert U.__file__.endswith("nconv_b200/train_utils.py")
    return U


def _sub(t):
    f = t.detach().cpu().reshape(-1)
    return f if f.numel() <= 20000 else f[::13]


@pytest.mark.parametrize("ci", range(len(LOSS_CASES)))
def test_multires_loss_vs_reference_golden(golden, U, ci):
    g = golden("loss_optim.npz")
    B, seed, scales = LOSS_CASES[ci]
    depths, gt = loss_case(B, seed, scales=scales)
    depths = [d.to(dev()).requires_grad_(True) for d in depths]
    loss = U.calculate_loss_multi_resolution(depths, gt.to(dev()), False)
    loss.backward()
    ref = float(g["mr%d_loss64" % ci])
    assert abs(float(loss.detach()) - ref) / abs(ref) < LOSS_TOL
    for k, d in enumerate(depths):
        r = T(g["mr%d_g64_%d" % (ci, k)])
        assert float((_sub(d.grad[0]).double() - r).abs().max()) / float(r.abs().max()) < GRAD_TOL64
        r = T(g["mr%d_g_%d" % (ci, k)])
        assert float((_sub(d.grad[0]) - r).abs().max()) / float(r.abs().max()) < GRAD_TOL
        assert float(d.grad[1:].abs().sum()) == 0.0          # utils.py:69: only sample 0 enters the loss


@pytest.mark.parametrize("ci", range(len(LOSS_CASES)))
def test_calculate_loss_vs_reference_golden(golden, U, ci):
    g = golden("loss_optim.npz")
    B, seed, _ = LOSS_CASES[ci]
    depths, gt = loss_case(B, seed, scales=(1,))
    d = depths[0].to(dev()).requires_grad_(True)
    loss = U.calculate_loss(d, gt.to(dev()), False)
    loss.backward()
    ref = float(g["cl%d_loss" % ci])
    assert abs(float(loss) - ref) / abs(ref) < 2e-6
    r = T(g["cl%d_g" % ci])
    assert float((_sub(d.grad) - r).abs().max()) / float(r.abs().max()) < GRAD_TOL


# odd sizes, non-integer and down-sampling ratios, all-masked and no-mask targets, scaled upstream gradient
@pytest.mark.parametrize("shape", [(2, 37, 53, (1, 2, 3)), (1, 8, 8, (1,)), (3, 45, 31, (7, 2)), (2, 64, 96, (1, 4, 8)),
                                   (1, 5, 300, (1, 2))])
@pytest.mark.parametrize("all_samples", [False, True])
@pytest.mark.parametrize("valid", [0.0, 0.6, 1.0])
def test_multires_loss_vs_oracle_fp64(U, shape, all_samples, valid):
    B, H, W, scales = shape
    depths, gt = loss_case(B, 500 + H, H=H, W=W, scales=scales, valid=valid)
    # one prediction larger than the target (down-sampling interpolation)
    g = torch.Generator().manual_seed(9)
    depths.append(torch.rand(B, 1, H + 3, W + 5, generator=g) * 10)
    d64 = [d.double().requires_grad_(True) for d in depths]
    ref = O.calculate_loss_multi_resolution(d64, gt.double(), all_samples=all_samples) * 3.0
    ref.backward()
    dc = [d.to(dev()).requires_grad_(True) for d in depths]
    loss = U.calculate_loss_multi_resolution(dc, gt.to(dev()), False, all_samples=all_samples) * 3.0
    loss.backward()
    assert abs(float(loss) - float(ref)) <= LOSS_TOL * abs(float(ref)) + 1e-12
    for a, b in zip(dc, d64):
        scale = float(b.grad.abs().max())
        if scale == 0.0:
            assert float(a.grad.abs().max()) == 0.0
        else:
            assert float((a.grad.double().cpu() - b.grad).abs().max()) / scale < GRAD_TOL64


def test_loss_is_deterministic_and_rejects_bad_input(U):
    depths, gt = loss_case(2, 3, scales=(1, 2, 4, 8))
    dc = [d.to(dev()).requires_grad_(True) for d in depths]
    vals, grads = [], []
    for _ in range(3):
        for d in dc:
            d.grad = None
        l = U.calculate_loss_multi_resolution(dc, gt.to(dev()), False, all_samples=True)
        l.backward()
        vals.append(float(l)); grads.append([d.grad.clone() for d in dc])
    assert vals[0] == vals[1] == vals[2]
    for k in range(len(dc)):
        assert torch.equal(grads[0][k], grads[1][k]) and torch.equal(grads[0][k], grads[2][k])
    with pytest.raises(ValueError):
        U.calculate_loss(dc[0].expand(-1, 2, -1, -1), gt.to(dev()).expand(-1, 2, -1, -1), True)   # Sobel is 1-channel
    with pytest.raises(RuntimeError):
        U.calculate_loss(depths[0], gt, False)               # CPU tensors: no CPU path
    with pytest.raises(ValueError):
        U.calculate_loss(dc[1], gt.to(dev()), False)         # shape mismatch


def test_full_size_loss_properties(U):
    """480x640, batch 8: loss(gt, gt) == 0 with zero gradients; scaling the residual by 2 scales the loss by 4."""
    depths, gt = loss_case(8, 21, scales=(1,))
    gtc = gt.to(dev())
    d = gtc.clone().requires_grad_(True)
    l = U.calculate_loss(d, gtc, False); l.backward()
    assert float(l) == 0.0 and float(d.grad.abs().max()) == 0.0
    m = (gtc != 0).float()
    r = torch.rand_like(gtc)
    l1 = float(U.calculate_loss(gtc + r * m, gtc, False))
    l2 = float(U.calculate_loss(gtc + 2 * r * m, gtc, False))
    assert abs(l2 / l1 - 4.0) < 1e-5
    expect = float(((r * m).double() ** 2).mean())
    assert abs(l1 - expect) / expect < 1e-5


def _frac_bad(a, ref, tol):
    ref = ref.double()
    return float(((a.double() - ref).abs() / float(ref.abs().max()) >= tol).float().mean())


@pytest.mark.parametrize("ci", range(len(LOSS_CASES)))
def test_gradient_loss_vs_reference_golden(golden, U, ci):
    """use_gradient_loss=True (train_step1.py:20): 0.8*RMSE + 0.2*Sobel-L1.  A Sobel response within rounding of 0
    can take the other sign in another summation order, which moves up to 9 gradient entries by O(1) of the
    Sobel term: isolated pixels (< 1e-4 of them) are allowed to differ, everything else is held to GRAD_TOL."""
    g = golden("loss_optim.npz")
    B, seed, scales = LOSS_CASES[ci]
    depths, gt = loss_case(B, seed, scales=(1,))
    d = depths[0].to(dev()).requires_grad_(True)
    loss = U.calculate_loss(d, gt.to(dev()), True)
    loss.backward()
    ref = float(g["sb%d_loss" % ci])
    assert abs(float(loss.detach()) - ref) / abs(ref) < 2e-6
    assert _frac_bad(_sub(d.grad), T(g["sb%d_g" % ci]), GRAD_TOL) < 1e-4
    depths, gt = loss_case(B, seed, scales=scales)
    depths = [x.to(dev()).requires_grad_(True) for x in depths]
    loss = U.calculate_loss_multi_resolution(depths, gt.to(dev()), True)
    loss.backward()
    ref = float(g["sbmr%d_loss" % ci])
    assert abs(float(loss.detach()) - ref) / abs(ref) < 2e-6
    for k, x in enumerate(depths):
        assert _frac_bad(_sub(x.grad[0]), T(g["sbmr%d_g_%d" % (ci, k)]), 2e-5) < 1e-4
        assert float(x.grad[1:].abs().sum()) == 0.0


@pytest.mark.parametrize("shape", [(2, 37, 53, (1, 2, 3)), (1, 8, 8, (1,)), (3, 45, 31, (7, 2)), (2, 64, 96, (1, 4, 8)),
                                   (1, 5, 300, (1, 2)), (1, 1, 1, (1,)), (2, 9, 33, (1,))])
@pytest.mark.parametrize("all_samples", [False, True])
@pytest.mark.parametrize("valid", [0.6, 1.0])
def test_gradient_loss_vs_oracle_fp64(U, shape, all_samples, valid):
    B, H, W, scales = shape
    depths, gt = loss_case(B, 700 + H, H=H, W=W, scales=[s for s in scales if H // s and W // s], valid=valid)
    if float(gt.abs().sum()) == 0.0:
        gt[..., 0, 0] = 3.0            # an all-masked target has rmse = 0: NaN gradient in the reference too
    d64 = [d.double().requires_grad_(True) for d in depths]
    ref = O.calculate_loss_multi_resolution(d64, gt.double(), all_samples=all_samples, use_gradient_loss=True) * 1.5
    ref.backward()
    dc = [d.to(dev()).requires_grad_(True) for d in depths]
    loss = U.calculate_loss_multi_resolution(dc, gt.to(dev()), True, all_samples=all_samples) * 1.5
    loss.backward()
    assert abs(float(loss.detach()) - float(ref.detach())) <= 3e-6 * abs(float(ref.detach()))
    for a, b in zip(dc, d64):
        assert _frac_bad(a.grad.cpu(), b.grad, GRAD_TOL64) < 2e-3


def test_gradient_loss_all_masked_is_nan_like_reference(U):
    """rmse == 0 -> d sqrt = inf * 0 = NaN in torch's autograd; the kernel follows the same arithmetic where the
    mask lets a gradient through, and an all-zero target masks everything (gradient exactly 0)."""
    gt = torch.zeros(1, 1, 16, 16, device=dev())
    d = torch.rand(1, 1, 16, 16, device=dev()).requires_grad_(True)
    l = U.calculate_loss(d, gt, True); l.backward()
    assert float(l.detach()) == 0.0 and float(d.grad.abs().max()) == 0.0
    gt2 = torch.rand(1, 1, 16, 16, device=dev()) + 1
    d2 = gt2.clone().requires_grad_(True)
    l2 = U.calculate_loss(d2, gt2, True); l2.backward()
    assert float(l2.detach()) == 0.0 and bool(torch.isnan(d2.grad).all())
    r64 = gt2.double().cpu().clone().requires_grad_(True)
    O.calculate_loss(r64, gt2.double().cpu(), True).backward()
    assert bool(torch.isnan(r64.grad).all())


def _adamw_net():
    gen = torch.Generator().manual_seed(77)
    return [torch.nn.Parameter(torch.randn(*s, generator=gen).to(dev())) for s in ADAMW_SHAPES]


def test_fused_adamw_vs_reference_golden(golden, U):
    g = golden("loss_optim.npz")

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.ps = torch.nn.ParameterList(_adamw_net())
    net = Net()
    opt = U.get_optimizer(net, "adam", 1e-2, 1e-2)
    gen = torch.Generator().manual_seed(78)
    for step in range(4):
        opt.zero_grad()
        for p in net.ps:
            gr = (torch.randn(p.shape, generator=gen) * (0.1 + step)).to(dev())
            if p.grad is None:
                p.grad = gr
            else:
                p.grad.add_(gr)                              # accumulates into the flat gradient buffer view
        opt.step()
        for k, p in enumerate(net.ps):
            ref = T(g["adamw_s%d_p%d" % (step, k)])
            assert float((p.detach().cpu() - ref).abs().max()) < 2e-6 * (1 + float(ref.abs().max())), (step, k)


def test_fused_adamw_vs_torch_with_unused_parameter_and_autograd():
    from nconv_b200 import ops
    torch.manual_seed(0)
    a = torch.nn.Sequential(torch.nn.Linear(33, 17), torch.nn.Tanh(), torch.nn.Linear(17, 5)).to(dev())
    unused_a = torch.nn.Parameter(torch.randn(9, device=dev()))
    import copy
    b = copy.deepcopy(a)
    unused_b = torch.nn.Parameter(unused_a.detach().clone())
    oa = torch.optim.AdamW(list(a.parameters()) + [unused_a], lr=3e-3, weight_decay=0.05)
    ob = ops.FusedAdamW(list(b.parameters()) + [unused_b], lr=3e-3, weight_decay=0.05)
    for step in range(6):
        x = torch.randn(64, 33, device=dev())
        for net, o in ((a, oa), (b, ob)):
            o.zero_grad()
            net(x).square().mean().backward()
            o.step()
        for pa, pb in zip(a.parameters(), b.parameters()):
            assert float((pa - pb).abs().max()) < 1e-6, step
    assert torch.equal(unused_a, unused_b)                   # torch.optim skips grad-less parameters: so do we
    assert ob.flat_param.numel() % 4 == 0 and all(p.data_ptr() % 16 == 0 for p in b.parameters())
    unused_b.grad = torch.ones_like(unused_b)
    with pytest.raises(RuntimeError):
        ob.step()
