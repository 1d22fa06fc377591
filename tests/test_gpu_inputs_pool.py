"""SURVEY 8f ranks 3 and 4 on the GPU: the input-pipeline kernels (bit-exact against the torch CPU ops the reference's
loaders run) and the opt-in confidence-argmax pooling (bit-exact values and argmax routing against
F.max_pool2d(return_indices=True), ties and NaN included)."""
import pytest
import torch
import torch.nn.functional as F

from oracle import nconv_oracle as O

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


@pytest.mark.parametrize("shape", [(2, 480, 640), (1, 37, 53), (3, 352, 1216)])
def test_decode_rgb_bit_exact(shape):
    """dataset/nyuloader.py:69: torch.FloatTensor(cv2.imread(path)).permute(2, 0, 1)."""
    from nconv_b200.ops import decode_rgb
    B, H, W = shape
    g = torch.Generator().manual_seed(H)
    img = torch.randint(0, 256, (B, H, W, 3), generator=g, dtype=torch.uint8)
    ref = img.float().permute(0, 3, 1, 2).contiguous()
    assert torch.equal(decode_rgb(img.to(dev())).cpu(), ref)


@pytest.mark.parametrize("dtype,scale", [(torch.uint8, 1 / 256.0), (torch.uint16, 1 / 256.0), (torch.float32, 1.0)])
def test_decode_depth_bit_exact(dtype, scale):
    """dataset/kittiloader.py:66-76: cv2.imread(...).astype(np.float32) / 256.0; nyuloader.py:71-75: np.load -> FloatTensor."""
    from nconv_b200.ops import decode_depth
    g = torch.Generator().manual_seed(3)
    if dtype == torch.float32:
        raw = torch.rand(2, 40, 50, generator=g) * 10
    else:
        hi = 256 if dtype == torch.uint8 else 65536
        raw = torch.randint(0, hi, (2, 40, 50), generator=g, dtype=torch.int32).to(dtype)
    ref = (raw.float() if dtype != torch.uint16 else raw.to(torch.int32).float()) * torch.tensor(scale, dtype=torch.float32)
    out = decode_depth(raw.to(dev()), scale)
    assert out.shape == (2, 1, 40, 50) and torch.equal(out.cpu()[:, 0], ref)


@pytest.mark.parametrize("use_mask,use_noise", [(True, True), (True, False), (False, True)])
def test_sparsify_bit_exact(use_mask, use_noise):
    """dataset/nyuloader.py:83-124 given the same random draws."""
    from nconv_b200.ops import sparsify
    g = torch.Generator().manual_seed(5)
    raw = torch.rand(2, 1, 48, 64, generator=g) * 9 + 0.5
    mask = (torch.rand(2, 1, 48, 64, generator=g) < 0.3).to(torch.uint8) if use_mask else None
    noise = None
    if use_noise:
        noise = torch.zeros(2 * 48 * 64)
        idx = torch.randperm(noise.numel(), generator=g)[: noise.numel() // 10]
        noise[idx] = torch.rand(idx.numel(), generator=g) * 0.2 - 0.1
        noise = noise.view_as(raw)
    ref = O.sparsify(raw, mask, noise)
    out = sparsify(raw.to(dev()), None if mask is None else mask.to(dev()), None if noise is None else noise.to(dev()))
    # the kernel contracts d + d*noise into one fma (single rounding); the reference's two-step result is within 1 ulp
    assert torch.allclose(out.cpu(), ref, rtol=2e-7, atol=0) and torch.equal(out.cpu() == 0, ref == 0)


@pytest.mark.parametrize("shape", [(2, 3, 8, 12), (1, 8, 9, 13), (2, 8, 60, 80)])
def test_pool_conf_argmax_bit_exact_with_ties_and_nan(shape):
    K = torch.ops.nconv_b200
    import nconv_b200.ops  # noqa: F401
    g = torch.Generator().manual_seed(sum(shape))
    c = torch.randint(0, 3, shape, generator=g).float()            # heavy ties: first maximum wins
    c[0, 0, 0, 1] = float("nan")
    x = torch.randn(shape, generator=g)
    xr, cr = x.clone().requires_grad_(True), c.clone().requires_grad_(True)
    xo_ref, co_ref = O.pool_conf_argmax(xr, cr)
    g1, g2 = torch.randn(xo_ref.shape, generator=g), torch.randn(xo_ref.shape, generator=g)
    (xo_ref * g1).sum().backward(retain_graph=True)
    gx_ref = xr.grad.clone()
    xd, cd = x.to(dev()).requires_grad_(True), c.to(dev()).requires_grad_(True)
    xo, co = K.pool_conf_argmax(xd, cd)
    nn = lambda t: torch.nan_to_num(t.detach().cpu(), nan=-7.0)        # noqa: E731
    assert torch.equal(nn(xo), nn(xo_ref)) and torch.equal(nn(co), nn(co_ref))
    ((xo * g1.to(dev())).sum() + (co * g2.to(dev())).sum()).backward()
    assert torch.equal(xd.grad.cpu(), gx_ref)                           # routing = the confidence's argmax, bit-exact
    # confidence gradient: g2 / 4 at the argmax
    _, idx = F.max_pool2d(c, 2, 2, return_indices=True)
    gc_ref = torch.zeros_like(c).flatten(2).scatter_(2, idx.flatten(2), (g2 * 0.25).flatten(2)).view_as(c)
    assert torch.equal(cd.grad.cpu(), gc_ref)


def test_dnet_conf_argmax_option_vs_oracle():
    """DNET(pool='conf_argmax') against the oracle graph with O.pool_conf_argmax in place of the two max-pools; the
    default stays the reference's independent pooling."""
    from models.step1 import DNET
    torch.manual_seed(4)
    m = DNET(32, pool="conf_argmax").to(dev()).eval()
    assert DNET(32).pool == "independent"
    S = (torch.rand(2, 1, 48, 64) * 9 + 1) * (torch.rand(2, 1, 48, 64) < 0.1)
    with torch.no_grad():
        out = m(S.to(dev()))
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}

    def layer(name, x, c, pad):
        return O.nconv2d(x, c, sd[name + ".weight"], sd[name + ".bias"], (pad, pad))
    c0 = (S > 0.01).float()
    x1, c1 = layer("nconv1", S, c0, 2)
    x1, c1 = layer("nconv2", x1, c1, 2)
    x2, c2 = layer("nconv_down1", *O.pool_conf_argmax(x1, c1), 2)
    x3, c3 = layer("nconv_down2", *O.pool_conf_argmax(x2, c2), 2)
    x4, c4 = layer("nconv_down3", *O.pool_conf_argmax(x3, c3), 2)
    up = lambda t, r: F.interpolate(t, r.shape[2:], mode="nearest")      # noqa: E731
    x34, c34 = layer("nconv4", torch.cat((x3, up(x4, x3)), 1), torch.cat((c3, up(c4, c3)), 1), 1)
    x23, c23 = layer("nconv5", torch.cat((x2, up(x34, x2)), 1), torch.cat((c2, up(c34, c2)), 1), 1)
    x6, c6 = layer("nconv6", torch.cat((up(x23, x1), x1), 1), torch.cat((up(c23, c1), c1), 1), 0)
    x7, _ = layer("nconv7", x6, c6, 2)
    assert O.max_rel_err(out, x7[:, :, 1:49, 1:65]) <= 1e-4
