"""CPU: the oracle restatement (oracle/nconv_oracle.py) against the golden vectors produced by
the unmodified reference (oracle/gen_golden.py).  This is what pins the oracle."""
import numpy as np
import pytest
import torch

from conftest import REL_TOL, T
from oracle import nconv_oracle as O
from nconv_b200.synthetic import (NCONV_LAYER_CASES, nconv_layer_case, sparse_depth_kitti, sparse_depth_nyu,
                                  setp2_inputs, randomise_bn_stats)

# The oracle issues the same ATen calls as the reference, so on the generating machine it is
# bit-identical; oneDNN may pick another ISA path on another CPU, hence a (tight) tolerance.
ORACLE_TOL = 2e-5


@pytest.mark.parametrize("idx", range(len(NCONV_LAYER_CASES)))
def test_nconv_layer_forward_backward(golden, idx):
    g = golden("nconv2d_layers.npz")
    ci, co, k, pad, B, H, W = NCONV_LAYER_CASES[idx]
    x, c, gy, gc = nconv_layer_case(idx)
    p = "case%d_" % idx
    w = T(g[p + "w"]).requires_grad_(True)
    b = T(g[p + "b"]).requires_grad_(True)
    x.requires_grad_(True); c.requires_grad_(True)
    y, cout = O.nconv2d(x, c, w, b, (pad, pad))
    assert O.max_rel_err(y, T(g[p + "y"])) < ORACLE_TOL
    assert O.max_rel_err(cout, T(g[p + "cout"])) < ORACLE_TOL
    ((y * gy).sum() + (cout * gc).sum()).backward()
    for name, t in (("gx", x.grad), ("gc", c.grad), ("gw", w.grad), ("gb", b.grad)):
        ref = T(g[p + name])
        scale = float(ref.abs().max()) + 1e-30
        assert float((t - ref).abs().max()) / scale < ORACLE_TOL, name


def _sd(g):
    return {k[3:]: T(g[k]) for k in g.files if k.startswith("sd_")}


def test_setp1_small_and_intermediates(golden):
    g = golden("setp1.npz")
    sd = _sd(g)
    S = T(g["small_S"])
    out, acts = O.dnet_forward(S, sd, return_all=True)
    ref = T(g["small_out"])                       # reference crop [1:481,1:641] of 66x98 -> 65x97
    assert ref.shape[-2:] == (65, 97)
    assert O.max_rel_err(out, ref[:, :, :64, :96]) < ORACLE_TOL
    assert O.max_rel_err(O.dnet_forward(S, sd, crop="reference"), ref) < ORACLE_TOL
    pairs = {"nconv2": ("x1", "c1"), "nconv_down1": ("x2_ds", "c2_ds"), "nconv_down2": ("x3_ds", "c3_ds"),
             "nconv_down3": ("x4_ds", "c4_ds"), "nconv4": ("x34_ds", "c34_ds"), "nconv5": ("x23_ds", "c23_ds"),
             "nconv6": ("x6", "c6"), "nconv7": ("x7", "c7")}
    for layer, (xn, cn) in pairs.items():
        assert O.max_rel_err(acts[xn], T(g["small_%s_x" % layer])) < ORACLE_TOL, layer
        assert O.max_rel_err(acts[cn], T(g["small_%s_c" % layer])) < ORACLE_TOL, layer


def test_setp1_odd_size(golden):
    g = golden("setp1.npz")
    out = O.dnet_forward(T(g["odd_S"]), _sd(g))
    assert O.max_rel_err(out, T(g["odd_out"])[:, :, :45, :70]) < ORACLE_TOL


def test_setp1_nyu_config1(golden):
    g = golden("setp1.npz")
    out = O.dnet_forward(sparse_depth_nyu(1, seed=0), _sd(g))
    ref = T(g["nyu_out"])
    assert out.shape == ref.shape == (1, 1, 480, 640)
    assert O.max_rel_err(out, ref) < ORACLE_TOL
    # KA3: the one-pixel border ring equals nconv7.bias exactly
    b7 = float(_sd(g)["d_net.nconv7.bias"][0])
    for ring in (out[0, 0, 0], out[0, 0, -1], out[0, 0, :, 0], out[0, 0, :, -1]):
        assert torch.all(ring == b7)


def test_setp1_kitti_overlap(golden):
    """KA2: the reference returns 353x640 for 352x1216; the generalised crop must agree on the
    overlap (rows 0..351, cols 0..639)."""
    g = golden("setp1.npz")
    out = O.dnet_forward(sparse_depth_kitti(1, seed=2000), _sd(g))
    ref = T(g["kitti_out"])
    assert ref.shape == (1, 1, 353, 640) and out.shape == (1, 1, 352, 1216)
    assert O.max_rel_err(out[..., :640], ref[:, :, :352]) < ORACLE_TOL


def test_setp2_export(golden):
    from models.step2 import SETP2_BP_EXPORT     # parameter container only: no CUDA needed to build
    g = golden("setp2_export.npz")
    torch.manual_seed(5)
    m = SETP2_BP_EXPORT()
    from test_boundary_cpu import state_fingerprint
    assert state_fingerprint(m.state_dict()) == str(g["init_fingerprint_seed5"])
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    sd = O.apply_enforce_pos(sd, prefix="step1.d_net.")
    m.load_state_dict(sd)
    m.eval()
    randomise_bn_stats(m, seed=55)
    assert state_fingerprint(m.state_dict()) == str(g["eval_fingerprint"])
    sd = {k: v.detach() for k, v in m.state_dict().items()}
    rgb0, d0, rgb1, d1 = setp2_inputs(1, seed=7)
    torch.set_num_threads(max(1, torch.get_num_threads()))
    with torch.no_grad():
        depths = O.setp2_forward(sd, rgb0, d0, rgb1, d1, export=False)
        out = O.setp2_forward(sd, rgb0, d0, rgb1, d1, export=True)
    assert O.max_rel_err(out[:1], T(g["out0"])) < ORACLE_TOL
    assert O.max_rel_err(out[1:], T(g["out1"])) < ORACLE_TOL
    for i, n in enumerate(("fuse0", "fuse1", "fuse2")):
        assert O.max_rel_err(depths[i], T(g[n + "_depth"])) < ORACLE_TOL
    assert torch.all(out[:, :, :45] == 0) and torch.all(out[:, :, -45:] == 0) and torch.all(out[..., :20] == 0)


# ---------------------------------------------------------------------------------------------
# training-loop callers: utils.py:63-71 / :138-151 (loss) and :53-55 (AdamW)
# ---------------------------------------------------------------------------------------------
import torch.nn.functional as F  # noqa: E402
from nconv_b200.synthetic import loss_case  # noqa: E402

LOSS_CASES = [(2, 11, (1, 2, 4, 8)), (1, 12, (1,)), (3, 13, (2, 4)), (2, 14, (1, 3, 5, 16))]   # = gen_golden.py
ADAMW_SHAPES = [(7,), (3, 5), (2, 3, 3, 3), (1,), (129,)]


def _sub_np(t):
    f = t.detach().cpu().reshape(-1)
    return f if f.numel() <= 20000 else f[::13]


@pytest.mark.parametrize("ci", range(len(LOSS_CASES)))
@pytest.mark.parametrize("dt", [torch.float32, torch.float64])
def test_multires_loss_oracle_vs_reference_golden(golden, ci, dt):
    g = golden("loss_optim.npz")
    B, seed, scales = LOSS_CASES[ci]
    tag = "" if dt == torch.float32 else "64"
    depths, gt = loss_case(B, seed, scales=scales)
    depths = [d.to(dt).requires_grad_(True) for d in depths]
    loss = O.calculate_loss_multi_resolution(depths, gt.to(dt))
    loss.backward()
    tol = ORACLE_TOL if dt == torch.float32 else 1e-12
    ref = float(g["mr%d_loss%s" % (ci, tag)])
    assert abs(float(loss) - ref) / abs(ref) < tol
    for k, d in enumerate(depths):
        r = T(g["mr%d_g%s_%d" % (ci, tag, k)])
        assert float((_sub_np(d.grad[0]) - r).abs().max()) / float(r.abs().max()) < tol
        assert float(d.grad[1:].abs().sum()) == 0.0


@pytest.mark.parametrize("ci", range(len(LOSS_CASES)))
def test_calculate_loss_oracle_vs_reference_golden(golden, ci):
    g = golden("loss_optim.npz")
    B, seed, _ = LOSS_CASES[ci]
    depths, gt = loss_case(B, seed, scales=(1,))
    d = depths[0].requires_grad_(True)
    loss = O.calculate_loss(d, gt)
    loss.backward()
    ref = float(g["cl%d_loss" % ci])
    assert abs(float(loss) - ref) / abs(ref) < ORACLE_TOL
    r = T(g["cl%d_g" % ci])
    assert float((_sub_np(d.grad) - r).abs().max()) / float(r.abs().max()) < ORACLE_TOL


def test_upsample_restatement_vs_aten():
    torch.manual_seed(4)
    for (h, w, H, W) in [(60, 80, 480, 640), (7, 5, 20, 33), (48, 64, 48, 64), (33, 20, 7, 5), (1, 1, 4, 4)]:
        x = torch.randn(2, 1, h, w, dtype=torch.float64)
        ref = F.interpolate(x, size=(H, W), mode="bilinear", align_corners=False)
        assert float((O.upsample_bilinear_af(x, H, W) - ref).abs().max()) < 1e-12


def test_adamw_oracle_vs_reference_golden(golden):
    g = golden("loss_optim.npz")
    gen = torch.Generator().manual_seed(77)
    ps = [torch.randn(*s, generator=gen) for s in ADAMW_SHAPES]
    ms = [torch.zeros_like(p) for p in ps]
    vs = [torch.zeros_like(p) for p in ps]
    gen = torch.Generator().manual_seed(78)
    for step in range(4):
        for k in range(len(ps)):
            grad = torch.randn(ps[k].shape, generator=gen) * (0.1 + step)
            ps[k], ms[k], vs[k] = O.adamw_step(ps[k], grad, ms[k], vs[k], step + 1, lr=1e-2, weight_decay=1e-2)
            ref = T(g["adamw_s%d_p%d" % (step, k)])
            assert float((ps[k] - ref).abs().max()) < 1e-6 * (1 + float(ref.abs().max()))


@pytest.mark.parametrize("ci", range(len(LOSS_CASES)))
def test_gradient_loss_oracle_vs_reference_golden(golden, ci):
    """use_gradient_loss=True: 0.8*RMSE + 0.2*Sobel-L1 (utils.py:95-147)."""
    g = golden("loss_optim.npz")
    B, seed, scales = LOSS_CASES[ci]
    depths, gt = loss_case(B, seed, scales=(1,))
    d = depths[0].requires_grad_(True)
    loss = O.calculate_loss(d, gt, True)
    loss.backward()
    ref = float(g["sb%d_loss" % ci])
    assert abs(float(loss.detach()) - ref) / abs(ref) < ORACLE_TOL
    r = T(g["sb%d_g" % ci])
    assert float((_sub_np(d.grad) - r).abs().max()) / float(r.abs().max()) < ORACLE_TOL
    depths, gt = loss_case(B, seed, scales=scales)
    depths = [x.requires_grad_(True) for x in depths]
    loss = O.calculate_loss_multi_resolution(depths, gt, use_gradient_loss=True)
    loss.backward()
    ref = float(g["sbmr%d_loss" % ci])
    assert abs(float(loss.detach()) - ref) / abs(ref) < ORACLE_TOL
    for k, x in enumerate(depths):
        r = T(g["sbmr%d_g_%d" % (ci, k)])
        # a Sobel response within rounding of 0 may take the other sign: allow isolated pixels
        bad = ((_sub_np(x.grad[0]) - r).abs() / float(r.abs().max()) >= ORACLE_TOL).float().mean()
        assert float(bad) < 1e-4
