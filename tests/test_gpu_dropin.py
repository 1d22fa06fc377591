"""Drop-in behaviour under the reference's own call sequences: nn.DataParallel wrap, stock torch
loss / optimizer / scheduler objects, copy.deepcopy, `module.`-prefixed checkpoints, eval-mode BatchNorm with
autograd, per-module caches, the fused optimizer as a torch.optim.Optimizer, streamed inference.

Restated from /root/reference/train_step1.py:24-125,152-153 and train_step2.py:24-107,134-135 (the scripts
themselves need datasets and matplotlib, neither of which exists on the box)."""
import copy
import os

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from oracle import nconv_oracle as O
from nconv_b200.synthetic import setp2_inputs, sparse_depth_nyu, randomise_bn_stats

pytestmark = pytest.mark.gpu
H, W = 64, 96


def dev():
    return torch.device("cuda:0")


def _masked_mse(est, gt):
    """utils.py:138-151 with stock torch ops — what the unchanged scripts run."""
    return F.mse_loss(est.masked_fill(gt == 0, 0), gt)


@pytest.mark.parametrize("ngpu", [1, 2])
def test_train_step1_script_sequence_under_dataparallel(ngpu, tmp_path, monkeypatch):
    if torch.cuda.device_count() < ngpu:
        pytest.skip("needs %d GPUs" % ngpu)
    from models.step1 import SETP1_NCONV
    from models.step2 import SETP2_BP_TRAIN
    torch.manual_seed(11)
    model = SETP1_NCONV()                                                  # train_step1.py:152
    model = nn.DataParallel(model, device_ids=list(range(ngpu)))           # :153
    model.to(dev())                                                        # :27
    optim = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=1e-7)     # reference utils.py:55
    sched = torch.optim.lr_scheduler.LinearLR(optim, start_factor=1.0, end_factor=0, total_iters=4)   # :38
    depth = sparse_depth_nyu(2 * ngpu, seed=3, H=H, W=W, points=400)
    gt = depth * 1.05
    losses = []
    for _ in range(3):
        model.train()                                                      # :50
        optim.zero_grad()
        est = model(depth.to(dev()))                                       # :61
        loss = _masked_mse(est, gt.to(dev()))                              # :63
        loss.backward()
        optim.step()
        sched.step()
        losses.append(float(loss))
    assert all(torch.isfinite(torch.tensor(losses)))
    w = model.module.d_net.nconv2.weight
    assert w.grad is not None and float(w.grad.abs().sum()) > 0
    model.eval()                                                           # utils.py:22-36 validation
    with torch.no_grad():
        val = model(depth.to(dev()))
    sd = {k: v.detach().cpu() for k, v in model.module.state_dict().items()}
    ref = O.dnet_forward(depth, sd)
    assert O.max_rel_err(val, ref) <= 1e-4
    best = copy.deepcopy(model)                                            # :98
    with torch.no_grad():
        assert torch.equal(best(depth.to(dev())), val)
    # save_checkpoint (utils.py:42-51) -> SETP2_BP_TRAIN's loader strips the `module.` prefix (step2.py:29-40)
    os.makedirs(tmp_path / "checkpoints")
    torch.save({"epoch": 1, "state_dict": model.state_dict(), "stats": []}, tmp_path / "checkpoints" / "s1.pth.tar")
    monkeypatch.chdir(tmp_path)
    m2 = SETP2_BP_TRAIN("s1")
    for k, v in model.module.state_dict().items():
        assert torch.equal(m2.step1.state_dict()[k].cpu(), v.cpu()), k
    assert not any(p.requires_grad for p in m2.step1.parameters())


@pytest.mark.parametrize("ngpu", [1, 2])
def test_train_step2_script_sequence_under_dataparallel(ngpu, tmp_path, monkeypatch):
    if torch.cuda.device_count() < ngpu:
        pytest.skip("needs %d GPUs" % ngpu)
    from models.step1 import SETP1_NCONV
    from models.step2 import SETP2_BP_TRAIN
    torch.manual_seed(3)
    os.makedirs(tmp_path / "checkpoints")
    torch.save({"state_dict": {"module." + k: v for k, v in SETP1_NCONV().state_dict().items()}},
               tmp_path / "checkpoints" / "s1.pth.tar")
    monkeypatch.chdir(tmp_path)
    torch.manual_seed(5)
    model = nn.DataParallel(SETP2_BP_TRAIN("s1"), device_ids=list(range(ngpu)))    # train_step2.py:134-135
    model.to(dev())
    optim = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-7)
    sched = torch.optim.lr_scheduler.LinearLR(optim, start_factor=1.0, end_factor=0, total_iters=4)
    rgb0, d0, _, _ = setp2_inputs(ngpu, seed=9, H=H, W=W)
    gt = (d0 * 1.1 + 0.5)
    for _ in range(2):
        model.train()                                                      # :60
        optim.zero_grad()
        depths, _ = model(rgb0.to(dev()), d0.to(dev()), rgb0.to(dev()), d0.to(dev()))     # :62
        loss = sum(_masked_mse(F.interpolate(d, size=(H, W), mode="bilinear", align_corners=False)[0], gt.to(dev())[0])
                   for d in depths) / len(depths)                           # utils.py:63-71
        loss.backward()
        optim.step()
        sched.step()
        assert torch.isfinite(loss)
    assert model.module.fuse3.fuse.fuse_conv3.conv.weight.grad is not None
    assert model.module.step1.d_net.nconv1.weight.grad is None             # frozen (step2.py:38-40)
    # validation pattern of utils.py:74-93 keeps train(); an eval()/no_grad pass takes the tensor-core trunk
    model.eval()
    with torch.no_grad():
        a, b = model(rgb0.to(dev()), d0.to(dev()), rgb0.to(dev()), d0.to(dev()))
    sd = {k: v.detach().cpu() for k, v in model.module.state_dict().items()}
    ref = O.setp2_forward(sd, rgb0, d0, rgb0, d0, export=False)
    B = rgb0.shape[0]
    for i in range(4):
        assert O.max_rel_err(torch.cat((a[i], b[i]), 0), ref[i], floor=0.1) <= 1e-4, i
        assert torch.equal(a[i][:B], b[i][:B])
    # per-module caches: not in the state_dict, not copied by deepcopy, rebuilt lazily
    assert not any("_nconv_rt" in k for k in model.state_dict())
    clone = copy.deepcopy(model)
    rt = clone.module.__dict__.get("_nconv_rt")
    assert rt is None or (not rt.ws and not rt.pack)
    with torch.no_grad():
        a2, _ = clone(rgb0.to(dev()), d0.to(dev()), rgb0.to(dev()), d0.to(dev()))
    assert torch.equal(a2[3], a[3])


@pytest.mark.parametrize("ngpu", [1, 2])
def test_tensor_core_dnet_under_dataparallel(ngpu):
    """The tcgen05 DNET pipeline (eligible from 32x224 up) through nn.DataParallel's thread-per-device replicas:
    per-(device, stream) workspaces, per-device launch attributes, tensor maps built per call."""
    if torch.cuda.device_count() < ngpu:
        pytest.skip("needs %d GPUs" % ngpu)
    from models.step1 import SETP1_NCONV
    torch.manual_seed(2)
    model = nn.DataParallel(SETP1_NCONV(), device_ids=list(range(ngpu))).to(dev())
    model.train()
    with torch.no_grad():
        model(torch.zeros(ngpu, 1, 16, 16, device=dev()))        # one training-mode forward: the hook applies softplus
    model.eval()
    S = sparse_depth_nyu(2 * ngpu, seed=8, H=64, W=256, points=600)
    with torch.no_grad():
        out = model(S.to(dev()))
        again = model(S.to(dev()))
    sd = {k: v.detach().cpu() for k, v in model.module.state_dict().items()}
    if ngpu == 1:                                               # with replicas the hook mutates the replicas, not the master
        assert O.max_rel_err(out, O.dnet_forward(S, sd)) <= 1e-4
    assert torch.equal(out, again)
    with torch.no_grad():
        halves = torch.cat([model.module(S[i:i + 2].to(dev())) for i in range(0, 2 * ngpu, 2)], 0)
    if ngpu == 1:
        assert torch.equal(out, halves)                        # batch sharding is bit-identical on the tensor-core path too


def test_two_models_never_share_packed_weights():
    """the pack cache used to be keyed by id(module); a second model allocated at the same addresses
    could hit the first one's packed weights.  The cache now lives on the module."""
    from models.step2 import SETP2_BP_EXPORT
    from nconv_b200.synthetic import synthetic_state_dict
    rgb0, d0, rgb1, d1 = (t.to(dev()) for t in setp2_inputs(1, seed=4, H=128, W=W))     # rows 45..82 survive step2.py:122-123
    outs = []
    for seed in (5, 6):
        m = SETP2_BP_EXPORT()
        m.load_state_dict(synthetic_state_dict(seed), strict=False)
        m = m.eval().to(dev())
        with torch.no_grad():
            outs.append(m(rgb0, d0, rgb1, d1)[0].clone())
        sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
        ref = O.setp2_forward(sd, rgb0.cpu(), d0.cpu(), rgb1.cpu(), d1.cpu(), export=True)
        assert O.max_rel_err(outs[-1], ref[:1], floor=0.1) <= 1e-4
        del m
    assert not torch.equal(outs[0], outs[1])


def test_pack_cache_follows_raw_pointer_updates():
    """Kernels that write parameters through raw pointers (fused AdamW) must invalidate the packed weights."""
    from models.step2 import SETP2_BP_EXPORT
    from nconv_b200 import ops
    from nconv_b200.synthetic import synthetic_state_dict
    m = SETP2_BP_EXPORT()
    m.load_state_dict(synthetic_state_dict(5), strict=False)
    m = m.eval().to(dev())
    inp = [t.to(dev()) for t in setp2_inputs(1, seed=4, H=128, W=W)]
    with torch.no_grad():
        o0 = m(*inp)[0].clone()
    p = m.fuse3.fuse.fuse_conv3.conv.weight
    opt = ops.FusedAdamW([p], lr=0.05, weight_decay=0.0)
    p.grad = torch.ones_like(p)
    opt.step()
    with torch.no_grad():
        o1 = m(*inp)[0].clone()
    assert not torch.equal(o0, o1)
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    ref = O.setp2_forward(sd, *[t.cpu() for t in inp], export=True)
    assert O.max_rel_err(o1, ref[:1], floor=0.1) <= 1e-4


def test_eval_mode_batchnorm_passes_gradients():
    """model.eval() without no_grad (frozen-statistics BatchNorm) used to raise; the eval-mode BatchNorm is
    a per-channel affine folded into the conv weights.  The four multi-scale depths of step2.py:60-75 and the
    gradients of every parameter against autograd through the CPU oracle in fp64."""
    from models.step2 import SETP2_BP_EXPORT, _setp2_trunk
    from nconv_b200.synthetic import synthetic_state_dict
    m = SETP2_BP_EXPORT()
    m.load_state_dict(synthetic_state_dict(5), strict=False)
    m = m.eval().to(dev())
    for p in m.step1.parameters():
        p.requires_grad = False                            # as SETP2_BP_TRAIN does (step2.py:38-40)
    hh, ww = 32, 48
    inp = setp2_inputs(1, seed=4, H=hh, W=ww)
    depths, _ = _setp2_trunk(m, *[t.to(dev()) for t in inp])
    gen = torch.Generator().manual_seed(1)
    Gs = [torch.randn(2, 1, hh // f, ww // f, generator=gen) for f in (8, 4, 2, 1)]
    sum((d * G.to(dev())).sum() for d, G in zip(depths, Gs)).backward()
    trainable = {k for k, p in m.named_parameters() if p.requires_grad}
    sd64 = {k: v.detach().cpu().double().requires_grad_(k in trainable) for k, v in m.state_dict().items()}
    ref = O.setp2_forward(sd64, *[t.double() for t in inp], export=False)
    sum((d * G.double()).sum() for d, G in zip(ref, Gs)).backward()
    for d, r in zip(depths, ref):
        # fp32 kernels against an fp64 run; at 32x48 the refined depths are ~0.1 m, so the bound is absolute here
        assert O.max_rel_err(d, r.detach().float(), floor=1.0) <= 1e-4
    checked = 0
    for name, p in m.named_parameters():
        if name.startswith("step1.") or p.grad is None:
            continue
        g64 = sd64[name].grad
        scale = float(g64.abs().max())
        if scale < 1e-9:
            continue
        err = float((p.grad.double().cpu() - g64).abs().max()) / scale
        assert err <= 2e-3, (name, err)                    # fp32 ReLU-chain noise of a whole-net gradient
        checked += 1
    assert checked > 30
    assert m.rgb_encoder1.encoder[1].weight.grad is not None          # BatchNorm gamma receives its gradient


def test_fused_adamw_is_a_torch_optimizer():
    """train_step1.py:36-38 / train_step2.py:35 hand the optimizer to LinearLR / ReduceLROnPlateau, which
    reject anything that is not a torch.optim.Optimizer; the scheduled lr must reach the kernel; optimizer state must
    round-trip in torch.optim.AdamW's own format."""
    from nconv_b200 import ops
    from nconv_b200 import train_utils as U
    torch.manual_seed(0)
    # (plain matmul layers: the comparison with torch.optim.AdamW over several steps needs bit-reproducible gradients,
    #  which cuDNN's convolution backward does not promise)
    net = nn.Sequential(nn.Flatten(), nn.Linear(3 * 12 * 12, 32), nn.Tanh(), nn.Linear(32, 4)).to(dev())
    ref = copy.deepcopy(net)
    opt = U.get_optimizer(net, "adam", 1e-2, 1e-3)
    assert isinstance(opt, torch.optim.Optimizer) and isinstance(opt, ops.FusedAdamW)
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=1e-3)
    s = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1.0, end_factor=0.0, total_iters=4)
    rs = torch.optim.lr_scheduler.LinearLR(ropt, start_factor=1.0, end_factor=0.0, total_iters=4)
    torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=0.1, patience=2)      # must construct
    x = torch.randn(2, 3, 12, 12, device=dev())
    for it in range(3):
        for n_, o_ in ((net, opt), (ref, ropt)):
            o_.zero_grad()
            n_(x).square().mean().backward()
            o_.step()
        s.step(); rs.step()
        assert abs(s.get_last_lr()[0] - rs.get_last_lr()[0]) < 1e-12
        for p, q in zip(net.parameters(), ref.parameters()):
            assert O.max_rel_err(p, q, floor=1e-3) <= 5e-6, it
    # checkpoint round trip: into a fresh FusedAdamW and into a stock AdamW
    sd = opt.state_dict()
    net2 = copy.deepcopy(net)
    opt2 = ops.FusedAdamW(net2.parameters(), lr=1.0)
    opt2.load_state_dict(sd)
    stock = torch.optim.AdamW(copy.deepcopy(net).parameters(), lr=1.0)
    stock.load_state_dict(sd)
    assert stock.param_groups[0]["lr"] == opt.param_groups[0]["lr"]
    for n_, o_ in ((net, opt), (net2, opt2)):
        o_.zero_grad()
        n_(x).square().mean().backward()
        o_.step()
    for p, q in zip(net.parameters(), net2.parameters()):
        assert torch.equal(p, q)


def test_streamed_inference_delivers_every_batch():
    """run() kept only the last `depth` results."""
    from nconv_b200.pipeline import StreamedInference
    model = lambda a, b: (a * 2 + b, a - b)             # noqa: E731
    batches = [[torch.full((4, 8), float(i)).pin_memory(), torch.full((4, 8), 0.5 * i).pin_memory()] for i in range(5)]
    out = StreamedInference(model, dev()).run(batches)
    assert len(out) == 5
    for i, (y0, y1) in enumerate(out):
        assert torch.all(y0 == 2.5 * i) and torch.all(y1 == 0.5 * i)
    seen = StreamedInference(model, dev(), depth=3).run(batches, consume=lambda i, outs: (i, float(outs[0][0, 0])))
    assert seen == [(i, 2.5 * i) for i in range(5)]


def test_graphed_inference_matches_eager_and_replays():
    """pipeline.GraphedInference: the SETP1+SETP2 EXPORT forward captured in a CUDA graph gives bit-identical outputs to the
    eager call, for the captured inputs and for new inputs of the same shape."""
    from models.step2 import SETP2_BP_EXPORT
    from nconv_b200.pipeline import GraphedInference
    from nconv_b200.synthetic import synthetic_state_dict
    m = SETP2_BP_EXPORT()
    m.load_state_dict(synthetic_state_dict(5), strict=False)
    m = m.eval().to(dev())
    a = [t.to(dev()) for t in setp2_inputs(2, seed=4, H=128, W=256)]
    b = [t.to(dev()) for t in setp2_inputs(2, seed=9, H=128, W=256)]
    with torch.no_grad():
        ea = [o.clone() for o in m(*a)]
        eb = [o.clone() for o in m(*b)]
    g = GraphedInference(m, a)
    for inp, ref in ((a, ea), (b, eb), (a, ea)):
        out = g(*inp)
        torch.cuda.synchronize()
        assert torch.equal(out[0], ref[0]) and torch.equal(out[1], ref[1])
