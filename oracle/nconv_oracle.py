"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's NConv hot path.

This file is the *checker*: only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
`cpu_baseline` / `--impl reference` legs may import it.  The product path
(`realtime-depth-estimation-nconv_b200/`) never does and has no CPU fallback.

What it is: a functional (state_dict in, tensors out) fp32 restatement, in plain
`torch` CPU ops, of
  * `NConv2d.forward`            /root/reference/models/step1.py:116-149
  * `EnforcePos._pos` (softplus) /root/reference/models/step1.py:197-207
  * `DNET.forward`               /root/reference/models/step1.py:51-94
  * `SETP2_BP_TRAIN/EXPORT.forward` and their blocks
                                 /root/reference/models/step2.py:60-77, 105-126, 134-297
  * `calculate_loss`, `calculate_loss_multi_resolution`, AdamW of `get_optimizer`
                                 /root/reference/utils.py:138-151, 63-71, 53-55
The arithmetic below the reference's call sites lives in ATen (torch 2.11.0 in this image;
the authors pinned pytorch 2.2.2, environment.yml:240) — a third-party dependency that is
not under /root/reference.  The restatement therefore issues the same ATen calls in the
same order as the reference's own lines, and is PINNED by `tests/golden/*.npz`, which were
produced by running the unmodified reference here (`oracle/gen_golden.py`).  The reference
repo itself holds no tests/golden vectors, so that is the only pin there is.

Two deliberate generalisations, both identical to the reference at 480x640:
  * crop: the reference hard-codes `[1:481, 1:641]` (step1.py:94); `crop="general"` crops
    `[1:H+1, 1:W+1]`.  `crop="reference"` reproduces the hard-coded slice.
  * SETP1 accepts two depth tensors stacked on the batch axis.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

EPS = 1e-7  # step1.py:103


def softplus_beta10(w: torch.Tensor) -> torch.Tensor:
    """step1.py:206-207 — F.softplus(p, beta=10) (ATen threshold 20: identity if 10*w > 20)."""
    return F.softplus(w, beta=10)


def nconv2d(data, conf, weight, bias, padding, stride=(1, 1), dilation=(1, 1), groups=1,
            eps=EPS):
    """step1.py:116-149, line by line."""
    denom = F.conv2d(conf, weight, None, stride, padding, dilation, groups)          # :119
    nomin = F.conv2d(data * conf, weight, None, stride, padding, dilation, groups)   # :121
    nconv = nomin / (denom + eps)                                                     # :123
    nconv = nconv + bias.view(1, -1, 1, 1)                                            # :127-131
    s = weight.reshape(weight.shape[0], -1).sum(dim=-1, keepdim=True)                 # :141-144
    sz = denom.shape
    cout = (denom.reshape(sz[0], sz[1], -1) / s).reshape(sz)                          # :146-147
    return nconv, cout


_DNET_LAYERS = {  # name: padding   (step1.py:38-49; nconv7 takes the ctor default (2,2), :98)
    "nconv1": (2, 2), "nconv2": (2, 2), "nconv_down1": (2, 2), "nconv_down2": (2, 2),
    "nconv_down3": (2, 2), "nconv4": (1, 1), "nconv5": (1, 1), "nconv6": (0, 0),
    "nconv7": (2, 2),
}


def dnet_forward(S, sd, prefix="d_net.", crop="general", return_all=False):
    """step1.py:51-94.  `sd` maps '<prefix><layer>.weight/bias' to tensors (weights used
    as stored — the EnforcePos hook is the caller's business, see `apply_enforce_pos`)."""

    def layer(name, x, c):
        return nconv2d(x, c, sd[prefix + name + ".weight"], sd[prefix + name + ".bias"],
                       _DNET_LAYERS[name])

    c0 = (S > 0.01).to(S.dtype)                               # :53 `.float()`; same for fp32 inputs, lets an fp64 checker run stay fp64
    x0 = S                                                    # :54
    x1, c1 = layer("nconv1", x0, c0)                          # :56
    x1, c1 = layer("nconv2", x1, c1)                          # :57
    c1_ds = F.max_pool2d(c1, 2, 2)                            # :62
    x1_ds = F.max_pool2d(x1, 2, 2)                            # :64
    x2_ds, c2_ds = layer("nconv_down1", x1_ds, c1_ds)         # :65
    c2_dss = F.max_pool2d(c2_ds, 2, 2)                        # :68
    x2_dss = F.max_pool2d(x2_ds, 2, 2)                        # :69
    x3_ds, c3_ds = layer("nconv_down2", x2_dss, c2_dss)       # :70
    c3_dss = F.max_pool2d(c3_ds, 2, 2)                        # :73
    x3_dss = F.max_pool2d(x3_ds, 2, 2)                        # :74
    x4_ds, c4_ds = layer("nconv_down3", x3_dss, c3_dss)       # :75
    x4 = F.interpolate(x4_ds, c3_ds.shape[2:], mode="nearest")     # :78
    c4 = F.interpolate(c4_ds, c3_ds.shape[2:], mode="nearest")     # :79
    x34_ds, c34_ds = layer("nconv4", torch.cat((x3_ds, x4), 1), torch.cat((c3_ds, c4), 1))  # :80
    x34 = F.interpolate(x34_ds, c2_ds.shape[2:], mode="nearest")   # :83
    c34 = F.interpolate(c34_ds, c2_ds.shape[2:], mode="nearest")   # :84
    x23_ds, c23_ds = layer("nconv5", torch.cat((x2_ds, x34), 1), torch.cat((c2_ds, c34), 1))  # :85
    x23 = F.interpolate(x23_ds, x0.shape[2:], mode="nearest")      # :88
    c23 = F.interpolate(c23_ds, c0.shape[2:], mode="nearest")      # :89
    x6, c6 = layer("nconv6", torch.cat((x23, x1), 1), torch.cat((c23, c1), 1))  # :90
    x7, c7 = layer("nconv7", x6, c6)                          # :92
    H, W = S.shape[2:]
    if crop == "reference":
        out = x7[:, :, 1:481, 1:641]                          # :94
    else:
        out = x7[:, :, 1:H + 1, 1:W + 1]
    if return_all:
        return out, dict(x1=x1, c1=c1, x2_ds=x2_ds, c2_ds=c2_ds, x3_ds=x3_ds, c3_ds=c3_ds,
                         x4_ds=x4_ds, c4_ds=c4_ds, x34_ds=x34_ds, c34_ds=c34_ds,
                         x23_ds=x23_ds, c23_ds=c23_ds, x6=x6, c6=c6, x7=x7, c7=c7)
    return out


def pool_conf_argmax(x, c):
    """Opt-in upstream-nconv pooling (NOT in the reference, which pools independently at step1.py:62-74; SURVEY 8f
    rank 4): confidence max-pooled with indices, data gathered at those indices, pooled confidence / 4."""
    co, idx = F.max_pool2d(c, 2, 2, return_indices=True)
    xo = torch.gather(x.flatten(2), 2, idx.flatten(2)).view_as(co)
    return xo, co / 4


def sparsify(raw, mask=None, noise=None):
    """dataset/nyuloader.py:83-124 given the random draws: noisy points `d += d * noise` (:107), then `d * mask` (:113)
    or the zeroed random points (:121-123), which is a 0/1 mask as well."""
    d = raw
    if noise is not None:
        d = d + d * noise
    if mask is not None:
        d = d * mask.to(d.dtype)
    return d


def apply_enforce_pos(sd, prefix="d_net."):
    """What one training-mode forward does to the stored NConv weights (step1.py:190-195):
    weight.data <- softplus(weight, beta=10), in place, for all nine layers."""
    out = dict(sd)
    for name in _DNET_LAYERS:
        k = prefix + name + ".weight"
        out[k] = softplus_beta10(sd[k])
    return out


# --------------------------------------------------------------------------------------
# SETP2 (step2.py)
# --------------------------------------------------------------------------------------
def _bn(x, sd, p, training, eps=1e-5):
    if training:
        return F.batch_norm(x, None, None, sd[p + ".weight"], sd[p + ".bias"], True, 0.1, eps)
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"],
                        sd[p + ".weight"], sd[p + ".bias"], False, 0.1, eps)


def _rgb_encoder(x, sd, p, stride, bn_training):
    """step2.py:134-154: ReLU(BN(conv3x3_s(x)+b)) + conv1x1_s(x)."""
    out = F.conv2d(x, sd[p + ".encoder.0.weight"], sd[p + ".encoder.0.bias"], stride, 1)
    out = F.relu(_bn(out, sd, p + ".encoder.1", bn_training))
    return out + F.conv2d(x, sd[p + ".downsample.0.weight"], None, stride, 0)


def _conv_block(x, sd, p):
    """step2.py:290-297: ReLU(conv3x3 p1 + bias)."""
    return F.relu(F.conv2d(x, sd[p + ".conv.weight"], sd[p + ".conv.bias"], 1, 1))


def _new_fusion(rgb, depth, sd, p):
    """step2.py:216-236."""
    rgb_feat = _conv_block(rgb, sd, p + ".rgb_conv")
    depth_feat = _conv_block(depth, sd, p + ".depth_conv")
    fused = torch.cat((rgb_feat, depth_feat), 1)
    fused = _conv_block(fused, sd, p + ".fuse_conv1")
    fused = _conv_block(fused, sd, p + ".fuse_conv2")
    return _conv_block(fused, sd, p + ".fuse_conv3")


def _fuse0(rgb, depth, sd, p, factor):
    """step2.py:262-278."""
    d = F.interpolate(depth, scale_factor=1 / factor, mode="bilinear", align_corners=True)
    f = _new_fusion(rgb, d, sd, p + ".fuse")
    return f, d + F.conv2d(f, sd[p + ".conv.weight"], None, 1, 1)


def _fuse_block(rgb, depth, prev_feat, prev_depth, sd, p, factor, bn_training):
    """step2.py:238-259 with UpCat :160-176, Basic2dTrans :197-214, Basic2d :178-195.
    (Call sites :72-74 pass (rgb, sparse, out_fusion, out_depth); the callee's parameter
    names are swapped but the data flow is cat([prev_depth(1ch), prev_feat(C)]).)"""
    u = F.conv_transpose2d(torch.cat([prev_depth, prev_feat], 1).contiguous(),
                           sd[p + ".upcat.upf.conv.weight"], None, 2, 1)
    u = F.relu(_bn(u, sd, p + ".upcat.upf.bn", bn_training))
    f0 = F.conv2d(torch.cat([u, rgb], 1), sd[p + ".upcat.conv.conv.conv.weight"], None, 1, 1)
    f0 = F.relu(_bn(f0, sd, p + ".upcat.conv.conv.bn", bn_training))
    d = F.interpolate(depth, scale_factor=1 / factor, mode="bilinear", align_corners=True)
    f = _new_fusion(f0, d, sd, p + ".fuse")
    return f, d + F.conv2d(f, sd[p + ".conv.weight"], None, 1, 1)


def setp2_forward(sd, rgb0, depth0, rgb1, depth1, export, bn_training=False,
                  crop="general"):
    """step2.py:60-77 (TRAIN) / :105-126 (EXPORT).  Returns the list of 4 multi-scale
    depths for the stacked batch (TRAIN; caller slices [:B]/[B:]) or the border-zeroed
    full-resolution depth (EXPORT)."""
    sparse = dnet_forward(torch.cat((depth0, depth1), 0), sd, prefix="step1.d_net.", crop=crop)
    rgb = torch.cat((rgb0, rgb1), dim=0)
    r0 = _rgb_encoder(rgb, sd, "rgb_encoder0", 1, bn_training)
    r1 = _rgb_encoder(r0, sd, "rgb_encoder1", 2, bn_training)
    r2 = _rgb_encoder(r1, sd, "rgb_encoder2", 2, bn_training)
    r3 = _rgb_encoder(r2, sd, "rgb_encoder3", 2, bn_training)
    f0, d0 = _fuse0(r3, sparse, sd, "fuse0", 8)
    f1, d1 = _fuse_block(r2, sparse, f0, d0, sd, "fuse1", 4, bn_training)
    f2, d2 = _fuse_block(r1, sparse, f1, d1, sd, "fuse2", 2, bn_training)
    f3, d3 = _fuse_block(r0, sparse, f2, d2, sd, "fuse3", 1, bn_training)
    if export:
        d3 = d3.clone()
        d3[:, :, :45, :] = 0          # :122
        d3[:, :, -45:, :] = 0         # :123
        d3[:, :, :, :20] = 0          # :124
        return d3
    return [d0, d1, d2, d3]


# --------------------------------------------------------------------------------------
# Training-loop callers (utils.py): masked multi-resolution MSE, AdamW
# --------------------------------------------------------------------------------------
def upsample_bilinear_af(x: torch.Tensor, H: int, W: int) -> torch.Tensor:
    """F.interpolate(x, size=(H, W), mode='bilinear', align_corners=False) — utils.py:67 — restated from
    ATen's upsample_bilinear2d (UpSample.h area_pixel_compute_source_index / compute_source_index_and_lambda):
    src = max((dst + 0.5) * in/out - 0.5, 0); i0 = floor(src); i1 = i0 + (i0 < in-1); lambda = src - i0.
    Works in the dtype of x (the scale is a float32 value in ATen for float32 input)."""
    h, w = x.shape[-2:]

    def axis(n_in, n_out):
        scale = torch.tensor(n_in / n_out, dtype=x.dtype)
        src = ((torch.arange(n_out, dtype=x.dtype) + 0.5) * scale - 0.5).clamp_min(0)
        i0 = src.floor().long().clamp_max(n_in - 1)
        i1 = i0 + (i0 < n_in - 1).long()
        lam = src - i0.to(x.dtype)
        return i0, i1, lam

    y0, y1, ly = axis(h, H)
    x0, x1, lx = axis(w, W)
    ly = ly.view(-1, 1)
    top = x[..., y0, :][..., x0] * (1 - lx) + x[..., y0, :][..., x1] * lx
    bot = x[..., y1, :][..., x0] * (1 - lx) + x[..., y1, :][..., x1] * lx
    return top * (1 - ly) + bot * ly


SOBEL_X = [[1., 0., -1.], [2., 0., -2.], [1., 0., -1.]]      # utils.py:99-103
SOBEL_Y = [[1., 2., 1.], [0., 0., 0.], [-1., -2., -1.]]      # utils.py:113-117


def gradient_loss(target, predicted):
    """utils.py:121-136: mean |sobel_x(target - predicted)| + mean |sobel_y(...)|, zero padding 1."""
    diff = target - predicted
    if diff.dim() == 3:
        diff = diff.unsqueeze(0)
    kx = torch.tensor([[SOBEL_X]], dtype=diff.dtype)
    ky = torch.tensor([[SOBEL_Y]], dtype=diff.dtype)
    return F.conv2d(diff, kx, padding=1).abs().mean() + F.conv2d(diff, ky, padding=1).abs().mean()


def calculate_loss(reconstructed, target, use_gradient_loss=False):
    """utils.py:138-151: mse over ALL elements after zeroing the prediction where target == 0;
    with use_gradient_loss: 0.8*sqrt(mse) + 0.2*gradient_loss(target, masked prediction)."""
    rec = reconstructed.masked_fill(target == 0, 0)
    mse = ((rec - target) ** 2).mean()
    if use_gradient_loss:
        return torch.sqrt(mse) * 0.8 + gradient_loss(target, rec) * 0.2
    return mse


def calculate_loss_multi_resolution(depths, target, all_samples=False, use_gradient_loss=False):
    """utils.py:63-71: upsample every scale to the target size, loss on sample 0 only ([0] at :69),
    mean over scales.  all_samples=True is the batch-mean generalisation used by the weak-scaling bench."""
    H, W = target.shape[-2:]
    total = 0.0
    for d in depths:
        up = upsample_bilinear_af(d, H, W)
        total = total + (calculate_loss(up, target, use_gradient_loss) if all_samples
                         else calculate_loss(up[0], target[0], use_gradient_loss))
    return total / len(depths)


def adamw_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2):
    """torch.optim.AdamW single-tensor update (torch/optim/adam.py _single_tensor_adam with
    decoupled_weight_decay; utils.py:55 constructs it with torch defaults).  Returns new (p, m, v)."""
    p = p * (1 - lr * weight_decay)
    m = m + (g - m) * (1 - beta1)
    v = v * beta2 + g * g * (1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = v.sqrt() / (bc2 ** 0.5) + eps
    return p - (lr / bc1) * (m / denom), m, v


# --------------------------------------------------------------------------------------
# Tolerance used by every parity test (BASELINE.json north_star: max-rel 1e-4 in fp32)
# --------------------------------------------------------------------------------------
def max_rel_err(a: torch.Tensor, b: torch.Tensor, floor: float = 1e-3) -> float:
    """max |a-b| / max(|b|, floor).  The floor keeps exact-zero confidence regions
    meaningful."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    if a.numel() == 0:
        return 0.0
    return float(((a - b).abs() / b.abs().clamp_min(floor)).max())


# --------------------------------------------------------------------------------------
# Synthetic weights for bench.py (both arms load the same tensors; no checkpoint exists offline)
# --------------------------------------------------------------------------------------
def synthetic_state_dict(seed: int = 5):
    """The input generator lives with the other seeded synthetic inputs in
    `nconv_b200/synthetic.py` (so the GPU arm of bench.py never imports this module); the CPU arm
    and the tests reach it through this alias."""
    from nconv_b200.synthetic import synthetic_state_dict as gen
    return gen(seed)
