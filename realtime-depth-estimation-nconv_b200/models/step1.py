"""Drop-in for the reference's `models/step1.py` — same import path, class names, constructor
signatures, `forward` contracts, RNG consumption at construction and `state_dict` keys — with
the arithmetic done by the hand-written sm_100a kernels behind `torch.ops.nconv_b200.*`.

Reference: /root/reference/models/step1.py
  SETP1_NCONV :15-27     DNET :30-94     NConv2d :97-172     EnforcePos :176-212

Differences from the reference, all conscious:
  * the output crop `[1:481, 1:641]` (:94) is generalised to `[1:H+1, 1:W+1]` — identical at
    480x640;
  * `SETP1_NCONV.forward(x0_d, x1_d=None)` also accepts the two-tensor call that
    `models/step2.py:62,107` and `void_train_step1.py:61` make (frames stacked on the batch axis);
  * tensors must live on a CUDA device: there is no CPU path.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.nn as nn
from scipy.stats import poisson
from torch.nn.modules.conv import _ConvNd

from nconv_b200 import ops as _ops  # noqa: F401  (registers torch.ops.nconv_b200.*)

_K = _ops.K          # torch.ops.nconv_b200 (ATen restatements while a TorchScript trace is recorded: nconv_b200/export.py)


class SETP1_NCONV(nn.Module):
    """step1.py:15-27.  forward(raw_depth [B,1,H,W]) -> estimated depth [B,1,H,W]."""

    def __init__(self):
        super().__init__()
        self.d_net = DNET(32)

    def forward(self, x0_d, x1_d=None):
        if x1_d is not None:
            x0_d = torch.cat((x0_d, x1_d), 0)
        return self.d_net(x0_d)

    def forward_with_confidence(self, x0_d):
        """Opt-in (not in the reference, which drops the last confidence at step1.py:92-94):
        returns (depth, confidence of nconv7) with the same crop."""
        return self.d_net(x0_d, return_confidence=True)


class DNET(nn.Module):
    """step1.py:30-94."""

    _ORDER = ("nconv1", "nconv2", "nconv_down1", "nconv_down2", "nconv_down3",
              "nconv4", "nconv5", "nconv6", "nconv7")

    def __init__(self, out_ch, pool="independent"):
        super().__init__()
        # pool: "independent" = the reference (two separate F.max_pool2d, step1.py:62-74);
        # "conf_argmax" = opt-in upstream-nconv behaviour (data gathered at the confidence argmax, pooled confidence / 4):
        # NOT what the reference computes, never the default
        if pool not in ("independent", "conf_argmax"):
            raise ValueError("DNET: pool must be 'independent' or 'conf_argmax'")
        self.pool = pool
        pos_fn = "softplus"
        nc = 8
        self.nconv1 = NConv2d(1, nc, (5, 5), pos_fn, 'p', padding=(2, 2))
        self.nconv2 = NConv2d(nc, nc, (5, 5), pos_fn, 'p', padding=(2, 2))
        self.nconv_down1 = NConv2d(nc, nc, (5, 5), pos_fn, 'p', padding=(2, 2))
        self.nconv_down2 = NConv2d(nc, nc, (5, 5), pos_fn, 'p', padding=(2, 2))
        self.nconv_down3 = NConv2d(nc, nc, (5, 5), pos_fn, 'p', padding=(2, 2))
        self.nconv4 = NConv2d(2 * nc, nc, (3, 3), pos_fn, 'p', padding=(1, 1))
        self.nconv5 = NConv2d(2 * nc, nc, (3, 3), pos_fn, 'p', padding=(1, 1))
        self.nconv6 = NConv2d(2 * nc, nc, (3, 3), pos_fn, 'p', padding=(0, 0))
        self.nconv7 = NConv2d(nc, 1, (1, 1), pos_fn, 'k')

    def _layers(self):
        return [getattr(self, n) for n in self._ORDER]

    def _fusable(self, S) -> bool:
        """The single-enqueue inference pipeline applies when nothing needs a gradient."""
        if torch.is_grad_enabled() and (S.requires_grad or any(
                p.requires_grad for l in self._layers() for p in (l.weight, l.bias))):
            return False
        return all(l.eps == 1e-7 for l in self._layers())

    def forward(self, S, return_confidence=False):
        if S.dim() != 4 or S.shape[1] != 1:
            raise ValueError("DNET expects a [B,1,H,W] sparse depth map")
        H, W = S.shape[2:]
        if not return_confidence and self.pool == "independent" and self._fusable(S):
            # One enqueue of 9 fused kernels.  The EnforcePos pre-hooks still fire (they mutate
            # the stored weights in training mode even when those are frozen, step1.py:190-195).
            for l in self._layers():
                for hook in l._forward_pre_hooks.values():
                    hook(l, ())
            params = []
            for l in self._layers():
                params += [l.weight, l.bias]
            # inference (no_grad) takes the tensor-core pipeline where the shape allows; a frozen step 1 inside a
            # recording (training) step stays on the fp32 CUDA-core pipeline
            fp32 = torch.is_grad_enabled() or bool(getattr(self, "fp32_only", False)) or bool(os.environ.get("NCONV_B200_FP32_TRUNK"))
            return _K.dnet_forward(S, params, fp32)

        # Staged graph (training): every stage is a custom op with an analytic backward.
        c0 = _K.confidence_mask(S, 0.01)
        x1, c1 = self.nconv1(S, c0)
        x1, c1 = self.nconv2(x1, c1)
        if self.pool == "conf_argmax":
            pool = _K.pool_conf_argmax
        else:
            pool = lambda x, c: (_K.maxpool2x2(x), _K.maxpool2x2(c))      # noqa: E731
        x2, c2 = self.nconv_down1(*pool(x1, c1))
        x3, c3 = self.nconv_down2(*pool(x2, c2))
        x4, c4 = self.nconv_down3(*pool(x3, c3))
        x34, c34 = self.nconv4(_K.upcat(x3, x4, True), _K.upcat(c3, c4, True))
        x23, c23 = self.nconv5(_K.upcat(x2, x34, True), _K.upcat(c2, c34, True))
        xo, co = self.nconv6(_K.upcat(x1, x23, False), _K.upcat(c1, c23, False))
        xo, co = self.nconv7(xo, co)
        out = xo[:, :, 1:H + 1, 1:W + 1]
        if return_confidence:
            return out, co[:, :, 1:H + 1, 1:W + 1]
        return out


class NConv2d(_ConvNd):
    """step1.py:97-172.  forward(data, conf) -> (nconv, cout), one fused kernel."""

    def __init__(self, in_channels, out_channels, kernel_size, pos_fn='softplus', init_method='k',
                 stride=(1, 1), padding=(2, 2), dilation=(1, 1), groups=1, bias=True):
        # _ConvNd.__init__ runs reset_parameters(), which consumes the global RNG exactly as the
        # reference's constructor does (RNG-identical initialisation).
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, False,
                         output_padding=(0, 0), groups=groups, bias=bias, padding_mode='zeros')
        self.eps = 1e-7
        self.pos_fn = pos_fn
        self.init_method = init_method
        self.init_parameters()
        # Unused by forward (call sites commented out at step1.py:133-134) but present in every
        # reference checkpoint, so they stay for state_dict compatibility.
        self.bnorm = nn.BatchNorm2d(out_channels)
        self.relu = nn.ReLU()
        if self.pos_fn is not None:
            EnforcePos.apply(self, 'weight', pos_fn)

    def forward(self, data, conf):
        if tuple(self.stride) != (1, 1) or tuple(self.dilation) != (1, 1) or self.groups != 1:
            raise NotImplementedError("NConv2d: only stride 1, dilation 1, groups 1 (all the reference uses)")
        return _K.nconv2d(data, conf, self.weight, self.bias, int(self.padding[0]), int(self.padding[1]),
                          float(self.eps))

    def init_parameters(self):
        if self.init_method == 'x':
            torch.nn.init.xavier_uniform_(self.weight)
        elif self.init_method == 'k':
            torch.nn.init.kaiming_uniform_(self.weight)
        elif self.init_method == 'p':
            # separable Poisson(mu=k/2) bump on 0..k-1, tiled over (out, in), plus U[0,1) noise
            k = self.kernel_size[0]
            pmf = poisson(k / 2).pmf(np.arange(0, k))
            bump = torch.Tensor(np.outer(pmf, pmf)).type_as(self.weight)
            w = bump.expand(self.out_channels, self.in_channels, k, k)
            self.weight.data = w + torch.rand(w.shape)
        self.bias = torch.nn.Parameter(torch.zeros(self.out_channels) + 0.01)


class EnforcePos(object):
    """step1.py:176-212: forward-pre-hook that, in training mode only, overwrites the stored
    parameter in place with pos_fn(parameter) (straight-through: no autograd involvement)."""

    def __init__(self, pos_fn, name):
        self.name = name
        self.pos_fn = pos_fn

    @staticmethod
    def apply(module, name, pos_fn):
        fn = EnforcePos(pos_fn, name)
        module.register_forward_pre_hook(fn)
        return fn

    def __call__(self, module, inputs):
        if module.training:
            weight = getattr(module, self.name)
            self._pos_(weight.data)

    def _pos_(self, w):
        mode = _ops.POS_MODES.get(self.pos_fn.lower())
        if mode is None:
            print('Undefined positive function!')
            return
        if not w.is_contiguous():
            raise RuntimeError("EnforcePos: parameter must be contiguous")
        group = w.shape[-1] * w.shape[-2] if w.dim() >= 2 else 1
        _K.enforce_pos_(w, mode, group)

    def _pos(self, p):
        """Out-of-place variant with the reference's signature (step1.py:197)."""
        q = p.detach().clone()
        self._pos_(q)
        return q
