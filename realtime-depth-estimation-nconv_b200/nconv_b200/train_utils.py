"""Caller-side helpers of the training loop, same names and argument meaning as /root/reference/utils.py,
running on the nconv_b200 CUDA kernels.  CUDA only: there is no CPU path.

  calculate_loss                   utils.py:138-151   masked MSE, or 0.8*RMSE + 0.2*Sobel-L1 (:95-136) with
                                                      use_gradient_loss=True (train_step1.py:20)
  calculate_loss_multi_resolution  utils.py:63-71     mean over scales of the loss of the bilinear upsample
  get_optimizer                    utils.py:53-61     'adam' -> AdamW as one fused kernel over a flat buffer

OPT-IN.  The reference's scripts do `from utils import *` (train_step1.py:3, train_step2.py:3) and need the rest of
that module too (`get_performance`, `save_checkpoint`, `save_depth`, the star-leaked `os` / `DataLoader`), so the
unchanged scripts keep importing the reference's own `utils.py` — with which the drop-in models work as they are
(stock `F.mse_loss` / `torch.optim.AdamW` on CUDA tensors).  To route the loss and the optimizer through the
library's kernels a maintainer adds ONE line after that import:

    from utils import *
    from nconv_b200.train_utils import *      # calculate_loss, calculate_loss_multi_resolution, get_optimizer

This module deliberately is not named `utils.py`: it must never shadow the reference's module on sys.path.
"""
from __future__ import annotations

import torch

from . import ops

__all__ = ["calculate_loss", "calculate_loss_multi_resolution", "get_optimizer"]


def _as4(t: torch.Tensor) -> torch.Tensor:
    if t.dim() == 4:
        return t
    if t.dim() == 3:            # [1,H,W]: one sample, as utils.py:69 passes it
        return t.unsqueeze(0)
    if t.dim() == 2:
        return t.unsqueeze(0).unsqueeze(0)
    raise ValueError("expected a [N,1,H,W], [1,H,W] or [H,W] tensor, got %s" % (tuple(t.shape),))


def calculate_loss(reconstructed_img, target_img, use_gradient_loss=False):
    """F.mse_loss(reconstructed.masked_fill(target == 0, 0), target) — utils.py:138-151; with
    use_gradient_loss: 0.8*sqrt(that) + 0.2*gradient_loss(target, masked reconstruction) (utils.py:121-136)."""
    r, t = _as4(reconstructed_img), _as4(target_img)
    if r.shape != t.shape:
        raise ValueError("calculate_loss: shapes differ: %s vs %s" % (tuple(r.shape), tuple(t.shape)))
    if r.shape[1] != 1:
        if use_gradient_loss:   # the reference's 1-channel Sobel kernel rejects it too (F.conv2d channel mismatch)
            raise ValueError("calculate_loss: use_gradient_loss needs single-channel images")
        # plain MSE: every element is its own pixel, fold channels into the batch axis
        r = r.reshape(-1, 1, r.shape[2], r.shape[3]); t = t.reshape(-1, 1, t.shape[2], t.shape[3])
    return ops.depth_loss_multires([r], t, use_gradient_loss=use_gradient_loss)


def calculate_loss_multi_resolution(reconstructed_img, target_img, use_gradient_loss=False, all_samples=False):
    """utils.py:63-71: every scale is upsampled (bilinear, align_corners=False) to the target's size and
    compared with the target; the mean over scales is returned.  Like the reference only sample 0 of the
    batch enters the loss (utils.py:69 indexes [0]); all_samples=True takes the mean over the whole batch
    instead.  The reference hard-codes size=(480, 640); here it is the target's own size."""
    t = _as4(target_img)
    return ops.depth_loss_multires([_as4(d) for d in reconstructed_img], t, rows=None if all_samples else 1,
                                   use_gradient_loss=use_gradient_loss)


def get_optimizer(net, optim_type, lr, weight_decay):
    """utils.py:53-61.  'adam' returns nconv_b200.ops.FusedAdamW (torch.optim.AdamW's update rule with its
    default betas/eps); 'sgd' and 'rmsprop' are plain torch.optim objects as in the reference."""
    if optim_type == "adam":
        return ops.FusedAdamW(net.parameters(), lr=lr, weight_decay=weight_decay)
    if optim_type == "sgd":
        return torch.optim.SGD(net.parameters(), lr=lr, weight_decay=weight_decay, momentum=0.9)
    if optim_type == "rmsprop":
        return torch.optim.RMSprop(net.parameters(), lr=lr, weight_decay=weight_decay, momentum=0.9)
    raise ValueError("Unsupported optimizer type. Choose 'adam', 'sgd', or 'rmsprop'.")
