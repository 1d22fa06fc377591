"""Export-time decomposition of the `nconv_b200::*` custom ops into ATen.

/root/reference/export_to_onnx.py:58-74 hands `SETP2_BP_EXPORT` to `torch.onnx.export` (opset 17, four inputs, two
outputs, dynamic batch).  An exporter cannot look inside a ctypes launch, so every op the EXPORT graph reaches has a
restatement here in plain ATen calls — the ops ONNX knows as Conv / ConvTranspose / Mul / Div / Add / Relu / MaxPool /
Resize / Concat / Slice / Pad — used in two ways:

* `torch.export` / `torch.onnx.export(dynamo=True)`: the module is traced with the custom ops as opaque nodes (their
  `register_fake` shape functions), then `export_decomposed()` runs `ExportedProgram.run_decompositions` with
  `DECOMPOSITIONS` (custom op overload -> function below) and returns a pure-ATen program;
* the TorchScript tracer (`torch.onnx.export(dynamo=False)`, `torch.jit.trace`): while tracing, the op namespace the
  models call through (`ops.K`) resolves to the functions below directly, so the trace records ATen nodes.

This is NOT a runtime path: outside tracing/export nothing here is called, and the models keep failing loudly on CPU
tensors.  The ONNX serialisation itself needs the `onnx` / `onnxscript` packages, which this image does not have; what
is tested (tests/test_export_cpu.py) is that both routes yield graphs free of `nconv_b200` nodes whose CPU execution
reproduces the reference's golden output.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from . import ops as _ops  # noqa: F401  (registers the op schemas)

_EPS_LAYERS = (  # (kernel padding) of DNET's nine layers, models/step1.py:38-49
    (2, 2), (2, 2), (2, 2), (2, 2), (2, 2), (1, 1), (1, 1), (0, 0), (2, 2))


def nconv2d(data, conf, weight, bias, ph: int, pw: int, eps: float):
    """NConv2d.forward, step1.py:116-149."""
    denom = F.conv2d(conf, weight, None, 1, (ph, pw))
    nomin = F.conv2d(data * conf, weight, None, 1, (ph, pw))
    y = nomin / (denom + eps) + bias.view(1, -1, 1, 1)
    s = weight.reshape(weight.shape[0], -1).sum(dim=-1).view(1, -1, 1, 1)
    return y, denom / s


def confidence_mask(s, thr: float):
    return (s > thr).to(s.dtype)                                   # step1.py:53


def maxpool2x2(x):
    return F.max_pool2d(x, 2, 2)                                   # step1.py:62-74


def pool_conf_argmax(x, c):
    co, idx = F.max_pool2d(c, 2, 2, return_indices=True)
    xo = torch.gather(x.flatten(2), 2, idx.flatten(2)).view_as(co)
    return xo, co / 4


def upcat(skip, low, skip_first: bool):
    up = F.interpolate(low, skip.shape[2:], mode="nearest")        # step1.py:78-89
    return torch.cat((skip, up), 1) if skip_first else torch.cat((up, skip), 1)


def dnet_forward(S, params: List[torch.Tensor], fp32_only: bool = False):
    """DNET.forward, step1.py:51-94, crop generalised to [1:H+1, 1:W+1]."""
    def layer(i, x, c):
        return nconv2d(x, c, params[2 * i], params[2 * i + 1], _EPS_LAYERS[i][0], _EPS_LAYERS[i][1], 1e-7)

    H, W = S.shape[2], S.shape[3]
    c0 = confidence_mask(S, 0.01)
    x1, c1 = layer(0, S, c0)
    x1, c1 = layer(1, x1, c1)
    x2, c2 = layer(2, maxpool2x2(x1), maxpool2x2(c1))
    x3, c3 = layer(3, maxpool2x2(x2), maxpool2x2(c2))
    x4, c4 = layer(4, maxpool2x2(x3), maxpool2x2(c3))
    x34, c34 = layer(5, upcat(x3, x4, True), upcat(c3, c4, True))
    x23, c23 = layer(6, upcat(x2, x34, True), upcat(c2, c34, True))
    x6, c6 = layer(7, upcat(x1, x23, False), upcat(c1, c23, False))
    x7, _ = layer(8, x6, c6)
    return x7[:, :, 1:H + 1, 1:W + 1]


def conv2d(a, b, weight, scale, shift, residual, stride: int, relu: bool):
    """act(scale * conv(cat[a, b]) + shift) + residual: ConvBlock / RGBEncoder / Basic2d / the depth head (step2.py)."""
    x = a if b is None else torch.cat((a, b), 1)
    y = F.conv2d(x, weight, None, stride, weight.shape[2] // 2)
    if scale is not None:
        y = y * scale.view(1, -1, 1, 1)
    if shift is not None:
        y = y + shift.view(1, -1, 1, 1)
    if relu:
        y = F.relu(y)
    if residual is not None:
        y = y + residual
    return y


def convT4x4s2(a, b, weight, scale, shift, relu: bool):
    """Basic2dTrans, step2.py:197-214: ConvTranspose2d(4, 2, 1) + folded BatchNorm + ReLU."""
    x = a if b is None else torch.cat((a, b), 1)
    y = F.conv_transpose2d(x, weight, None, 2, 1)
    if scale is not None:
        y = y * scale.view(1, -1, 1, 1)
    if shift is not None:
        y = y + shift.view(1, -1, 1, 1)
    return F.relu(y) if relu else y


def bilinear_ac(x, HO: int, WO: int):
    return F.interpolate(x, size=(HO, WO), mode="bilinear", align_corners=True)      # step2.py:252, 271


def bn_fold(gamma, beta, mean, var, conv_bias, eps: float):
    scale = gamma * torch.rsqrt(var + eps)
    shift = beta - mean * scale
    if conv_bias is not None:
        shift = shift + conv_bias * scale
    return scale, shift


def border_zeroed(x, top: int, bottom: int, left: int):
    """step2.py:122-124 as Slice + Pad (no in-place scatter in the exported graph)."""
    H, W = x.shape[2], x.shape[3]
    top, bottom, left = min(top, H), min(bottom, H), min(left, W)
    if top + bottom >= H or left >= W:
        return torch.zeros_like(x)
    return F.pad(x[:, :, top:H - bottom, left:], (left, 0, top, bottom))


# op name -> ATen restatement (what `ops.K.<name>` resolves to while a TorchScript trace is being recorded)
ATEN = {
    "nconv2d": nconv2d, "confidence_mask": confidence_mask, "maxpool2x2": maxpool2x2, "upcat": upcat,
    "pool_conf_argmax": pool_conf_argmax,
    "dnet_forward": dnet_forward, "conv2d": conv2d, "convT4x4s2": convT4x4s2, "bilinear_ac": bilinear_ac,
    "bn_fold": bn_fold, "border_zeroed": border_zeroed,
}

# custom op overload -> ATen restatement, for ExportedProgram.run_decompositions
DECOMPOSITIONS: Dict[object, object] = {getattr(torch.ops.nconv_b200, name).default: fn for name, fn in ATEN.items()}


def export_decomposed(model: torch.nn.Module, args, dynamic_batch: bool = False):
    """torch.export the module (custom ops opaque, shapes from their fake functions), then decompose every
    `nconv_b200::*` node into ATen.  Returns an ExportedProgram with no custom op left: the input a downstream
    exporter (`torch.onnx.export(ep, ...)`) or any ATen backend needs.  `dynamic_batch` marks dim 0 of every
    input as one shared symbolic size (export_to_onnx.py:66-73)."""
    dyn = None
    if dynamic_batch:
        b = torch.export.Dim("batch_size", min=1)
        dyn = tuple({0: b} for _ in args)
    was_training = model.training
    model.eval()
    try:
        with torch.no_grad():
            ep = torch.export.export(model, tuple(args), dynamic_shapes=dyn, strict=False)
    finally:
        model.train(was_training)
    table = torch.export.default_decompositions()
    table.update(DECOMPOSITIONS)
    ep = ep.run_decompositions(table)
    left = custom_nodes(ep.graph_module)
    if left:
        raise RuntimeError("export_decomposed: undecomposed custom ops remain: %s" % sorted(set(left)))
    return ep


def custom_nodes(gm) -> List[str]:
    """Names of nconv_b200 ops still present in a torch.fx graph module."""
    return [str(n.target) for n in gm.graph.nodes if n.op == "call_function" and "nconv_b200" in str(n.target)]


def to_onnx(model: torch.nn.Module, args, path: str, dynamic_batch: bool = True, **kw):
    """export_to_onnx.py:58-74 for the drop-in model.  Needs the `onnx` + `onnxscript` packages (absent from this image:
    untested here); the decomposed program above is what gets serialised."""
    ep = export_decomposed(model, args, dynamic_batch=dynamic_batch)
    return torch.onnx.export(ep, tuple(args), path, dynamo=True, **kw)
