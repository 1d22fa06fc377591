"""CPU: export-time decomposition of the custom ops (/root/reference/export_to_onnx.py:24-74).

`SETP2_BP_EXPORT` is captured on CPU tensors — no kernel runs: torch.export sees the custom ops through their fake
shape functions, the TorchScript tracer sees their ATen restatements — and the captured graph, which must not
contain a single `nconv_b200` node, is then executed by ATen on the CPU and compared with the reference's golden
output (tests/golden/setp2_export.npz) at the SETP2 tolerances of tests/test_gpu_parity.py.  The final `onnx`
serialisation is not exercised: the `onnx` / `onnxscript` packages are not in this image."""
import numpy as np
import pytest
import torch

from conftest import REL_TOL, T
from oracle import nconv_oracle as O
from nconv_b200.synthetic import randomise_bn_stats, setp2_inputs


def _golden_model():
    from models.step2 import SETP2_BP_EXPORT
    torch.manual_seed(5)
    m = SETP2_BP_EXPORT()
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    m.load_state_dict(O.apply_enforce_pos(sd, prefix="step1.d_net."))
    m.eval()
    randomise_bn_stats(m, seed=55)
    return m


def _check(o0, o1, g):
    for o, key in ((o0, "out0"), (o1, "out1")):
        ref = T(g[key])
        assert o.shape == ref.shape
        assert O.max_rel_err(o, ref, floor=0.1) <= REL_TOL
        assert float((o - ref).abs().max() / ref.abs().max()) <= 1e-5
        noise = float(g["ref32_vs_64_rel_floor1e3_" + key[-1]])
        assert O.max_rel_err(o, ref, floor=1e-3) <= max(1e-4, 2.5 * noise)
    assert torch.all(o0[:, :, :45] == 0) and torch.all(o0[:, :, -45:] == 0) and torch.all(o0[..., :20] == 0)


def test_decomposition_table_covers_every_op_of_the_export_graph():
    from nconv_b200 import export as E
    import nconv_b200.ops  # noqa: F401
    for name in ("dnet_forward", "conv2d", "convT4x4s2", "bilinear_ac", "bn_fold", "border_zeroed", "nconv2d",
                 "confidence_mask", "maxpool2x2", "upcat"):
        assert getattr(torch.ops.nconv_b200, name).default in E.DECOMPOSITIONS


def test_torch_export_decomposes_to_pure_aten_and_matches_golden(golden):
    from nconv_b200 import export as E
    g = golden("setp2_export.npz")
    m = _golden_model()
    inputs = setp2_inputs(1, seed=7)
    # the opaque capture really goes through the custom ops (that is what the decomposition table has to remove)
    with torch.no_grad():
        raw = torch.export.export(m, tuple(inputs), strict=False)
    seen = set(E.custom_nodes(raw.graph_module))
    assert any("dnet_forward" in s for s in seen) and any("conv2d" in s for s in seen) and any("convT4x4s2" in s for s in seen)
    ep = E.export_decomposed(m, inputs)
    assert E.custom_nodes(ep.graph_module) == []
    targets = {str(n.target) for n in ep.graph_module.graph.nodes if n.op == "call_function"}
    for op in ("aten.convolution", "aten.max_pool2d", "aten.div", "aten.cat", "aten.relu"):
        assert any(op in t for t in targets), (op, sorted(targets))
    with torch.no_grad():
        o0, o1 = ep.module()(*inputs)
    _check(o0, o1, g)


def test_torch_export_dynamic_batch():
    """export_to_onnx.py:66-73 marks the batch axis of all four inputs dynamic: the decomposed program must accept a
    batch it was not traced with.  Small frames: this checks shapes and frame independence, not arithmetic."""
    from nconv_b200 import export as E
    m = _golden_model()
    ep = E.export_decomposed(m, setp2_inputs(2, seed=3, H=128, W=64), dynamic_batch=True)
    mod = ep.module()
    a = setp2_inputs(3, seed=4, H=128, W=64)
    with torch.no_grad():
        o0, o1 = mod(*a)
        p0, p1 = mod(*[t[1:2] for t in a])
    assert o0.shape == (3, 1, 128, 64) and o1.shape == (3, 1, 128, 64)
    assert torch.allclose(o0[1:2], p0, rtol=0, atol=1e-5) and torch.allclose(o1[1:2], p1, rtol=0, atol=1e-5)


def test_torchscript_trace_records_aten_only_and_matches_golden(golden):
    """The route torch.onnx.export(dynamo=False) takes: while tracing, ops.K resolves to the ATen restatements."""
    g = golden("setp2_export.npz")
    m = _golden_model()
    inputs = setp2_inputs(1, seed=7)
    with torch.no_grad():
        traced = torch.jit.trace(m, tuple(inputs), check_trace=False)
    text = str(traced.inlined_graph)
    assert "nconv_b200::" not in text and ("aten::_convolution" in text or "aten::conv2d" in text)
    with torch.no_grad():
        o0, o1 = traced(*inputs)
    _check(o0, o1, g)


def test_untraced_cpu_call_still_refuses():
    """The decompositions are export-time only: an ordinary call on CPU tensors still fails loudly."""
    m = _golden_model()
    with torch.no_grad(), pytest.raises((RuntimeError, NotImplementedError)):
        m(*setp2_inputs(1, seed=7, H=64, W=64))
