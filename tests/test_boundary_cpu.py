"""CPU: the drop-in boundary — the C-ABI library loads and exports every symbol the header
declares, the module surface (class names, ctor signatures, state_dict keys, RNG-identical
init, parameter counts) matches the reference, and the product path refuses CPU tensors instead
of falling back.  No compute call is made here (there is no GPU in the container)."""
import ctypes
import hashlib
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, T
from oracle import ref_shim


def state_fingerprint(sd) -> str:
    """Same definition as oracle/gen_golden.py:state_fingerprint."""
    h = hashlib.sha256()
    for k in sorted(sd):
        v = sd[k]
        h.update(k.encode())
        a = (v.double() if v.is_floating_point() else v).detach().cpu().numpy()
        h.update(a.tobytes())
    return h.hexdigest()


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "nconv_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(nconv_b200_[a-zA-Z0-9_]+)\s*\(", txt)))


def test_cabi_exports_every_declared_symbol():
    from nconv_b200 import _lib
    syms = _header_symbols()
    assert len(syms) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(lib, s), "libnconv_b200.so does not export %s" % s
    # the ctypes table binds exactly the header's entry points
    assert sorted(_lib.SIGNATURES) == syms
    loaded = _lib.load()
    assert loaded.nconv_b200_abi_version() >= 1
    assert b"unsupported" in loaded.nconv_b200_error_string(100002)


def test_cabi_struct_layout_matches_header():
    """The ctypes mirrors of the two descriptor structs must have the C layout (13/7 pointers
    followed by ints/floats, natural alignment)."""
    from nconv_b200._lib import ConvDesc, FwdDesc
    assert ctypes.sizeof(FwdDesc) == 13 * 8 + 14 * 4          # 12 ints + 2 floats
    assert ctypes.sizeof(ConvDesc) == 7 * 8 + 9 * 4 + 4        # 9 ints, padded to 8


def test_setp1_surface_and_rng_identical_init(golden):
    from models.step1 import DNET, SETP1_NCONV, EnforcePos, NConv2d
    g = golden("setp1.npz")
    torch.manual_seed(0)
    m = SETP1_NCONV()
    # KA1: 10 259 parameters (10 129 used + 130 in the unused bnorm modules)
    assert sum(p.numel() for p in m.parameters()) == 10259
    assert state_fingerprint(m.state_dict()) == str(g["init_fingerprint_seed0"])
    keys = set(m.state_dict())
    for layer in DNET._ORDER:
        for suffix in ("weight", "bias", "bnorm.weight", "bnorm.bias", "bnorm.running_mean",
                       "bnorm.running_var", "bnorm.num_batches_tracked"):
            assert "d_net.%s.%s" % (layer, suffix) in keys
    assert isinstance(m.d_net.nconv7, NConv2d) and tuple(m.d_net.nconv7.padding) == (2, 2)
    assert tuple(m.d_net.nconv6.padding) == (0, 0)
    hooks = list(m.d_net.nconv1._forward_pre_hooks.values())
    assert len(hooks) == 1 and isinstance(hooks[0], EnforcePos)
    import copy
    m2 = copy.deepcopy(m)                       # train_step1.py:98 deep-copies the best model
    assert state_fingerprint(m2.state_dict()) == state_fingerprint(m.state_dict())


@pytest.mark.skipif(not ref_shim.reference_available(), reason="reference not mounted")
def test_state_dict_keys_match_reference():
    r1, r2 = ref_shim.load_reference()
    from models.step1 import SETP1_NCONV
    from models.step2 import SETP2_BP_EXPORT
    torch.manual_seed(1)
    ours, ref = SETP1_NCONV(), None
    torch.manual_seed(1)
    ref = r1.SETP1_NCONV()
    assert list(ours.state_dict()) == list(ref.state_dict())
    assert state_fingerprint(ours.state_dict()) == state_fingerprint(ref.state_dict())
    torch.manual_seed(2)
    ours2 = SETP2_BP_EXPORT()
    torch.manual_seed(2)
    ref2 = r2.SETP2_BP_EXPORT()
    assert list(ours2.state_dict()) == list(ref2.state_dict())
    assert state_fingerprint(ours2.state_dict()) == state_fingerprint(ref2.state_dict())
    assert sum(p.numel() for p in ours2.parameters()) == 988595          # KA1
    # reference checkpoints load into ours and vice versa
    ours2.load_state_dict(ref2.state_dict())
    ref2.load_state_dict(ours2.state_dict())


@pytest.mark.skipif(not ref_shim.reference_available(), reason="reference not mounted")
def test_setp2_train_construction_order(tmp_path, monkeypatch):
    """SETP2_BP_TRAIN loads ./checkpoints/<name>.pth.tar, strips `module.`, freezes step 1 and
    builds 9 encoders (step2.py:29-51); the RNG stream must match the reference's."""
    r1, r2 = ref_shim.load_reference()
    from models.step2 import SETP2_BP_TRAIN
    torch.manual_seed(3)
    ck = r1.SETP1_NCONV()
    os.makedirs(tmp_path / "checkpoints")
    sd = {"module." + k: v for k, v in ck.state_dict().items()}          # DataParallel prefix
    torch.save({"epoch": 0, "state_dict": sd, "stats": {}}, tmp_path / "checkpoints" / "s1.pth.tar")
    monkeypatch.chdir(tmp_path)
    torch.manual_seed(4)
    ours = SETP2_BP_TRAIN("s1")
    torch.manual_seed(4)
    ref = r2.SETP2_BP_TRAIN("s1")
    assert list(ours.state_dict()) == list(ref.state_dict())
    assert state_fingerprint(ours.state_dict()) == state_fingerprint(ref.state_dict())
    assert all(not p.requires_grad for p in ours.step1.parameters())
    assert "rgb_encoder4.encoder.0.weight" in ours.state_dict()


def test_no_cpu_fallback():
    """The product path must fail loudly on CPU tensors: there is no CPU implementation."""
    from models.step1 import SETP1_NCONV, NConv2d
    m = SETP1_NCONV().eval()
    with pytest.raises((RuntimeError, NotImplementedError)):
        m(torch.zeros(1, 1, 16, 16))
    layer = NConv2d(2, 2, (3, 3), 'softplus', 'p', padding=(1, 1)).eval()
    with pytest.raises((RuntimeError, NotImplementedError)):
        layer(torch.zeros(1, 2, 8, 8), torch.zeros(1, 2, 8, 8))


def test_product_path_never_imports_oracle():
    pkg = os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "/root/reference" not in re.sub(r'""".*?"""', "", src, flags=re.S).replace(
                    "# /root/reference", ""), f


def test_fake_kernels_give_shapes():
    """register_fake: shape inference works without a device (meta tensors)."""
    import nconv_b200.ops  # noqa: F401
    x = torch.empty(2, 8, 20, 36, device="meta")
    w = torch.empty(8, 8, 5, 5, device="meta")
    b = torch.empty(8, device="meta")
    y, c = torch.ops.nconv_b200.nconv2d(x, x, w, b, 2, 2, 1e-7)
    assert y.shape == (2, 8, 20, 36) and c.shape == y.shape
    assert torch.ops.nconv_b200.maxpool2x2(x).shape == (2, 8, 10, 18)
    lo = torch.empty(2, 8, 10, 18, device="meta")
    assert torch.ops.nconv_b200.upcat(x, lo, True).shape == (2, 16, 20, 36)
