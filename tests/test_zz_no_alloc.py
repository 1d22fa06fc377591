"""The no-allocation entry point `nconv_b200_relu_bias_bwd_noalloc` (include/nconv_b200.h; no counterpart in the reference)
against the existing `nconv_b200_relu_bias_bwd`, through the C ABI.

Collected last on purpose (a failure here hides nothing after it).  The static scratch array is
shared by every call and keeps the previous call's sums, so the check is made on two calls in a row with different inputs
and channel counts; it holds in a fresh process and after the other tests (nothing else uses that array)."""
import ctypes

import pytest
import torch

from conftest import report

pytestmark = pytest.mark.gpu

MAX_C = 1024                      # NCONV_B200_NOALLOC_MAX_C
ERR_STATIC_SCRATCH = 100004       # NCONV_B200_ERR_STATIC_SCRATCH


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _both(lib, gy, y, relu):
    """(g, gbias) from the workspace entry point and from the no-allocation one."""
    B, C, H, W = gy.shape
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ws = torch.empty(int(lib.nconv_b200_train_ws_bytes(C)), dtype=torch.uint8, device=gy.device)
    g0, b0 = torch.empty_like(gy), torch.empty(C, device=gy.device)
    g1, b1 = torch.empty_like(gy), torch.empty(C, device=gy.device)
    assert lib.nconv_b200_relu_bias_bwd(_p(gy), _p(y), _p(g0), _p(b0), _p(ws), B, C, H, W, int(relu), st) == 0
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    assert lib.nconv_b200_relu_bias_bwd_noalloc(_p(gy), _p(y), _p(g1), _p(b1), B, C, H, W, int(relu), st) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_reserved() == reserved0 and free1 >= free0      # nothing was allocated by the call
    return (g0, b0), (g1, b1)


def test_relu_bias_bwd_noalloc_matches_workspace_entry_point_twice():
    from nconv_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.manual_seed(1234)
    worst = 0.0
    # two calls in a row: the second has FEWER channels and other data, so sums left behind by the first would show
    for k, (shape, relu) in enumerate([((3, 64, 37, 52), True), ((2, 24, 60, 80), True), ((1, 40, 16, 24), False)]):
        gy = torch.randn(*shape, device=dev)
        y = torch.relu(torch.randn(*shape, device=dev))
        (g0, b0), (g1, b1) = _both(lib, gy, y if relu else None, relu)
        ref = (gy * (y > 0) if relu else gy)
        if relu:
            assert torch.equal(g0, g1) and torch.equal(g1, ref)              # the masked gradient: bit-exact
        scale = float(ref.abs().sum((0, 2, 3)).max())
        # per-channel sums: fp64 atomics in either entry point (order varies run to run) -> 1e-6 of the channel's L1 mass,
        # the bound the bias-gradient tests use; against torch's fp64 sum as well
        err = float((b0.double() - b1.double()).abs().max()) / scale
        err64 = float((b1.double() - ref.double().sum((0, 2, 3))).abs().max()) / scale
        worst = max(worst, err, err64)
        assert err <= 1e-6 and err64 <= 1e-6, (k, err, err64)
    report("relu_bias_bwd_noalloc", max_err_over_channel_l1=worst, calls=3)


def test_relu_bias_bwd_noalloc_rejects_sizes_beyond_the_static_array():
    from nconv_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    C = MAX_C + 8
    gy = torch.zeros(1, C, 2, 2, device=dev)
    gb = torch.full((C,), 5.0, device=dev)
    rc = lib.nconv_b200_relu_bias_bwd_noalloc(_p(gy), None, None, _p(gb), 1, C, 2, 2, 0, st)
    assert rc == ERR_STATIC_SCRATCH
    assert "static scratch" in _lib.error_string(rc)
    torch.cuda.synchronize()
    assert bool((gb == 5).all())                                             # nothing was launched
    # the largest size it accepts
    gy = torch.ones(1, MAX_C, 2, 2, device=dev)
    gb = torch.empty(MAX_C, device=dev)
    assert lib.nconv_b200_relu_bias_bwd_noalloc(_p(gy), None, None, _p(gb), 1, MAX_C, 2, 2, 0, st) == 0
    torch.cuda.synchronize()
    assert bool((gb == 4).all())
