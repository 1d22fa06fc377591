"""ctypes binding of `csrc/libnconv_b200.so` (the C ABI declared in include/nconv_b200.h).

There is no fallback: if the library is missing this module raises, and every op in
`nconv_b200.ops` is CUDA-only.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_longlong, c_size_t, c_ulonglong, c_void_p

PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(PKG_ROOT, "csrc", "libnconv_b200.so")

F_THRESH, F_POOL, F_TAIL = 1, 2, 4


class FwdDesc(Structure):
    """nconv_b200_fwd_desc"""
    _fields_ = [
        ("xa", c_void_p), ("ca", c_void_p), ("xb", c_void_p), ("cb", c_void_p),
        ("w", c_void_p), ("bias", c_void_p),
        ("y", c_void_p), ("cout", c_void_p), ("y_pool", c_void_p), ("c_pool", c_void_p),
        ("w_tail", c_void_p), ("b_tail", c_void_p), ("out_tail", c_void_p),
        ("B", c_int), ("CA", c_int), ("CB", c_int), ("CO", c_int), ("H", c_int), ("W", c_int),
        ("HB", c_int), ("WB", c_int), ("K", c_int), ("pad", c_int),
        ("a_first", c_int), ("flags", c_int), ("eps", c_float), ("thr", c_float),
    ]


class ConvDesc(Structure):
    """nconv_b200_conv_desc"""
    _fields_ = [
        ("a", c_void_p), ("b", c_void_p), ("w", c_void_p), ("scale", c_void_p),
        ("shift", c_void_p), ("residual", c_void_p), ("y", c_void_p),
        ("B", c_int), ("CA", c_int), ("CB", c_int), ("CO", c_int), ("H", c_int), ("W", c_int),
        ("K", c_int), ("stride", c_int), ("relu", c_int),
    ]


class TcDesc(Structure):
    """nconv_b200_tc_desc"""
    _fields_ = [
        ("a", c_void_p), ("b", c_void_p), ("wpack", c_void_p), ("scale", c_void_p), ("shift", c_void_p),
        ("y", c_void_p), ("yq", c_void_p),
        ("B", c_int), ("CA", c_int), ("CB", c_int), ("CO", c_int), ("H", c_int), ("W", c_int), ("relu", c_int),
        ("mode", c_int), ("has_1x1", c_int),
        ("y_f32", c_void_p), ("y_ctot", c_int), ("y_c0", c_int), ("y_phase_c", c_int),
    ]


class NconvTcDesc(Structure):
    """nconv_b200_nconv_tc_desc"""
    _fields_ = [
        ("a", c_void_p), ("b", c_void_p), ("wpack", c_void_p), ("w", c_void_p), ("bias", c_void_p),
        ("out_n8", c_void_p), ("out_pool", c_void_p), ("out_up", c_void_p), ("y", c_void_p), ("cout", c_void_p),
        ("w_tail", c_void_p), ("b_tail", c_void_p), ("out_tail", c_void_p),
        ("B", c_int), ("H", c_int), ("W", c_int), ("K", c_int), ("CB", c_int), ("eps", c_float),
    ]


# name -> (restype, argtypes); must list every symbol include/nconv_b200.h declares
SIGNATURES = {
    "nconv_b200_abi_version": (c_int, []),
    "nconv_b200_error_string": (c_char_p, [c_int]),
    "nconv_b200_launch_count": (c_ulonglong, []),
    "nconv_b200_profile_enable": (c_int, [c_int]),
    "nconv_b200_profile_count": (c_int, []),
    "nconv_b200_profile_get": (c_int, [c_int, ctypes.c_char_p, c_int, POINTER(c_float), POINTER(ctypes.c_double),
                                       POINTER(ctypes.c_double)]),
    "nconv_b200_train_ws_bytes": (c_size_t, [c_int]),
    "nconv_b200_bn_train_fwd": (c_int, [c_void_p] * 12 + [c_int] * 4 + [c_float, c_float, c_int, c_void_p]),
    "nconv_b200_bn_train_bwd": (c_int, [c_void_p] * 11 + [c_int] * 5 + [c_void_p]),
    "nconv_b200_relu_bias_bwd": (c_int, [c_void_p] * 5 + [c_int] * 5 + [c_void_p]),
    "nconv_b200_relu_bias_bwd_noalloc": (c_int, [c_void_p] * 4 + [c_int] * 5 + [c_void_p]),
    "nconv_b200_conv_wgrad_ws_bytes": (c_size_t, [c_int] * 3),
    "nconv_b200_conv_wgrad": (c_int, [c_void_p] * 4 + [c_size_t] + [c_int] * 8 + [c_void_p]),
    "nconv_b200_conv_dgrad_ws_bytes": (c_size_t, [c_int] * 3),
    "nconv_b200_conv_dgrad_gather": (c_int, [c_void_p] * 4 + [c_int] * 9 + [c_void_p]),
    "nconv_b200_wflip": (c_int, [c_void_p] * 2 + [c_int] * 3 + [c_void_p]),
    "nconv_b200_nconv2d_fwd": (c_int, [c_void_p] * 6 + [c_int] * 9 + [c_float, c_void_p]),
    "nconv_b200_nconv2d_fwd_fused": (c_int, [POINTER(FwdDesc), c_void_p]),
    "nconv_b200_nconv2d_bwd_workspace_bytes": (c_size_t, [c_int] * 9),
    "nconv_b200_nconv2d_bwd": (c_int, [c_void_p] * 12 + [c_int] * 9 + [c_float, c_void_p, c_size_t, c_void_p]),
    "nconv_b200_enforce_pos": (c_int, [c_void_p, c_longlong, c_int, c_int, c_void_p]),
    "nconv_b200_confidence_mask": (c_int, [c_void_p, c_void_p, c_longlong, c_float, c_void_p]),
    "nconv_b200_maxpool2x2_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "nconv_b200_maxpool2x2_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "nconv_b200_pool_conf_argmax_fwd": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_void_p]),
    "nconv_b200_pool_conf_argmax_bwd": (c_int, [c_void_p] * 5 + [c_int] * 3 + [c_void_p]),
    "nconv_b200_upcat_fwd": (c_int, [c_void_p] * 3 + [c_int] * 8 + [c_void_p]),
    "nconv_b200_upcat_bwd": (c_int, [c_void_p] * 3 + [c_int] * 8 + [c_void_p]),
    "nconv_b200_dnet_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "nconv_b200_dnet_forward": (c_int, [c_void_p, POINTER(c_void_p), c_void_p, c_int, c_int, c_int,
                                        c_void_p, c_size_t, c_void_p]),
    "nconv_b200_debug_buffer": (c_int, [c_void_p]),
    "nconv_b200_conv_wgrad_tc_ws_bytes": (c_size_t, [c_int] * 5),
    "nconv_b200_conv_wgrad_tc": (c_int, [c_void_p] * 4 + [c_size_t] + [c_int] * 5 + [c_void_p]),
    "nconv_b200_rgb_u8_to_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "nconv_b200_depth_decode": (c_int, [c_void_p, c_int, c_void_p, c_longlong, c_float, c_void_p]),
    "nconv_b200_sparsify": (c_int, [c_void_p] * 4 + [c_longlong, c_void_p]),
    "nconv_b200_n8_bytes": (c_size_t, [c_int] * 3),
    "nconv_b200_n8_zero_ring": (c_int, [c_void_p] + [c_int] * 3 + [c_void_p]),
    "nconv_b200_n8_pack": (c_int, [c_void_p] * 3 + [c_int] * 3 + [c_void_p]),
    "nconv_b200_n8_unpack": (c_int, [c_void_p] * 3 + [c_int] * 3 + [c_void_p]),
    "nconv_b200_nconv1_n8": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_float, c_float, c_void_p]),
    "nconv_b200_nconv_tc_wpack_bytes": (c_size_t, [c_int] * 2),
    "nconv_b200_nconv_tc_pack_weights": (c_int, [c_void_p] * 2 + [c_int] * 3 + [c_void_p]),
    "nconv_b200_nconv_tc": (c_int, [POINTER(NconvTcDesc), c_void_p]),
    "nconv_b200_dnet_forward_fp32": (c_int, [c_void_p, POINTER(c_void_p), c_void_p, c_int, c_int, c_int,
                                             c_void_p, c_size_t, c_void_p]),
    "nconv_b200_conv2d_fwd": (c_int, [POINTER(ConvDesc), c_void_p]),
    "nconv_b200_convT4x4s2_fwd": (c_int, [c_void_p] * 6 + [c_int] * 7 + [c_void_p]),
    "nconv_b200_bilinear_ac_fwd": (c_int, [c_void_p, c_void_p] + [c_int] * 5 + [c_void_p]),
    "nconv_b200_zero_border": (c_int, [c_void_p] + [c_int] * 6 + [c_void_p]),
    "nconv_b200_bn_fold": (c_int, [c_void_p] * 7 + [c_int, c_float, c_void_p]),
    "nconv_b200_p8_bytes": (c_size_t, [c_int] * 4),
    "nconv_b200_p8_pack": (c_int, [c_void_p, c_void_p] + [c_int] * 4 + [c_void_p]),
    "nconv_b200_p8_unpack": (c_int, [c_void_p, c_void_p] + [c_int] * 4 + [c_void_p]),
    "nconv_b200_p8_zero_ring": (c_int, [c_void_p] + [c_int] * 4 + [c_void_p]),
    "nconv_b200_relu_bias_bwd_p8": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p]),
    "nconv_b200_p8_unpack_into": (c_int, [c_void_p] * 2 + [c_int] * 6 + [c_void_p]),
    "nconv_b200_p8_pack_ex": (c_int, [c_void_p] * 2 + [c_int] * 7 + [c_void_p]),
    "nconv_b200_convT_phase_weights": (c_int, [c_void_p] * 2 + [c_int] * 2 + [c_void_p]),
    "nconv_b200_convS2_phase_weights": (c_int, [c_void_p] * 2 + [c_int] * 2 + [c_void_p]),
    "nconv_b200_convS2_phase_wgrad": (c_int, [c_void_p] * 2 + [c_int] * 2 + [c_void_p]),
    "nconv_b200_convT_phase_wgrad": (c_int, [c_void_p] * 2 + [c_int] * 3 + [c_void_p]),
    "nconv_b200_p8q_bytes": (c_size_t, [c_int] * 4),
    "nconv_b200_p8q_zero_ring": (c_int, [c_void_p] + [c_int] * 4 + [c_void_p]),
    "nconv_b200_p8q_pack": (c_int, [c_void_p, c_void_p] + [c_int] * 4 + [c_void_p]),
    "nconv_b200_tc_wpack_bytes": (c_size_t, [c_int] * 5),
    "nconv_b200_tc_pack_weights": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "nconv_b200_conv_tc": (c_int, [POINTER(TcDesc), c_void_p]),
    "nconv_b200_conv_small_p8": (c_int, [c_void_p] * 7 + [c_int] * 6 + [c_void_p]),
    "nconv_b200_head_p8": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p]),
    "nconv_b200_masked_mse_ws_bytes": (c_size_t, [c_int] * 5),
    "nconv_b200_masked_mse_bilinear_fwd": (c_int, [c_void_p] * 4 + [c_size_t] + [c_int] * 5 + [c_float, c_int, c_void_p]),
    "nconv_b200_masked_mse_bilinear_bwd": (c_int, [c_void_p] * 5 + [c_size_t] + [c_int] * 5 + [c_float, c_void_p]),
    "nconv_b200_sobel_loss_ws_bytes": (c_size_t, [c_int] * 5),
    "nconv_b200_sobel_loss_fwd": (c_int, [c_void_p] * 5 + [c_size_t] + [c_int] * 5 + [c_float, c_int, c_void_p]),
    "nconv_b200_sobel_loss_bwd": (c_int, [c_void_p] * 6 + [c_size_t] + [c_int] * 5 + [c_float, c_void_p]),
    "nconv_b200_adamw_step": (c_int, [c_void_p] * 4 + [c_longlong] + [c_float] * 5 + [c_int, c_void_p]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load the C-ABI library; raise (loudly) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "nconv_b200: %s not found. Build it with `python __graft_entry__.py build` "
            "(nvcc, sm_100a). There is no CPU or PyTorch fallback for this path." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def error_string(code: int) -> str:
    return load().nconv_b200_error_string(code).decode()


def check(code: int, what: str) -> None:
    if code != 0:
        raise RuntimeError("%s failed: [%d] %s" % (what, code, error_string(code)))


def launch_count() -> int:
    return int(load().nconv_b200_launch_count())


def profile_enable(on: bool) -> None:
    """Per-launch CUDA-event timing inside the library (bench.py's roofline leg)."""
    check(load().nconv_b200_profile_enable(int(on)), "profile_enable")


def profile_records():
    """[(label, ms, flops, bytes)] for every launch since profile_enable(True).  The caller
    must have synchronised the device."""
    lib = load()
    out = []
    buf = ctypes.create_string_buffer(64)
    ms, fl, by = c_float(), ctypes.c_double(), ctypes.c_double()
    for i in range(lib.nconv_b200_profile_count()):
        check(lib.nconv_b200_profile_get(i, buf, 64, ctypes.byref(ms), ctypes.byref(fl), ctypes.byref(by)),
              "profile_get")
        out.append((buf.value.decode(), float(ms.value), float(fl.value), float(by.value)))
    return out
