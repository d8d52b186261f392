"""ctypes binding of ``libzkb200.so`` (the C ABI declared in ``include/zkb200.h``).

The library is the product: there is no CPU or PyTorch fallback.  If the shared object is missing
or a call returns a non-zero status, an exception is raised.
"""
from __future__ import annotations

import ctypes
from ctypes import c_double, c_float, c_int, c_longlong, c_ulonglong, c_void_p
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "libzkb200.so"


class ZkbError(RuntimeError):
    """A libzkb200 entry point returned a non-zero status."""

    def __init__(self, fn: str, status: int, message: str):
        super().__init__(f"{fn} failed with status {status}: {message}")
        self.fn = fn
        self.status = status
        self.message = message


P = c_void_p
I = c_int
L = c_longlong
U = c_ulonglong
F = c_float
D = c_double

# name -> argument ctypes, in the order of include/zkb200.h
SIGNATURES = {
    "zkb_preprocess_u8": [P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, U, U, P, P],
    "zkb_sign_pack": [P, P, L, I, P],
    "zkb_bgemm_popc": [P, P, P, L, I, I, I, P],
    "zkb_conv_first_fwd": [P, P, P, I, I, I, I, I, I, P],
    "zkb_conv_first_wgrad": [P, P, P, I, I, I, I, I, I, P],
    "zkb_conv_first_bn_stats": [P, P, I, I, I, I, I, I, P, P, P],
    "zkb_conv_first_bn_sign": [P, P, I, I, I, I, I, I, P, P, P, P, P, P],
    "zkb_conv_first_bwd_reduce": [P, P, P, I, I, I, I, I, I, P, P, P, P, P, P],
    "zkb_conv_first_bwd_wgrad": [P, P, P, I, I, I, I, I, I, P, P, P, P, P, P, P],
    "zkb_bconv2d_fwd": [P, P, P, I, I, I, I, I, I, P],
    "zkb_bdense_fwd": [P, P, P, L, I, I, I, P],
    "zkb_bconv2d_fwd_i8": [P, P, P, I, I, I, I, I, I, P],
    "zkb_bdense_fwd_i8": [P, P, P, L, I, I, I, P],
    "zkb_bconv2d_fwd_i8_sign": [P, P, P, P, I, I, I, I, I, I, P],
    "zkb_bdense_fwd_i8_sign": [P, P, P, P, L, I, I, I, P],
    "zkb_bn_eval_rstd": [P, F, I, P, P],
    "zkb_bn_sign_thresholds": [P, P, P, I, P, P],
    "zkb_maxpool2_sign_i8": [P, I, I, I, I, P, P],
    "zkb_bconv2d_fwd_i8_stats": [P, P, P, I, I, I, I, I, I, P, P, P],
    "zkb_bdense_fwd_i8_stats": [P, P, P, L, I, I, I, P, P, P],
    "zkb_bconv2d_dgrad": [P, P, P, P, I, I, I, I, I, I, P],
    "zkb_bdense_dgrad": [P, P, P, P, I, I, I, P],
    "zkb_bconv2d_dgrad_bn": [P, P, P, P, I, I, I, I, I, I, P, P, P, P, P, P, P],
    "zkb_bconv2d_dgrad_first_bwd": [P, P, P, I, I, I, I, I, I, P, I, I, P, P, P, P, P, P],
    "zkb_conv_first_bwd_finalize": [P, P, P, P, P, D, D, P, P, I, P],
    "zkb_conv_first_gram": [P, I, I, I, P, P],
    "zkb_conv_first_stats_from_gram": [P, P, I, P, P, P],
    "zkb_bconv2d_wgrad": [P, P, P, P, I, I, I, I, I, I, P],
    "zkb_bdense_wgrad": [P, P, P, P, L, I, I, I, I, P],
    "zkb_bn_stats": [P, I, L, I, P, P, P],
    "zkb_bn_finalize": [P, P, D, F, F, I, I, P, P, P, P, P],
    "zkb_bn_apply_sign": [P, I, L, I, P, P, P, P, P, P, P],
    "zkb_maxpool2_i16": [P, I, I, I, I, P, P, P, P, P],
    "zkb_softmax_ce": [P, I, P, I, I, F, P, P, P, P],
    "zkb_bn_bwd_reduce": [P, P, I, L, I, P, P, P, I, P, P, P],
    "zkb_bn_bwd_finalize": [P, P, D, D, I, P, P, P, P],
    "zkb_bn_bwd_apply": [P, P, I, L, I, P, P, P, P, P, I, I, P, P, P, P, I, I, P],
    "zkb_prep_conv_weights": [P, I, I, I, P, P, P, P],
    "zkb_prep_dense_weights": [P, I, I, I, P, P, P, P],
    "zkb_sign_f32": [P, P, L, P],
    "zkb_adam_clip": [P, P, P, P, L, F, F, F, F, F, I, P, P],
    "zkb_weight_abs_mean": [P, L, I, P, P],
    "zkb_bn_gain_finalize": [P, P, D, P, P, F, F, I, I, P, P, P, P, P, P],
    "zkb_bn_gain_add": [P, L, I, P, P, P, P, P, P],
    "zkb_bn_gain_bwd_apply": [P, P, L, I, P, P, P, P, P, P, P, P],
    "zkb_approx_sign_bwd_add": [P, P, P, P, L, P],
    "zkb_global_avgpool_fwd": [P, I, I, I, P, P],
    "zkb_global_avgpool_bwd": [P, I, I, I, P, P],
    "zkb_dense_f32_fwd": [P, P, P, P, I, I, I, I, P],
    "zkb_dense_f32_bwd": [P, P, P, I, P, P, P, I, I, I, P],
    "zkb_maxpool3s2_f32_fwd": [P, I, I, I, I, P, P, P],
    "zkb_maxpool3s2_f32_bwd": [P, P, I, I, I, I, P, P],
    "zkb_maxpool3s2_fwd": [P, I, I, I, I, I, I, P, P, P],
    "zkb_maxpool3s2_bwd": [P, P, I, I, I, I, I, P, P, P, P],
    "zkb_bconv2d_fwd_i8_pool": [P, P, P, P, I, I, I, I, I, I, P, P, P],
    "zkb_bconv2d_fwd_i8_s2": [P, P, P, I, I, I, I, I, I, P],
    "zkb_bconv2d_fwd_s2": [P, P, P, I, I, I, I, I, I, P],
    "zkb_conv_stem_f32_fwd": [P, P, P, I, I, I, I, I, I, I, P],
    "zkb_conv_stem_f32_wgrad": [P, P, P, I, I, I, I, I, I, I, P],
    "zkb_bconv2d_wgrad_s2": [P, P, P, P, I, I, I, I, I, I, P],
    "zkb_dilate2_hi_lo": [P, P, I, I, I, I, I, P, P, P],
    "zkb_avgpool2_f32_fwd": [P, I, I, I, I, P, P],
    "zkb_avgpool2_f32_bwd": [P, I, I, I, I, P, P],
    "zkb_bn_gain_add_f32": [P, L, I, P, P, P, P, P, P],
    "zkb_bn_gain_bwd_apply_f32": [P, P, L, I, P, P, P, P, P, P, P],
    "zkb_bn_train_sign_static": [P, I, L, I, P, F, F, I, P, P, P, P, P],
}

_lib = None


def load() -> ctypes.CDLL:
    """dlopen the library once; fail loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ZkbError(
            "dlopen", -100,
            f"{LIB_PATH} does not exist; run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)",
        )
    lib = ctypes.CDLL(str(LIB_PATH))
    lib.zkb_last_error_string.restype = ctypes.c_char_p
    lib.zkb_last_error_string.argtypes = []
    lib.zkb_abi_version.restype = c_int
    if hasattr(lib, "zkb_bconv2d_fwd_i8_pool_ok"):
        lib.zkb_bconv2d_fwd_i8_pool_ok.restype = c_int
        lib.zkb_bconv2d_fwd_i8_pool_ok.argtypes = [c_int, c_int, c_int]
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue  # reported by tests/test_abi.py; calling it raises below
        fn.restype = c_int
        fn.argtypes = argtypes
    _lib = lib
    return lib


def call(name: str, *args) -> None:
    lib = load()
    fn = getattr(lib, name, None)
    if fn is None:
        raise ZkbError(name, -101, "symbol is not exported by libzkb200.so")
    status = fn(*args)
    if status != 0:
        raise ZkbError(name, status, lib.zkb_last_error_string().decode(errors="replace"))


def ptr(t) -> int:
    """Device (or host) address of a torch tensor, 0 for None."""
    return 0 if t is None else t.data_ptr()
