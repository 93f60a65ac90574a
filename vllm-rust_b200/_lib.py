"""ctypes binding of libw4a16_b200.so (include/w4a16_b200.h).

The product has no CPU fallback: if the shared library is missing, cannot be
loaded, or a compute entry point is called without an sm_100 device, this
module raises -- it never routes to the oracle or to eager PyTorch.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libw4a16_b200.so")
CSRC = os.path.join(_HERE, "csrc")
HEADER = os.path.join(os.path.dirname(_HERE), "include", "w4a16_b200.h")


class W4A16Error(RuntimeError):
    """Mirror of candle_core::Error::Msg raised by the reference's launchers
    (marlin_cuda.rs:546-564): every failure carries the library's message."""


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile libw4a16_b200.so in-tree with nvcc for sm_100a (no GPU needed)."""
    args = ["make", "-C", CSRC, "-j8"]
    if force:
        subprocess.check_call(["make", "-C", CSRC, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(args, stdout=None if verbose else subprocess.DEVNULL)
    if not os.path.exists(LIB_PATH):
        raise W4A16Error(f"build finished but {LIB_PATH} is missing")
    return LIB_PATH


_vp, _i64, _i32, _u64 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_uint64

# name -> (restype, argtypes); MUST list every symbol include/w4a16_b200.h declares
# (tests/test_abi.py parses the header and compares).
SIGNATURES = {
    "w4a16_last_error": (ctypes.c_char_p, []),
    "w4a16_abi_version": (_i32, []),
    "w4a16_check_shape": (_i32, [_i64, _i64, _i32]),
    "w4a16_create": (_i32, [ctypes.POINTER(_vp), _i64, _i64, _i32, _i32, _i32, _i32]),
    "w4a16_create_shard": (_i32, [ctypes.POINTER(_vp), _i64, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32]),
    "w4a16_destroy": (_i32, [_vp]),
    "w4a16_load_awq": (_i32, [_vp, _vp, _vp, _vp, _i32, _vp, _i32, _vp]),
    "w4a16_load_gptq": (_i32, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _vp]),
    "w4a16_load_awq_part": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp, _i32, _vp, _i32, _vp]),
    "w4a16_load_gptq_part": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _vp]),
    "w4a16_set_epilogue": (_i32, [_vp, _i32]),
    "w4a16_save_tiled": (_i32, [_vp, ctypes.c_char_p, _u64]),
    "w4a16_load_tiled": (_i32, [_vp, ctypes.c_char_p, _u64, _vp]),
    "w4a16_forward_into": (_i32, [_vp, _vp, _vp, _i64, _vp]),
    "w4a16_forward_decode_static": (_i32, [_vp, _vp, _vp, _i64, _vp]),
    "w4a16_forward_host": (_i32, [_vp, _vp, _vp, _i64, _vp]),
    "w4a16_forward_row_allreduce": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "w4a16_peer_create": (_i32, [ctypes.POINTER(_vp), _i32, _i32, _i64]),
    "w4a16_peer_export": (_i32, [_vp, _vp]),
    "w4a16_peer_attach": (_i32, [_vp, _vp]),
    "w4a16_peer_destroy": (_i32, [_vp]),
    "w4a16_forward_row_fused": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp, _vp]),
    "w4a16_nccl_unique_id": (_i32, [_vp]),
    "w4a16_nccl_comm_init": (_i32, [ctypes.POINTER(_vp), _i32, _vp, _i32]),
    "w4a16_nccl_comm_destroy": (_i32, [_vp]),
    "w4a16_reserve": (_i32, [_vp, _i64]),
    "w4a16_dequant_into": (_i32, [_vp, _vp, _vp]),
    "w4a16_dequant_d1_into": (_i32, [_vp, _vp, _i32, _vp, _vp]),
    "w4a16_row_perm": (_i32, [_vp, _vp]),
    "w4a16_unpack_into": (_i32, [_vp, _vp, _vp, _vp]),
    "w4a16_scales_into": (_i32, [_vp, _vp, _vp]),
    "w4a16_in_features": (_i64, [_vp]),
    "w4a16_out_features": (_i64, [_vp]),
    "w4a16_weight_rows": (_i64, [_vp]),
    "w4a16_has_bias": (_i32, [_vp]),
    "w4a16_weight_bytes": (_i64, [_vp]),
    "w4a16_kernel_launches": (_u64, []),
    "w4a16_kernel_for": (ctypes.c_char_p, [_vp, _i64]),
    "w4a16_set_dispatch": (_i32, [ctypes.c_char_p]),
    "w4a16_debug_stamps": (_i32, [_vp, _i64, ctypes.POINTER(_u64)]),
    "w4a16_chain_create": (_i32, [ctypes.POINTER(_vp), _i32, ctypes.POINTER(_vp), ctypes.POINTER(_vp), ctypes.POINTER(_vp), _i64]),
    "w4a16_chain_set_peer": (_i32, [_vp, _vp]),
    "w4a16_chain_run": (_i32, [_vp, _vp]),
    "w4a16_chain_run_host": (_i32, [_vp, _vp, _vp, _vp]),
    "w4a16_chain_destroy": (_i32, [_vp]),
    "w4a16_chain_kernel": (ctypes.c_char_p, [_vp]),
    "w4a16_chain_debug_stamps": (_i32, [_vp, _vp, _i64]),
}

class ChainPrologue(ctypes.Structure):
    """w4a16_chain_prologue (include/w4a16_b200.h)."""
    _fields_ = [("norm_weight", _vp), ("residual_in", _vp), ("residual_out", _vp), ("normed_out", _vp),
                ("eps", ctypes.c_float), ("add_running_residual", _i32)]


SIGNATURES["w4a16_chain_create_ex"] = (_i32, [ctypes.POINTER(_vp), _i32, ctypes.POINTER(_vp), ctypes.POINTER(_vp),
                                               ctypes.POINTER(_vp), _i64, ctypes.POINTER(ChainPrologue)])

_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise W4A16Error(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). There is no CPU fallback.")
        try:
            handle = ctypes.CDLL(LIB_PATH)
        except OSError as e:  # loud, never silent
            raise W4A16Error(f"cannot load {LIB_PATH}: {e}") from e
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        if handle.w4a16_abi_version() != 1:
            raise W4A16Error("libw4a16_b200.so ABI version mismatch")
        _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise W4A16Error(lib().w4a16_last_error().decode("utf-8", "replace"))
