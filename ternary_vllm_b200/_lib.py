"""ctypes binding of ``libternary_b200.so`` (the C ABI declared in ``include/ternary_b200.h``).

There is deliberately no fallback: if the shared library is missing the import fails with
instructions to build it, and on a machine without an sm_100 GPU every compute call raises
``TernaryB200Error`` (TB200_ENODEV).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# TB200_LIB selects another build of the same ABI (the experiment library of `make ABLATE=1`); never a fallback
LIB_PATH = os.environ.get("TB200_LIB") or os.path.join(_HERE, "libternary_b200.so")

TB200_F32, TB200_BF16, TB200_F16 = 0, 1, 2
Q_TERNARY, Q_SIGN, Q_STEP = 0, 1, 2
FMT_A, FMT_B, FMT_C = 0, 1, 2
GEMV_MAX_M = 32
GEMV_FUSED_M = 8          # TB200_GEMV_FUSED_M: tokens per pass of the decode GEMVs
STATIC_SCRATCH_BYTES = 8 << 20    # TB200_STATIC_SCRATCH_BYTES (tb200_linear_forward_static)
FLAG_PDL = 1
ACT_NONE, ACT_GELU_TANH, ACT_RELU = 0, 1, 2
CHAIN_MAX_M, CHAIN_MAX_LAYERS = 8, 128


class ChainLayerDesc(C.Structure):
    """``tb200_chain_layer`` of include/ternary_b200.h"""
    _fields_ = [("packed", C.c_void_p), ("ld_packed", C.c_int64), ("N", C.c_int64), ("K", C.c_int64),
                ("scale", C.c_void_p), ("bias", C.c_void_p), ("clip_val", C.c_float), ("act", C.c_int32)]

ERRORS = {-1: "TB200_EINVAL", -2: "TB200_ECUDA", -3: "TB200_ENODEV", -4: "TB200_ENOTSUP"}


class TernaryB200Error(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile the library in-tree with nvcc (sm_100a).  Used by ``__graft_entry__.build``."""
    import subprocess

    csrc = os.path.join(_HERE, "csrc")
    out = subprocess.run(["make", "-C", csrc, "-j8"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("building libternary_b200.so failed:\n" + out.stdout[-4000:] + out.stderr[-4000:])
    if verbose:
        print(out.stdout[-2000:])
    return LIB_PATH


_i64, _i32, _f32, _vp, _sz = C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); mirrors include/ternary_b200.h one to one
SIGNATURES = {
    "tb200_abi_version": (_i32, []),
    "tb200_last_error": (C.c_char_p, []),
    "tb200_launch_count": (C.c_uint64, []),
    "tb200_debug_set_trace": (None, [_vp]),
    "tb200_quantize_pack": (_i32, [_vp, _i32, _i64, _i64, _i64, _i32, _vp, _i64, _vp, _vp]),
    "tb200_unpack": (_i32, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "tb200_pack_format": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp]),
    "tb200_unpack_format": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp]),
    "tb200_clip_latent": (_i32, [_vp, _i32, _i64, _vp]),
    "tb200_b1_pitch": (_i64, [_i64]),
    "tb200_quantize_pack_b1": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _vp]),
    "tb200_expand_b1": (_i32, [_vp, _i64, _i64, _i64, _i32, _vp, _i64, _vp]),
    "tb200_gemv_b1_fused": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, C.c_float, _i32, _vp, _i32, _i64, _i32, _vp]),
    "tb200_gemv_b1": (_i32, [_vp, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _vp, C.c_float, _i32, _vp, _i32, _i64, _i32, _vp]),
    "tb200_train_workspace": (C.c_size_t, []),
    "tb200_sumsq_accumulate": (_i32, [_vp, _i32, _i64, _i32, _vp, _vp, _vp]),
    "tb200_biobjective_apply": (_i32, [_vp, _vp, _i32, _i64, _vp, _vp, _vp]),
    "tb200_ternary_init_workspace": (_sz, [_i64]),
    "tb200_ternary_init": (_i32, [_vp, _i32, _i64, _i64, _i32, _i32, _vp, _vp, _vp, _vp]),
    "tb200_binary_init": (_i32, [_vp, _i32, _i64, _i64, _i32, _i32, _f32, _i32, _vp, _vp, _vp, _vp]),
    "tb200_quant_act_i8": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _i32, _vp]),
    "tb200_gemv": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32,
                          _vp, _i32, _i64, _vp, _vp, _vp, _i32, _vp]),
    "tb200_gemv_q": (_i32, [_vp, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32,
                            _vp, _i32, _i64, _vp, _i32, _vp]),
    "tb200_gemv_f32": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32,
                              _vp, _i32, _i64, _vp]),
    "tb200_gemv_grouped": (_i32, [_vp, _i32, _i64, _i64, _i64, _i64, _vp, _i64, _i64, _i64,
                                  _vp, _vp, _i64, _f32, _vp, _i32, _i64, _i64, _i32, _i32, _vp]),
    "tb200_gemv_stream_workspace": (_sz, [_i64, _i64, _i32]),
    "tb200_gemv_stream": (_i32, [_vp, _i32, _i64, _i64, _i64, _i64, _vp, _i64, _i64, _i64,
                                 _vp, _vp, _i64, _f32, _vp, _i32, _i64, _i64, _i32, _vp, _vp, _vp]),
    "tb200_gemv_stream_mc": (_i32, [_vp, _i32, _i64, _i64, _i64, _i64, _vp, _i64, _i64, _i64,
                                    _vp, _vp, _i64, _f32, _vp, _i32, _i64, _i64, _i32, _vp, _vp]),
    "tb200_linear_forward_mc": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32, _vp, _i32, _i64, _vp, _vp]),
    "tb200_gemv_stream_tp": (_i32, [_vp, _i32, _i64, _i64, _i64, _i64, _vp, _i64, _i64, _i64,
                                    _vp, _vp, _i64, _f32, _vp, _i32, _i32, _i64, _i64, _i32, _vp, _vp]),
    "tb200_gemv_stream_host": (_i32, [_vp, _i32, _i64, _i64, _vp, _i64, _i64, _i64, _vp, _vp, _i64, _f32,
                                      _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    "tb200_chain_plan_bytes": (_sz, [_i32]),
    "tb200_chain_plan_build": (_i32, [_vp, _i32, _i64, _i32, _vp, _vp]),
    "tb200_chain_forward": (_i32, [_vp, _vp, _vp, _i32, _i64, _i64, _vp, _i32, _i64, _vp, _i32, _vp]),
    "tb200_gemm_i8": (_i32, [_vp, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32,
                             _vp, _i32, _i64, _vp, _vp]),
    "tb200_linear_workspace": (_sz, [_i64, _i64]),
    "tb200_linear_forward": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32,
                                    _vp, _i32, _i64, _vp, _vp]),
    "tb200_linear_forward_static": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32,
                                           _vp, _i32, _i64, _vp]),
    "tb200_linear_forward_tp": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32,
                                       _vp, _i32, _i32, _i64, _vp, _vp]),
    "tb200_linear_forward_host": (_i32, [_vp, _i32, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _f32,
                                         _vp, _i32, _vp, _vp, _vp, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise TernaryB200Error(
                f"{LIB_PATH} is missing: build it with `make -C {os.path.join(_HERE, 'csrc')}` "
                "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)     # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().tb200_last_error().decode("utf-8", "replace")
        raise TernaryB200Error(f"{what}: {ERRORS.get(rc, rc)}: {msg}")
