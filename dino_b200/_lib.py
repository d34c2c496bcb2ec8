"""ctypes binding of libdino_b200.so (the C ABI declared in include/dino_b200.h).

There is no fallback: if the shared library has not been built, or a call fails, this module
raises.  Build it with ``python -c "import __graft_entry__ as g; g.build()"`` or
``make -C dino_b200/csrc``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdino_b200.so")

ACT_LINEAR, ACT_SOFTPLUS = 0, 1
JAC_NONE, JAC_FULL, JAC_OUT_REDUCED, JAC_SKETCHED = 0, 1, 2, 3
PREC_FP32, PREC_TF32, PREC_FP32X3 = 0, 1, 2
STAT_L2_SUM, STAT_H1_SUM, STAT_EQ_SUM, STAT_NQ_SUM, STAT_JROW_SUM, STAT_COUNT, STAT_JROWS = range(7)
STAT_SLOTS = 8
NOALLOC_MAX_BATCH, NOALLOC_MAX_WIDTH = 256, 1024

JAC_MODES = {"none": JAC_NONE, "full": JAC_FULL, "out_reduced": JAC_OUT_REDUCED, "sketched": JAC_SKETCHED}
PRECISIONS = {"fp32": PREC_FP32, "tf32": PREC_TF32, "fp32x3": PREC_FP32X3}


class DinoLayerSpec(C.Structure):
    _fields_ = [("n_out", C.c_int32), ("act", C.c_int32), ("has_bias", C.c_int32),
                ("trainable", C.c_int32), ("skip_from", C.c_int32)]


class DinoNetSpec(C.Structure):
    _fields_ = [("d_in", C.c_int32), ("n_layers", C.c_int32), ("layers", C.POINTER(DinoLayerSpec)),
                ("jac_mode", C.c_int32), ("seed_dim", C.c_int32), ("max_batch", C.c_int32),
                ("precision", C.c_int32)]


class DinoBatch(C.Structure):
    _fields_ = [("m", C.c_void_p), ("q", C.c_void_p), ("J", C.c_void_p), ("q_sqnorm", C.c_void_p),
                ("J_sqnorm", C.c_void_p), ("idx", C.c_void_p), ("U", C.c_void_p), ("V", C.c_void_p),
                ("seed", C.c_void_p), ("ld_J", C.c_int64), ("J_row_sqnorm", C.c_void_p), ("n_samples", C.c_int64)]


# name -> (restype, argtypes): every symbol include/dino_b200.h declares
SIGNATURES = {
    "dino_plan_create": (C.c_int, [C.POINTER(DinoNetSpec), C.c_int, C.POINTER(C.c_void_p)]),
    "dino_plan_destroy": (None, [C.c_void_p]),
    "dino_plan_param_count": (C.c_int64, [C.c_void_p]),
    "dino_plan_layer_offsets": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "dino_plan_workspace_bytes": (C.c_size_t, [C.c_void_p]),
    "dino_plan_bind_workspace": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "dino_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(DinoBatch), C.c_int32, C.c_void_p, C.c_void_p]),
    "dino_jacobian": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(DinoBatch), C.c_int32, C.c_void_p, C.c_void_p,
                                C.c_void_p]),
    "dino_jvp": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(DinoBatch), C.c_int32, C.c_void_p, C.c_int32, C.c_void_p,
                           C.c_void_p]),
    "dino_h1_fwd_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(DinoBatch), C.c_int32, C.c_float, C.c_float,
                                  C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "dino_h1_eval": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(DinoBatch), C.c_int32, C.c_void_p, C.c_void_p]),
    "dino_adam": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float,
                            C.c_float, C.c_float, C.c_int64, C.c_void_p]),
    "dino_stage_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "dino_adam_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float,
                                C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    "dino_roll_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "dino_sgd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p]),
    "dino_sample_sqnorms": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]),
    "dino_forward_noalloc": (C.c_int, [C.POINTER(DinoNetSpec), C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                       C.c_void_p]),
    "dino_gemm_nt": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                               C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "dino_gemm_nt_acc": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32,
                                   C.c_int32, C.c_int32, C.c_void_p]),
    "dino_gemm_tn_acc": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32,
                                   C.c_int32, C.c_int32, C.c_void_p]),
    "dino_plan_profile": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "dino_debug_chain_trace": (C.c_int, [C.c_void_p]),
    "dino_launch_count": (C.c_int64, []),
    "dino_last_error": (C.c_char_p, []),
    "dino_version": (C.c_char_p, []),
}

_lib = None


class DinoLibraryError(RuntimeError):
    pass


def load():
    """Load the shared library (once) and attach the prototypes.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DinoLibraryError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built "
            "(run `make -C dino_b200/csrc` or `__graft_entry__.build()`); there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().dino_last_error()
        raise DinoLibraryError(f"{what} failed: {msg.decode() if msg else 'unknown error'}")


def launch_count() -> int:
    return int(load().dino_launch_count())
