"""ctypes binding of libplaincv_b200.so (the C ABI declared in include/plaincv_b200.h).

There is deliberately no fallback: if the shared library is missing or a symbol cannot be resolved the import of
any compute path raises, and a kernel returning a non-zero status raises `KernelError`.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_double, c_float, c_int, c_longlong, c_void_p
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libplaincv_b200.so")
if os.environ.get("PCV_LIB_PATH"):  # debug / A-B builds of the same library (e.g. -DPCV_HANG_DEBUG); still no fallback
    LIB_PATH = os.path.abspath(os.environ["PCV_LIB_PATH"])

PCV_EPI_BF16 = 0
PCV_EPI_F32_ATOMIC = 1
PCV_EPI_F32 = 2

_ERR = {
    -1: "PCV_ERR_ARG (bad shape / null pointer / unsupported combination)",
    -2: "PCV_ERR_ALIGN (pointer or leading dimension violates the 16-byte rule)",
    -3: "PCV_ERR_TMAP (cuTensorMapEncodeTiled failed)",
    -4: "PCV_ERR_DRIVER (driver entry point unavailable)",
    -5: "PCV_ERR_SCRATCH (problem larger than the built-in scratch of a *_static entry point)",
}

# name -> (restype, argtypes); mirrors include/plaincv_b200.h one to one
SIGNATURES = {
    "pcv_version": (c_int, []),
    "pcv_launch_count": (c_longlong, []),
    "pcv_gemm_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                              c_int, c_int, c_float, c_void_p]),
    "pcv_gemm_set_mode": (c_int, [c_int]),
    "pcv_gemm_pair_launch_count": (c_longlong, []),
    "pcv_set_pdl": (c_int, [c_int]),
    "pcv_gemm_bf16_ce": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                 c_void_p, c_void_p]),
    "pcv_ce_from_stats": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                  c_float, c_void_p]),
    "pcv_cast_pad_bf16": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "pcv_embed_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "pcv_embed_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "pcv_rmsnorm_residual_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                         c_float, c_void_p]),
    "pcv_rmsnorm_residual_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                         c_int, c_void_p]),
    "pcv_swiglu_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "pcv_swiglu_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "pcv_gemm_bf16_rope": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                   c_int, c_int, c_void_p]),
    "pcv_gemm_bf16_swiglu": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                     c_int, c_void_p]),
    "pcv_gemm_bf16_dswiglu": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                      c_int, c_void_p]),
    "pcv_rope_qk": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "pcv_attn_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "pcv_attn_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                             c_int, c_int, c_int, c_void_p]),
    "pcv_attn_bwd_static": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                    c_int, c_void_p]),
    "pcv_softmax_ce": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float,
                               c_void_p]),
    "pcv_mean2": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "pcv_act_fwd": (c_int, [c_void_p, c_void_p, c_longlong, c_int, c_void_p]),
    "pcv_act_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_longlong, c_int, c_void_p]),
    "pcv_add_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_longlong, c_void_p]),
    "pcv_cast_f32_bf16": (c_int, [c_void_p, c_void_p, c_longlong, c_int, c_void_p]),
    "pcv_sumsq_partials": (c_int, [c_void_p, c_longlong, c_void_p, c_int, c_void_p]),
    "pcv_adamw_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_longlong, c_void_p, c_int, c_void_p,
                               c_double, c_double, c_double, c_double, c_double, c_double, c_int, c_void_p]),
    "pcv_rmsnorm_residual_jvp": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_int, c_int, c_void_p]),
    "pcv_swiglu_jvp": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "pcv_act_jvp": (c_int, [c_void_p, c_void_p, c_void_p, c_longlong, c_int, c_void_p]),
    "pcv_attn_jvp": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                             c_void_p]),
    "pcv_ce_hessian_vec": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p]),
}


class KernelError(RuntimeError):
    pass


_lib: Optional[ctypes.CDLL] = None


def load() -> ctypes.CDLL:
    """Load the shared library (once) and type every entry point.  Raises if it is absent: there is no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C plaincv_b200/csrc`). plaincv_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    # measurement switches (A/B runs only): PCV_GEMM_MODE = 0 auto | 1 single-CTA tiles | 2 CTA-pair tiles,
    # PCV_GEMM_TAIL = 0 keeps the last partial wave of a CTA-pair GEMM as full 256x256 tiles
    if "PCV_GEMM_MODE" in os.environ:
        lib.pcv_gemm_set_mode(int(os.environ["PCV_GEMM_MODE"]))
    if "PCV_GEMM_TAIL" in os.environ:
        lib.pcv_gemm_set_mode(4 if os.environ["PCV_GEMM_TAIL"] != "0" else 3)
    # PCV_GEMM_DBG (leave out operand loads / epilogue stores) exists only in a -DPCV_MEASURE build of the library
    # (make XFLAGS=-DPCV_MEASURE); the shipped library rejects it, so a stray variable cannot corrupt results
    if "PCV_GEMM_DBG" in os.environ:
        if lib.pcv_gemm_set_mode(100 + int(os.environ["PCV_GEMM_DBG"])) != 0:
            raise RuntimeError("PCV_GEMM_DBG needs a measurement build of libplaincv_b200.so (make XFLAGS=-DPCV_MEASURE)")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = _ERR.get(rc, f"cudaError_t {rc}" if rc > 0 else f"status {rc}")
        raise KernelError(f"{what} failed: {msg}")
