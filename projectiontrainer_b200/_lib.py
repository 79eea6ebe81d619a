"""ctypes binding of libpt_stage1.so (the C ABI in include/pt_stage1.h).

There is deliberately no fallback: if the shared library is missing, or a call
fails, a RuntimeError is raised.  Nothing here routes to PyTorch math.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

PKG = Path(__file__).resolve().parent
# PT_STAGE1_LIB: A/B probes load another build of the same library (tools/); never set in the product path
LIB_PATH = Path(os.environ.get("PT_STAGE1_LIB") or PKG / "libpt_stage1.so")

PT_BF16, PT_F32, PT_I64, PT_I32 = 0, 1, 2, 3
BAND_NONE, BAND_SCORES, BAND_K_BY_QUERY, BAND_K_BY_KEY = 0, 1, 2, 3
ACT_NONE, ACT_GELU_TANH, ACT_GELU_ERF, ACT_GEGLU, ACT_GEGLU_BWD, ACT_SOFTMAX_BWD = 0, 1, 2, 3, 4, 5
ACT_SWIGLU, ACT_SWIGLU_BWD, ACT_GELU_ERF_BWD, ACT_ATTN_DELTA, ACT_CE_STATS, ACT_QKNORM_ROPE = 6, 7, 8, 9, 10, 11


class GemmArgs(C.Structure):
    _fields_ = [
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("nz", C.c_int32 * 3),
        ("A", C.c_void_p), ("lda", C.c_int64), ("a_zs", C.c_int64 * 3), ("a_mn_major", C.c_int32),
        ("B", C.c_void_p), ("ldb", C.c_int64), ("b_zs", C.c_int64 * 3), ("b_mn_major", C.c_int32),
        ("C", C.c_void_p), ("c_dtype", C.c_int32), ("ldc", C.c_int64), ("c_zs", C.c_int64 * 3),
        ("C_pre", C.c_void_p),
        ("bias", C.c_void_p),
        ("residual", C.c_void_p), ("r_dtype", C.c_int32), ("ldr", C.c_int64), ("r_zs", C.c_int64 * 3),
        ("act", C.c_int32), ("alpha", C.c_float),
        ("kg", C.c_int32), ("a_kgs", C.c_int64), ("b_kgs", C.c_int64),
        ("band", C.c_int32), ("band_window", C.c_int32),
        ("aux", C.c_void_p), ("ld_aux", C.c_int64),
        ("bn_hint", C.c_int32),
        ("aux2", C.c_void_p), ("aux_i0", C.c_int32), ("aux_i1", C.c_int32),
        ("A2", C.c_void_p), ("lda2", C.c_int64), ("B2", C.c_void_p), ("ldb2", C.c_int64), ("K2", C.c_int32),
        ("ksplit", C.c_int32),
        ("eps", C.c_float),
    ]


_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raise loudly if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m projectiontrainer_b200.build` "
            "(or __graft_entry__.build()). There is no CPU / PyTorch fallback for this path.")
    L = C.CDLL(str(LIB_PATH))
    L.pt_last_error.restype = C.c_char_p
    L.pt_image_preprocess_workspace.restype = C.c_longlong
    _lib = L
    return L


def last_error() -> str:
    return lib().pt_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed (status {rc}): {last_error()}")


# kernels launched per C-ABI call (for the bench's gpu_launches claim)
KERNELS_PER_CALL = {"pt_ce_fwd_bwd": 3, "pt_ce_from_partials": 3, "pt_image_preprocess": 3, "pt_jpeg_decode_batch": 3}
launch_count = 0


def call(name: str, *args) -> None:
    global launch_count
    fn = getattr(lib(), name)
    check(fn(*args), name)
    launch_count += KERNELS_PER_CALL.get(name, 1)
