"""ctypes binding of ``libcyten_b200.so`` (the C-ABI declared in ``include/cyten_b200.h``).

There is NO fallback: if the library is missing, or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_HERE = Path(__file__).resolve().parent
import os as _os
# CYB_LIB: tuning aid (a library built from the same sources with other compile-time tile parameters)
LIB_PATH = Path(_os.environ["CYB_LIB"]) if _os.environ.get("CYB_LIB") else _HERE / "libcyten_b200.so"

CYB_F64, CYB_C128 = 0, 1
MAX_RANK = 8
COPY_JOB_WORDS = 3 + 3 * MAX_RANK   # int64 words of struct cyb_copy_job

_vp, _i64, _int, _dbl = C.c_void_p, C.c_int64, C.c_int, C.c_double

# name -> (restype, argtypes); must list every symbol of include/cyten_b200.h
SIGNATURES = {
    "cyb_version": (_int, []),
    "cyb_last_error": (C.c_char_p, []),
    "cyb_device_info": (_int, [C.POINTER(_int)] * 3),
    "cyb_lexsort_rows": (_int, [_vp, _i64, _i64, _vp, _vp]),
    "cyb_take_rows": (_int, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _vp]),
    "cyb_row_differences": (_int, [_vp, _i64, _i64, _vp, _vp, _vp]),
    "cyb_pack_rows": (_int, [_vp, _i64, _i64, _vp, _vp, _vp]),
    "cyb_common_sorted": (_int, [_vp, _i64, _vp, _i64, _vp, _vp, _vp]),
    "cyb_copy_blocks_workspace": (_i64, [_vp, _i64]),
    "cyb_copy_blocks": (_int, [_vp, _vp, _vp, _vp, _i64, _int, _int, _vp]),
    "cyb_copy_blocks_prepare": (_int, [_vp, _i64, _vp, _i64, _vp]),
    "cyb_copy_blocks_run": (_int, [_vp, _vp, _vp, _i64, _int, _int, _vp]),
    "cyb_grouped_gemm_workspace": (_i64, [_i64]),
    "cyb_grouped_gemm_prepare": (_int, [_int, _vp, _vp, _vp, _vp, _i64, _vp, _i64, _vp]),
    "cyb_grouped_gemm_run": (_int, [_vp, _vp, _vp, _int, _vp, _vp, _vp, _vp, _i64, _vp, _int, _vp]),
    "cyb_grouped_gemm_skinny_rows": (_i64, []),
    "cyb_compose_plan_workspace": (_i64, [_i64, _i64]),
    "cyb_compose_plan": (_int, [_vp, _i64, _i64, _i64, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _i64,
                                _vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp,
                                _vp, _vp, _i64, _vp]),
    "cyb_fuse_charges": (_int, [_vp, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _vp, _vp]),
    "cyb_grouped_gemm": (_int, [_vp, _vp, _vp, _int, _vp, _vp, _vp, _vp, _vp, _i64, _vp]),
    "cyb_batched_svd": (_int, [_vp, _vp, _vp, _vp, _vp, _i64, _int, _vp, _vp]),
    "cyb_batched_qr": (_int, [_vp, _vp, _vp, _vp, _i64, _int, _vp]),
    "cyb_batched_eigh": (_int, [_vp, _vp, _vp, _vp, _i64, _int, _vp, _vp]),
    "cyb_fill_eye": (_int, [_vp, _vp, _vp, _i64, _int, _vp]),
    "cyb_truncate_select": (_int, [_vp, _i64, _i64, _i64, _dbl, _dbl, _dbl, _int, _vp, _vp, _vp]),
    "cyb_truncate_select_static": (_int, [_vp, _i64, _i64, _i64, _dbl, _dbl, _dbl, _int, _vp, _vp, _vp]),
    "cyb_sumsq": (_int, [_vp, _i64, _int, _vp, _vp]),
    "cyb_axpby": (_int, [_vp, _vp, _i64, _int, _dbl, _dbl, _vp]),
    "cyb_cutoff_inverse": (_int, [_vp, _vp, _i64, _dbl, _vp]),
    "cyb_middle_apply_max_rows": (_i64, []),
    "cyb_middle_apply_unit_cols": (_i64, []),
    "cyb_middle_apply": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _int, _vp]),
    "cyb_scale_rows_cols": (_int, [_vp, _vp, _vp, _vp, _vp, _i64, _int, _int, _vp]),
    "cyb_gather_rows_cols": (_int, [_vp, _vp, _vp, _vp, _i64, _int, _int, _vp]),
    "cyb_blas1_workspace_bytes": (_i64, []),
    "cyb_blas1_chunk_elems": (_i64, [_int]),
    "cyb_block_dot": (_int, [_vp, _vp, _vp, _vp, _i64, _i64, _int, _int, _vp, _vp, _vp]),
    "cyb_block_lincomb": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _int, _vp, _vp, _vp, _vp, _vp]),
    "cyb_scale_axis": (_int, [_vp, _vp, _vp, _vp, _i64, _i64, _vp, _int, _int, _vp]),
}

_lib = None


class CytenB200Error(RuntimeError):
    pass


def load():
    """Loads the shared library (once). Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise CytenB200Error(
                f"{LIB_PATH} is missing: build it with `python -m cyten_b200.build` "
                "(there is no CPU fallback)")
        lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().cyb_last_error().decode(errors="replace")
        raise CytenB200Error(f"{what or 'cyten_b200 call'} failed (code {rc}): {msg}")


def call(name: str, *args):
    """Calls an int-returning entry point and raises on a non-zero status."""
    check(getattr(load(), name)(*args), name)
