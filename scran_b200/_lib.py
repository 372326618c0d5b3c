"""ctypes binding of libscranb200.so (the C ABI in include/scranb200.h).

The library is built in-tree by ``__graft_entry__.build()`` (scran_b200/csrc/Makefile). There is
no CPU fallback: if the shared object is missing, or no CUDA device is present when a context is
created, the call fails loudly.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libscranb200.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "scranb200.h")
NCOLS = 19
NA_INT = np.iinfo(np.int32).min

_lib = None

c_ctx = C.c_void_p
c_mat = C.c_void_p
c_plan = C.c_void_p
_p = C.c_void_p  # raw pointer (host numpy buffer or device address)


class ScranB200Error(RuntimeError):
    pass


def _signatures():
    i64, i32, dbl, u64 = C.c_int64, C.c_int32, C.c_double, C.c_uint64
    PP = C.POINTER(C.c_void_p)
    return {
        "scranb200_last_error": (C.c_char_p, []),
        "scranb200_version": (C.c_int, []),
        "scranb200_ctx_create": (C.c_int, [C.c_int, PP]),
        "scranb200_ctx_destroy": (None, [c_ctx]),
        "scranb200_ctx_sync": (C.c_int, [c_ctx]),
        "scranb200_ctx_stream": (u64, [c_ctx]),
        "scranb200_ctx_launches": (i64, [c_ctx]),
        "scranb200_ctx_last_timing": (C.c_int, [c_ctx, _p, C.c_int]),
        "scranb200_upload_csc": (C.c_int, [c_ctx, i64, i64, _p, _p, _p, PP]),
        "scranb200_upload_csc32": (C.c_int, [c_ctx, i64, i64, _p, _p, _p, PP]),
        "scranb200_mat_begin": (C.c_int, [c_ctx, i64, i64, i64, PP]),
        "scranb200_mat_append_columns": (C.c_int, [C.c_void_p, i64, _p, _p, _p]),
        "scranb200_mat_append_columns32": (C.c_int, [C.c_void_p, i64, _p, _p, _p]),
        "scranb200_mat_finish": (C.c_int, [C.c_void_p, PP]),
        "scranb200_mat_abort": (None, [C.c_void_p]),
        "scranb200_upload_dense": (C.c_int, [c_ctx, i64, i64, _p, PP]),
        "scranb200_wrap_device_csc": (C.c_int, [c_ctx, i64, i64, i64, _p, _p, _p, PP]),
        "scranb200_mat_free": (None, [c_mat]),
        "scranb200_mat_nnz": (i64, [c_mat]),
        "scranb200_download_csc": (C.c_int, [c_ctx, c_mat, _p, _p, _p]),
        "scranb200_blocked_stats": (C.c_int, [c_ctx, c_mat, _p, C.c_int, _p, _p, _p]),
        "scranb200_pairwise_wilcox": (C.c_int, [c_ctx, c_mat, _p, C.c_int, _p, C.c_int, C.c_double, C.c_int, _p, _p, _p]),
        "scranb200_pairwise_t_tests": (C.c_int, [c_ctx, c_mat, _p, C.c_int, _p, C.c_int, C.c_double, C.c_int, C.c_int, _p, _p, _p]),
        "scranb200_combine_markers": (C.c_int, [c_ctx, i64, C.c_int, _p, _p, C.c_int, C.c_double, _p, _p, _p, _p, _p]),
        "scranb200_blocked_stats_transformed": (C.c_int, [c_ctx, c_mat, _p, C.c_int, C.c_int, _p, C.c_double, _p, _p]),
        "scranb200_overlap_exprs_paired": (C.c_int, [c_ctx, c_mat, _p, _p, i64, C.c_int, _p, _p, dbl, _p, _p]),
        "scranb200_overlap_exprs": (C.c_int, [c_ctx, c_mat, C.c_int, _p, _p, dbl, _p, _p]),
        "scranb200_plan_create": (C.c_int, [c_ctx, i64, _p, C.c_int, _p, _p, _p, i64, dbl, PP]),
        "scranb200_plan_destroy": (None, [c_plan]),
        "scranb200_plan_info": (C.c_int, [c_plan, _p, C.c_int]),
        "scranb200_plan_tables": (C.c_int, [c_plan, _p, _p, _p, _p]),
        "scranb200_plan_pairs": (C.c_int, [c_plan] + [_p] * 12),
        "scranb200_score_markers": (C.c_int, [c_ctx, c_mat, c_plan, _p, _p]),
        "scranb200_summary_marker_stats": (C.c_int, [c_ctx, c_mat, c_plan, _p]),
        "scranb200_score_effects_dev": (C.c_int, [c_ctx, c_mat, c_plan, i64, i64, _p, _p]),
        "scranb200_score_minrank_dev": (C.c_int, [c_ctx, c_plan, i64, i64, _p, i64, i64, _p]),
        "scranb200_score_finish_dev": (C.c_int, [c_ctx, c_plan, i64, i64, _p, _p]),
        "scranb200_fill_i32": (C.c_int, [c_ctx, _p, i64, i32]),
        "scranb200_compute_min_rank": (C.c_int, [c_ctx, _p, i64, i64, _p]),
        "scranb200_compute_min_rank_static": (C.c_int, [c_ctx, _p, i64, i64, _p]),
        "scranb200_comm_unique_id": (C.c_int, [_p]),
        "scranb200_ctx_comm_init": (C.c_int, [c_ctx, C.c_int, C.c_int, _p]),
        "scranb200_ctx_comm_init_all": (C.c_int, [_p, C.c_int]),
        "scranb200_exchange_plan": (C.c_int, [i64, C.c_int, C.c_int, _p, _p, _p, _p, _p, _p]),
        "scranb200_balance_rows": (C.c_int, [_p, i64, C.c_int, _p]),
        "scranb200_score_markers_sharded": (C.c_int, [c_ctx, c_mat, c_plan, i64, _p, C.c_int, _p, _p]),
        "scranb200_synth_row_weights": (C.c_int, [i64, i64, dbl, u64, _p]),
        "scranb200_synth_csc": (C.c_int, [c_ctx, i64, i64, i64, i64, _p, C.c_int, dbl, u64, PP]),
    }


def declared_symbols():
    """Names of the functions declared in include/scranb200.h (for the export test)."""
    import re
    text = open(HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(scranb200_[a-z0-9_]+)\s*\(", text)))


def lib():
    """Load the shared library (once). Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ScranB200Error(
                "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)" % LIB_PATH)
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _signatures().items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        raise ScranB200Error(lib().scranb200_last_error().decode())


def ptr(a):
    """Raw address of a numpy array / torch tensor / int / None."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    raise TypeError(type(a))
