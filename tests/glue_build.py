"""Builds glue/scran_b200_glue.cpp + tests/glue_driver.cpp against oracle/stubs/Rcpp.h into a scratch
shared library that links libscranb200.so (test scaffolding shared by test_glue_cpu.py / test_gpu_glue.py)."""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None


def build():
    global _lib
    if _lib is not None:
        return _lib
    out = os.path.join(tempfile.mkdtemp(prefix="glue_"), "libglue.so")
    libdir = os.path.join(ROOT, "scran_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "oracle", "stubs"),
                           "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "glue", "scran_b200_glue.cpp"),
                           os.path.join(ROOT, "tests", "glue_driver.cpp"), "-o", out, "-L" + libdir, "-lscranb200",
                           "-Wl,-rpath," + libdir])
    lib = C.CDLL(out)
    lib.glue_last_error.restype = C.c_char_p
    _lib = lib
    return lib


def mat_args(m):
    """scipy CSC or dense [nrow, ncol] -> the driver's matrix arguments (kept alive by the caller)."""
    import scipy.sparse as sp
    if sp.issparse(m):
        m = m.tocsc()
        m.sort_indices()
        p = np.ascontiguousarray(m.indptr, np.int64)
        i = np.ascontiguousarray(m.indices, np.int32)
        x = np.ascontiguousarray(m.data, np.float64)
        return (m.shape[0], m.shape[1], p, i, x, 1)
    x = np.ascontiguousarray(np.asarray(m, np.float64).T).ravel()
    return (m.shape[0], m.shape[1], np.zeros(1, np.int64), np.zeros(1, np.int32), x, 0)


def call(fn, *args):
    conv = []
    for a in args:
        if isinstance(a, np.ndarray):
            conv.append(a.ctypes.data_as(C.c_void_p))
        elif isinstance(a, float):
            conv.append(C.c_double(a))
        elif a is None:
            conv.append(None)
        else:
            conv.append(C.c_int64(int(a)))
    return fn(*conv)
