"""CPU: the Rcpp-side glue (glue/scran_b200_glue.cpp: the bodies a maintainer puts behind the reference's
.Call symbols, /root/reference/src/RcppExports.cpp:52-89, 162-186) goes through a compiler. R and Rcpp are
not installed, so it is compiled against the stand-in Rcpp.h that also compiles the reference's own sources,
linked against libscranb200.so, and its entry points are looked up. Without a GPU a call must fail with the
library's "no CUDA device" message, not crash."""
import numpy as np
import pytest


def test_glue_compiles_links_and_fails_loudly_without_a_gpu():
    import torch

    from glue_build import build, call, mat_args
    lib = build()
    for name in ("glue_blocked_stats", "glue_overlap_paired", "glue_overlap", "glue_score_markers"):
        assert hasattr(lib, name)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the calls themselves are tested in test_gpu_glue.py")
    x = np.eye(4)
    mean, var = np.zeros((1, 4)), np.zeros((1, 4))
    rc = call(lib.glue_blocked_stats, *mat_args(x), np.zeros(4, np.int32), 1, None, 1.0, mean, var)
    assert rc != 0
    assert "no CUDA device" in lib.glue_last_error().decode()
