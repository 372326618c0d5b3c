"""CPU: the C-ABI library loads, exports every symbol include/scranb200.h declares, and refuses
to compute without a GPU (no CPU fallback). No compute calls are made here."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from scran_b200 import _lib


def test_library_is_built_in_tree():
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"


def test_every_declared_symbol_is_exported_and_bound():
    declared = _lib.declared_symbols()
    assert len(declared) >= 28
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    missing = [s for s in declared if s not in exported]
    assert not missing, missing
    bound = set(_lib._signatures())
    assert set(declared) == bound, (sorted(set(declared) - bound), sorted(bound - set(declared)))
    handle = _lib.lib()
    for name in declared:
        assert getattr(handle, name) is not None


def test_abi_has_no_torch_or_cxx_types():
    import re
    text = re.sub(r"/\*.*?\*/", "", open(_lib.HEADER_PATH).read(), flags=re.S)  # declarations only
    for banned in ("torch", "at::", "std::", "Rcpp", "SEXP", "&"):
        assert banned not in text, banned


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = _lib.lib().scranb200_ctx_create(0, C.byref(h))
    assert rc != 0
    msg = _lib.lib().scranb200_last_error().decode()
    assert "no CUDA device" in msg and "no CPU fallback" in msg
    import scran_b200
    with pytest.raises(scran_b200.ScranB200Error):
        scran_b200.compute_blocked_stats_none(np.zeros((3, 4)), np.zeros(4, np.int32), 1)


def test_product_never_imports_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "scran_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower().replace("no cpu fallback", ""), os.path.join(dirpath, f)
