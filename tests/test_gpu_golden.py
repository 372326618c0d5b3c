"""GPU: the CUDA path against the committed golden vectors (tests/golden/*.npz), i.e. against what
the reference's own object code returned for these inputs when the fixtures were made
(tests/golden/make_golden.py). Needs neither /root/reference nor the oracle at run time."""
import glob
import os

import numpy as np
import pytest
import scipy.sparse as sp

pytestmark = pytest.mark.gpu

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


@pytest.fixture(scope="module")
def sb():
    import scran_b200
    return scran_b200


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_device_matches_golden_vectors(sb, path):
    z = np.load(path)
    x, g, block, lfc = z["x"], z["g"], z["block"], float(z["lfc"])
    G = int(g.max()) + 1
    groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(G)]
    m = sp.csc_matrix(x)
    # compute_blocked_stats_none, sparse and dense input (block 3 is empty, some cells are NA)
    mean, var = sb.compute_blocked_stats_none(m, block, 4)
    assert np.array_equal(mean, z["mean"], equal_nan=True) and np.array_equal(var, z["var"], equal_nan=True)
    mean, var = sb.compute_blocked_stats_none(x, block, 4)
    assert np.array_equal(mean, z["dense_mean"], equal_nan=True) and np.array_equal(var, z["dense_var"], equal_nan=True)
    # overlap_exprs_paired and overlap_exprs: U and the tie term, bit for bit
    assert np.array_equal(sb.overlap_exprs_paired(m, z["left"], z["right"], groups, lfc), z["u_paired"])
    U, T = sb.overlap_exprs(m, groups, lfc)
    assert np.array_equal(U, z["U"], equal_nan=True) and np.array_equal(T, z["T"], equal_nan=True)
