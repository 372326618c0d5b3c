"""GPU: the reference's .Call-level functions through the compiled Rcpp glue (glue/scran_b200_glue.cpp, built
against the stand-in Rcpp.h) against the reference's own object code: the SEXP-level signatures of
/root/reference/src/RcppExports.cpp:52-89, 162-186 exercised end to end - dgCMatrix slots and ordinary
matrices in, list(mean, var) / npairs x ngenes matrix / ragged list of matrices out."""
import numpy as np
import pytest

from helpers import all_pairs, as_csc, make_matrix

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("sparse", [True, False])
def test_blocked_stats_and_overlap_through_the_glue(oracle_lib, sparse):
    from glue_build import build, call, mat_args
    from oracle import native
    lib = build()
    x = make_matrix(120, 211, seed=71)
    m = as_csc(x) if sparse else x
    rng = np.random.default_rng(5)
    block = rng.integers(0, 4, 211).astype(np.int32)
    block[::13] = native.NA_INT
    args = mat_args(m)
    mean, var = np.zeros((4, 120)), np.zeros((4, 120))
    assert call(lib.glue_blocked_stats, *args, block, 4, None, 1.0, mean, var) == 0, lib.glue_last_error()
    rmean, rvar = oracle_lib.blocked_stats(m, block, 4)
    assert np.array_equal(mean.T, rmean, equal_nan=True) and np.array_equal(var.T, rvar, equal_nan=True)
    sf = np.exp(rng.normal(0, 0.3, 211))
    assert call(lib.glue_blocked_stats, *args, block, 4, sf, 1.0, mean, var) == 0, lib.glue_last_error()
    rmean, rvar = oracle_lib.blocked_stats(m, block, 4, transform=1, sf=sf, pseudo=1.0)
    np.testing.assert_allclose(mean.T, rmean, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(var.T, rvar, rtol=1e-12, atol=1e-13)

    g = rng.integers(0, 5, 211)
    groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(5)]
    gptr = np.zeros(6, np.int64)
    gptr[1:] = np.cumsum([len(c) for c in groups])
    cells = np.concatenate(groups).astype(np.int32)
    left, right = all_pairs(5)
    for lfc in (0.0, 0.5):
        out = np.zeros((120, len(left)))
        assert call(lib.glue_overlap_paired, *args, left, right, len(left), 5, gptr, cells, lfc, out) == 0, lib.glue_last_error()
        assert np.array_equal(out, oracle_lib.overlap_paired(m, left, right, groups, lfc))
        u, t = np.zeros((5, 5, 120)), np.zeros((5, 5, 120))
        assert call(lib.glue_overlap, *args, 5, gptr, cells, lfc, u, t) == 0, lib.glue_last_error()
        ru, rt = oracle_lib.overlap(m, groups, lfc)
        assert np.array_equal(u, ru, equal_nan=True) and np.array_equal(t, rt, equal_nan=True)
    # the reference's error text survives the boundary (src/overlap_exprs.cpp:25)
    bad = cells.copy()
    bad[3] = 999
    out = np.zeros((120, len(left)))
    assert call(lib.glue_overlap_paired, *args, left, right, len(left), 5, gptr, bad, 0.0, out) != 0
    assert "indices in 'groups' out of range" in lib.glue_last_error().decode()


def test_fused_score_markers_through_the_glue():
    import scran_b200 as sb
    from glue_build import build, call, mat_args
    lib = build()
    x = make_matrix(150, 180, seed=72)
    rng = np.random.default_rng(6)
    groups = rng.integers(0, 4, 180).astype(np.int32)
    block = rng.integers(0, 2, 180).astype(np.int32)
    out = np.zeros((4, 19, 150))
    assert call(lib.glue_score_markers, *mat_args(as_csc(x)), groups, 4, block, 0.0, out) == 0, lib.glue_last_error()
    want = sb.score_markers(as_csc(x), groups, block=block)
    for k in range(4):
        for c, name in enumerate(sb.COLUMNS):
            w = np.asarray(want[k][name], np.float64)
            g = out[k, c]
            if name.startswith("rank."):
                g = np.where(np.isfinite(g), g, np.iinfo(np.int32).max)
            assert np.array_equal(g, w, equal_nan=True), (k, name)
