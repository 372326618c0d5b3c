"""Seeded inputs shaped like the reference's own test data (tests/testthat/test-score-markers.R:
4-9: Poisson-ish log-counts, a few hundred genes x cells), plus the degenerate rows its tests
exercise: all-zero, all-equal, integer-valued (ties), negative values."""
import numpy as np
import scipy.sparse as sp


def make_matrix(ngenes, ncells, seed, rate=0.6, integer_rows=True):
    rng = np.random.default_rng(seed)
    lam = rate * np.exp(rng.normal(0, 1, ngenes))[:, None] * np.ones((1, ncells))
    x = rng.poisson(lam).astype(np.float64)
    nz = x > 0
    x[nz] = np.log2(x[nz] + rng.random(nz.sum()))
    if ngenes >= 8:
        x[0] = 0.0                                   # all-zero gene
        x[1] = 1.5                                   # all-equal gene
        if integer_rows:
            x[2] = rng.poisson(1.0, ncells)          # heavy ties
            x[3] = rng.poisson(3.0, ncells)          # dense, ties
        x[4] = -x[4]                                 # negative values
        x[5, ::2] = -x[5, ::2]                       # mixed signs
        x[6] = (rng.random(ncells) < 0.02) * 2.0     # ultra sparse
    return x


def as_csc(x):
    return sp.csc_matrix(x)


def all_pairs(G):
    left, right = [], []
    for a in range(G):
        for b in range(G):
            if a != b:
                left.append(a + 1)
                right.append(b + 1)
    return np.array(left, np.int32), np.array(right, np.int32)


def assert_tables_match(got, want, rtol=1e-12, skip_selfother=False):
    """got/want: OrderedDict group -> OrderedDict column -> array, as score_markers returns."""
    assert list(got.keys()) == list(want.keys())
    for g in want:
        wcols = [c for c in want[g] if not c.endswith(".right")]
        for c in wcols:
            a, b = got[g][c], want[g][c]
            if skip_selfother and c.startswith(("self.", "other.")):
                continue
            if c.startswith("rank."):
                assert np.array_equal(np.asarray(a), np.asarray(b)), "min-rank differs: %r %s" % (g, c)
            else:
                a = np.asarray(a, np.float64)
                b = np.asarray(b, np.float64)
                assert a.shape == b.shape, (g, c, a.shape, b.shape)
                ok = np.isclose(a, b, rtol=rtol, atol=1e-12, equal_nan=True) | (a == b)
                assert ok.all(), "%r %s: max abs diff %g" % (g, c, np.nanmax(np.abs(a - b)))
        if any(c.endswith(".right") for c in want[g]):
            for c in want[g]:
                if c.endswith(".right"):
                    assert list(got[g][c]) == list(want[g][c])
