"""CPU: pins the oracle. The plain-C port must agree bit for bit with the reference's own object
code (when it is built here), with the committed golden vectors that object code produced, with
scipy/numpy as the independent check the reference's tests use (wilcox.test, rowMeans, rowVars),
and with the hand-worked vector of SURVEY.md Appendix D."""
import glob
import os

import numpy as np
import pytest
import scipy.sparse as sp
import scipy.stats as st

from helpers import all_pairs, as_csc, make_matrix
from oracle import native, rlevel

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _groups(g, G):
    return [np.flatnonzero(g == k).astype(np.int32) for k in range(G)]


def test_appendix_d_known_answers(port_lib):
    x = np.array([[0, 0, 1.5, 2.0, 0, 1.5, 3.0]])
    groups = [np.arange(4, dtype=np.int32), np.arange(4, 7, dtype=np.int32)]
    for m in (x, sp.csc_matrix(x)):
        mean, var = port_lib.blocked_stats(m, np.array([0, 0, 0, 0, 1, 1, 1], np.int32), 2)
        assert mean.tolist() == [[0.875, 1.5]] and var.tolist() == [[1.0625, 2.25]]
        u, t = port_lib.overlap_paired(m, [1], [2], groups, 0.0, want_ties=True)
        assert (u[0, 0], t[0, 0]) == (4.5, 30.0)
        u, t = port_lib.overlap_paired(m, [1], [2], groups, 0.5, want_ties=True)
        assert (u[0, 0], t[0, 0]) == (2.5, 12.0)
        u, t = port_lib.overlap_paired(m, [1], [2], groups, -0.5, want_ties=True)
        assert (u[0, 0], t[0, 0]) == (6.0, 6.0)
    d = rlevel.cohen_d(np.array([[0.875, 1.5]]), np.array([[1.0625, 2.25]]), np.array([0]), np.array([1]), 0.0)
    assert d[0, 0] == -0.48564293117863205
    f = rlevel.lfc_detected(np.array([[2.0, 2.0]]), np.array([0]), np.array([1]), np.array([4.0]), np.array([3.0]))
    assert abs(f[0, 0] - (-0.2630344058337938)) < 1e-15
    assert abs(rlevel.get_sigma(4.0, 3.0, 30.0) ** 2 - 7.285714285714286) < 1e-12


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_port_matches_golden_vectors(port_lib, path):
    z = np.load(path)
    x, g, block, lfc = z["x"], z["g"], z["block"], float(z["lfc"])
    G = int(g.max()) + 1
    m = sp.csc_matrix(x)
    mean, var = port_lib.blocked_stats(m, block, 4)
    assert np.array_equal(mean, z["mean"], equal_nan=True) and np.array_equal(var, z["var"], equal_nan=True)
    mean, var = port_lib.blocked_stats(x, block, 4)
    assert np.array_equal(mean, z["dense_mean"], equal_nan=True) and np.array_equal(var, z["dense_var"], equal_nan=True)
    assert np.array_equal(port_lib.overlap_paired(m, z["left"], z["right"], _groups(g, G), lfc), z["u_paired"])
    U, T = port_lib.overlap(m, _groups(g, G), lfc)
    assert np.array_equal(U, z["U"], equal_nan=True) and np.array_equal(T, z["T"], equal_nan=True)


def test_golden_vectors_exist():
    assert len(GOLDEN) >= 4


@pytest.mark.skipif(not os.path.exists("/root/reference/src/overlap_exprs.cpp"), reason="reference sources absent")
@pytest.mark.parametrize("lfc", [0.0, 0.5, -1.0, 1.0])
def test_port_matches_reference_object_code(port_lib, oracle_lib, lfc):
    assert oracle_lib.kind == "ref"
    x = make_matrix(150, 211, seed=42)
    rng = np.random.default_rng(1)
    g = rng.integers(0, 5, 211)
    groups = _groups(g, 5) + [np.zeros(0, np.int32)]
    left, right = all_pairs(6)
    block = rng.integers(0, 4, 211).astype(np.int32)
    block[::13] = native.NA_INT
    for m in (x, as_csc(x)):
        assert np.array_equal(port_lib.overlap_paired(m, left, right, groups, lfc),
                              oracle_lib.overlap_paired(m, left, right, groups, lfc))
        for a, b in zip(port_lib.overlap(m, groups, lfc), oracle_lib.overlap(m, groups, lfc)):
            assert np.array_equal(a, b, equal_nan=True)
        for a, b in zip(port_lib.blocked_stats(m, block, 5), oracle_lib.blocked_stats(m, block, 5)):
            assert np.array_equal(a, b, equal_nan=True)
    sf = rng.random(211) + 0.5
    counts = rng.poisson(0.7, (60, 211)).astype(np.float64)
    for tr, ps in ((1, 1.0), (1, 2.0), (2, 1.0)):
        for m in (counts, as_csc(counts)):
            for a, b in zip(port_lib.blocked_stats(m, block, 5, tr, sf, ps),
                            oracle_lib.blocked_stats(m, block, 5, tr, sf, ps)):
                assert np.array_equal(a, b, equal_nan=True)


def test_u_equals_mann_whitney(port_lib):
    """test-pairwise-wilcox.R:4-60 uses wilcox.test(...)$statistic; scipy's is the same number."""
    x = make_matrix(40, 90, seed=7)
    g = np.arange(90) % 3
    groups = _groups(g, 3)
    left, right = all_pairs(3)
    u, t = port_lib.overlap_paired(x, left, right, groups, 0.0, want_ties=True)
    for k, (a, b) in enumerate(zip(left, right)):
        for r in range(40):
            xa, xb = x[r, groups[a - 1]], x[r, groups[b - 1]]
            assert u[r, k] == st.mannwhitneyu(xa, xb).statistic
            _, cnt = np.unique(np.concatenate([xa, xb]), return_counts=True)
            assert t[r, k] == float((cnt * (cnt.astype(float) ** 2 - 1)).sum())
    # lfc: brute force count (test-pairwise-wilcox.R:70-113 uses mu=)
    u = port_lib.overlap_paired(x, left, right, groups, 0.4)
    for k, (a, b) in enumerate(zip(left, right)):
        xa, xb = x[:, groups[a - 1]] - 0.4, x[:, groups[b - 1]]
        brute = (xa[:, :, None] > xb[:, None, :]).sum((1, 2)) + 0.5 * (xa[:, :, None] == xb[:, None, :]).sum((1, 2))
        assert np.array_equal(u[:, k], brute)


def test_blocked_stats_equal_row_means_and_vars(port_lib):
    """test-model-var.R:14-57: mean == rowMeans, total == rowVars, per block; NA rules."""
    x = make_matrix(80, 140, seed=9)
    block = (np.arange(140) % 4).astype(np.int32)
    block[block == 3] = 2
    block[5] = 3
    mean, var = port_lib.blocked_stats(as_csc(x), block, 5)
    for b in range(3):
        assert np.allclose(mean[:, b], x[:, block == b].mean(1), rtol=1e-13, atol=1e-15)
        assert np.allclose(var[:, b], x[:, block == b].var(1, ddof=1), rtol=1e-10, atol=1e-15)
    assert np.array_equal(mean[:, 3], x[:, 5]) and np.isnan(var[:, 3]).all()
    assert np.isnan(mean[:, 4]).all() and np.isnan(var[:, 4]).all()


# ----------------------------------------------------------------------------------------------
# R-level restatement vs independent formulas (the reference's own test expectations)
# ----------------------------------------------------------------------------------------------

def test_score_markers_effects_against_definitions(port_lib):
    """test-score-markers.R:11-35 (Cohen's d), :65-89 (AUC), :177-194 (summaries), :196-214."""
    x = make_matrix(120, 200, seed=12)
    g = np.random.default_rng(2).integers(0, 4, 200)
    out = rlevel.score_markers(as_csc(x), g, full_stats=True, lib=port_lib)
    for a in range(4):
        others = [b for b in range(4) if b != a]
        xa = x[:, g == a]
        with np.errstate(invalid="ignore", divide="ignore"):
            d = np.stack([(xa.mean(1) - x[:, g == b].mean(1)) /
                          np.sqrt((xa.var(1, ddof=1) + x[:, g == b].var(1, ddof=1)) / 2) for b in others], 1)
        auc = np.stack([np.array([st.mannwhitneyu(xa[r], x[r, g == b]).statistic for r in range(120)]) /
                        (xa.shape[1] * (g == b).sum()) for b in others], 1)
        ok = np.isfinite(d).all(1)
        assert np.allclose(out[a]["full.logFC.cohen"][ok], d[ok], rtol=1e-9, atol=1e-12)
        assert np.allclose(out[a]["full.AUC"], auc, rtol=1e-14, atol=1e-15)
        assert np.allclose(out[a]["mean.AUC"], auc.mean(1), rtol=1e-13)
        assert np.allclose(out[a]["median.AUC"], np.median(auc, 1), rtol=1e-13)
        assert np.allclose(out[a]["min.AUC"], auc.min(1), rtol=1e-14) and np.allclose(out[a]["max.AUC"], auc.max(1), rtol=1e-14)
        assert np.array_equal(out[a]["rank.AUC"], rlevel.compute_min_rank(out[a]["full.AUC"]))
        assert np.allclose(out[a]["self.average"], xa.mean(1), rtol=1e-13, atol=1e-15)
        assert np.allclose(out[a]["self.detected"], (xa > 0).mean(1), rtol=1e-13, atol=1e-15)
        assert np.allclose(out[a]["other.average"], np.mean([x[:, g == b].mean(1) for b in others], 0), rtol=1e-12)
    # all-zero and all-equal rows: AUC 0.5, d = 0 (test-score-markers.R:37-63, 91-119)
    assert (out[0]["full.AUC"][:2] == 0.5).all() and (out[0]["full.logFC.cohen"][:2] == 0).all()


def test_score_markers_blocking_is_weighted_mean_of_blocks(port_lib):
    """test-score-markers.R:216-243: blocked effect = sum(w e) / sum(w), w = nL nR per block."""
    x = make_matrix(60, 300, seed=13)
    rng = np.random.default_rng(3)
    g = rng.integers(0, 3, 300)
    b = rng.integers(0, 2, 300)
    both = rlevel.score_markers(x, g, block=b, full_stats=True, lib=port_lib)
    per = [rlevel.score_markers(x[:, b == k], g[b == k], full_stats=True, lib=port_lib) for k in range(2)]
    for a in range(3):
        for eff in ("logFC.cohen", "AUC", "logFC.detected"):
            num = 0
            den = 0
            for k in range(2):
                w = per[k][a]["full." + eff + ".weight"]
                num = num + per[k][a]["full." + eff] * w[None, :]
                den = den + w
            assert np.allclose(both[a]["full." + eff], num / den[None, :], rtol=1e-12, atol=1e-13, equal_nan=True)


def test_lfc_is_a_shift_of_the_left_group(port_lib):
    """test-score-markers.R:390-415."""
    x = make_matrix(50, 120, seed=14) + 0.0
    g = np.arange(120) % 2
    shifted = x.copy()
    shifted[:, g == 1] -= 0.5
    a = rlevel.score_markers(x, g, lfc=0.5, full_stats=True, lib=port_lib)
    c = rlevel.score_markers(shifted, g, full_stats=True, lib=port_lib)
    assert np.allclose(a[1]["full.logFC.cohen"], c[1]["full.logFC.cohen"], rtol=1e-10, atol=1e-12)


def test_min_rank_definition():
    x = np.array([[3.0, 1.0], [2.0, np.nan], [3.0, 5.0], [np.nan, np.nan]])
    assert rlevel.compute_min_rank(x).tolist() == [1, 3, 1, np.iinfo(np.int32).max]


def test_pairwise_layers_match_scipy(port_lib):
    """test-pairwise-wilcox.R:4-60 and test-pairwise-t.R:4-95: p-values vs wilcox.test / t.test."""
    x = make_matrix(40, 100, seed=15)
    g = np.arange(100) % 2
    w = [q for q in rlevel.pairwise_wilcox(x, g, lib=port_lib) if (q["first"], q["second"]) == (1, 0)][0]
    ref = np.array([st.mannwhitneyu(x[r, g == 1], x[r, g == 0], use_continuity=True, method="asymptotic").pvalue
                    for r in range(40)])
    ok = ~np.isnan(ref)
    assert np.allclose(np.exp(w["log_p"][ok]), ref[ok], rtol=1e-9, atol=1e-300)
    t = [q for q in rlevel.pairwise_ttests(x, g, lib=port_lib) if (q["first"], q["second"]) == (1, 0)][0]
    with np.errstate(invalid="ignore", divide="ignore"):
        ref = st.ttest_ind(x[:, g == 1], x[:, g == 0], axis=1, equal_var=False).pvalue
    ok = np.isfinite(ref) & (x.var(1) > 1e-6)
    assert np.allclose(np.exp(t["log_p"][ok]), ref[ok], rtol=1e-8)


@pytest.mark.parametrize("sparse", [True, False])
def test_transformed_blocked_stats_pin(port_lib, oracle_lib, sparse):
    """compute_blocked_stats_lognorm / _norm (src/compute_blocked_stats.cpp:129-185): the port agrees bit
    for bit with the reference's object code, and both with rowMeans / rowVars of the transformed
    matrix (what tests/testthat/test-model-var.R checks through modelGeneVarByPoisson's callers)."""
    rng = np.random.default_rng(23)
    x = np.where(rng.random((40, 150)) < 0.35, rng.poisson(3.0, (40, 150)).astype(float), 0.0)
    block = rng.integers(0, 3, 150).astype(np.int32)
    sf = np.exp(rng.normal(0, 0.3, 150))
    m = sp.csc_matrix(x) if sparse else x
    for transform, pseudo in ((2, 1.0), (1, 1.0), (1, 0.5)):
        pm, pv = port_lib.blocked_stats(m, block, 3, transform=transform, sf=sf, pseudo=pseudo)
        om, ov = oracle_lib.blocked_stats(m, block, 3, transform=transform, sf=sf, pseudo=pseudo)
        assert np.array_equal(pm, om, equal_nan=True) and np.array_equal(pv, ov, equal_nan=True)
        y = x / sf if transform == 2 else np.log(x / sf + pseudo) / np.log(2.0)
        for b in range(3):
            keep = block == b
            assert np.allclose(pm[:, b], y[:, keep].mean(1), rtol=1e-12, atol=1e-14)
            assert np.allclose(pv[:, b], y[:, keep].var(1, ddof=1), rtol=1e-10, atol=1e-14)


def test_combine_markers_helpers_port_equals_reference_object_code(port_lib):
    """compute_Top_statistic_from_ranks / choose_middle_effect_size (src/choose_effect_size.cpp:35-97): the C
    restatement against the reference's own object code, and the R-level restatement of combineMarkers against
    the definitions the reference's tests use (tests/testthat/test-combine-markers.R:19-137)."""
    from oracle import native, rlevel
    from scipy.stats import false_discovery_control
    rng = np.random.default_rng(23)
    lp = np.log(rng.random((9, 200)))
    lp[2, 5] = np.nan
    lp[:, 7] = np.nan
    lp[:, 50:60] = np.round(lp[:, 50:60], 1)
    eff = rng.normal(size=(9, 200))
    eff[4, 9] = np.nan
    ranks = np.stack([rlevel.rank_first(lp[c]) for c in range(9)])
    libs = [port_lib] + ([native.load("ref")] if native.have_ref() else [])
    for prop in (0.0, 0.25, 0.5, 1.0):
        tops = [lib.top_from_ranks(ranks, prop) for lib in libs]
        mids = [lib.choose_middle(lp, eff, prop) for lib in libs]
        assert all(np.array_equal(t, tops[0]) for t in tops)
        assert all(np.array_equal(m, mids[0], equal_nan=True) for m in mids)
    assert tops[0][7] == native.NA_INT
    ok = ~np.isnan(lp).any(0)
    res = rlevel.combine_markers_one(lp, eff, "any", lib=port_lib)
    want = np.array([np.log(false_discovery_control(np.exp(lp[:, g])).min()) for g in np.flatnonzero(ok)])
    np.testing.assert_allclose(res["log_p"][ok], want, rtol=1e-12)
    assert np.array_equal(res["Top"][ok], ranks[:, ok].min(0))
    assert np.isnan(res["log_fdr"]).all()           # one NA p-value makes .logBH's cummin all NA
    res = rlevel.combine_markers_one(lp[:, ok], eff[:, ok], "all", lib=port_lib)
    np.testing.assert_allclose(res["log_p"], lp[:, ok].max(0))
    np.testing.assert_allclose(np.exp(res["log_fdr"]), false_discovery_control(np.exp(res["log_p"])), rtol=1e-12)
