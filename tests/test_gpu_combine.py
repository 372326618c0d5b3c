"""GPU: combineMarkers on the device (scranb200_combine_markers) against the restatement of
R/combineMarkers.R:232-330 whose Top statistic and "some" summary effect come from the reference's own
src/choose_effect_size.cpp (oracle/_ref), and against the definitions the reference's tests use
(tests/testthat/test-combine-markers.R:19-137: min of the BH-adjusted p-values, pmax, the I-th sorted
Holm-adjusted p-value, pmin / I-th of the per-contrast ranks with ties.method = "first")."""
import numpy as np
import pytest

from helpers import as_csc, make_matrix

pytestmark = pytest.mark.gpu


def _lists(rng, ngroups, ngenes, with_na=True, ties=False):
    out = []
    for a in range(ngroups):
        for b in range(ngroups):
            if a == b:
                continue
            p = np.log(rng.random(ngenes))
            if ties:
                p = np.round(p, 1)
            e = rng.normal(size=ngenes)
            if with_na:
                p[rng.random(ngenes) < 0.03] = np.nan
                e[rng.random(ngenes) < 0.02] = np.nan
            out.append(dict(first=a, second=b, log_p=p, effect=e))
    if with_na:
        for q in out:
            q["log_p"][7] = np.nan        # a gene without any p-value
    return out


@pytest.mark.parametrize("pval_type,min_prop", [("any", None), ("any", 0.5), ("some", None), ("some", 0.25), ("all", None)])
@pytest.mark.parametrize("ties", [False, True])
def test_combine_markers_matches_the_restatement(oracle_lib, pval_type, min_prop, ties):
    import scran_b200 as sb
    from oracle import native, rlevel
    rng = np.random.default_rng(17)
    lists = _lists(rng, 6, 700, with_na=True, ties=ties)
    got = sb.combine_markers(lists, pval_type, min_prop, sorted_rows=False)
    for f in range(6):
        cur = [q for q in lists if q["first"] == f]
        want = rlevel.combine_markers_one(np.stack([q["log_p"] for q in cur]), np.stack([q["effect"] for q in cur]),
                                          pval_type, min_prop, lib=oracle_lib)
        g = got[f]
        if pval_type == "any":
            assert np.array_equal(g["Top"], want["Top"])
        np.testing.assert_allclose(g["log.p.value"], want["log_p"], rtol=1e-13, atol=0, equal_nan=True)
        np.testing.assert_allclose(g["log.FDR"], want["log_fdr"], rtol=1e-13, atol=0, equal_nan=True)
        assert np.array_equal(g["summary.effect"], want["summary"], equal_nan=True)
        assert np.array_equal(g["order"], want["order"])


def test_combine_markers_matches_the_reference_tests_definitions():
    import scran_b200 as sb
    from scipy.stats import false_discovery_control
    rng = np.random.default_rng(18)
    ngroups, ngenes = 10, 100                      # the shapes of test-combine-markers.R:7-15
    lists = _lists(rng, ngroups, ngenes, with_na=False)
    any_ = sb.combine_markers(lists, "any", sorted_rows=False)
    mid = sb.combine_markers(lists, "any", 0.5)
    all_ = sb.combine_markers(lists, "all", sorted_rows=False)
    some = sb.combine_markers(lists, "some", sorted_rows=False)
    quart = sb.combine_markers(lists, "some", 0.25, sorted_rows=False)
    for f in range(ngroups):
        cur = [q for q in lists if q["first"] == f]
        pmat = np.exp(np.stack([q["log_p"] for q in cur]))          # contrasts x genes
        emat = np.stack([q["effect"] for q in cur])
        any_p = np.array([false_discovery_control(pmat[:, g]).min() for g in range(ngenes)])
        np.testing.assert_allclose(np.exp(any_[f]["log.p.value"]), any_p, rtol=1e-12)
        np.testing.assert_allclose(np.exp(any_[f]["log.FDR"]), false_discovery_control(any_p), rtol=1e-12)
        ranks = np.stack([np.argsort(np.argsort(pmat[c], kind="stable"), kind="stable") + 1 for c in range(pmat.shape[0])])
        min_rank = ranks.min(0)
        assert np.array_equal(any_[f]["Top"], min_rank)
        assert np.array_equal(any_[f]["order"], np.argsort(min_rank, kind="stable"))
        mid_rank = np.sort(ranks, axis=0)[4]                          # sort(x)[5] of 9 contrasts
        assert np.array_equal(mid[f]["Top"], mid_rank[np.argsort(mid_rank, kind="stable")])
        assert np.array_equal(any_[f]["summary.effect"], emat[pmat.argmin(0), np.arange(ngenes)])
        np.testing.assert_allclose(np.exp(all_[f]["log.p.value"]), pmat.max(0), rtol=1e-12)
        assert np.array_equal(all_[f]["order"], np.argsort(all_[f]["log.p.value"], kind="stable"))
        assert np.array_equal(all_[f]["summary.effect"], emat[pmat.argmax(0), np.arange(ngenes)])

        def holm(p):
            o = np.argsort(p)
            n = len(p)
            adj = np.maximum.accumulate(np.minimum(1.0, p[o] * (n - np.arange(n))))
            return adj                                                # sorted Holm-adjusted p-values
        for res, prop in ((some, 0.5), (quart, 0.25)):
            idx = int(np.ceil((ngroups - 1) * prop)) - 1
            want = np.array([holm(pmat[:, g])[idx] for g in range(ngenes)])
            np.testing.assert_allclose(np.exp(res[f]["log.p.value"]), want, rtol=1e-12)
            eff = np.array([emat[np.argsort(pmat[:, g])[idx], g] for g in range(ngenes)])
            assert np.array_equal(res[f]["summary.effect"], eff)


def test_find_markers_wilcox_end_to_end(port_lib):
    """findMarkers(test.type = "wilcox") = pairwiseWilcox + combineMarkers (R/findMarkers.R:123-156)."""
    import scran_b200 as sb
    from oracle import rlevel
    rng = np.random.default_rng(19)
    x = make_matrix(300, 260, seed=119, rate=1.0)
    groups = rng.integers(0, 4, 260)
    block = rng.integers(0, 2, 260)
    got = sb.find_markers_wilcox(as_csc(x), groups, block=block, pval_type="any")
    pairs = rlevel.pairwise_wilcox(as_csc(x), groups, block=block, lib=port_lib)
    for f in range(4):
        cur = [q for q in pairs if q["first"] == f]
        want = rlevel.combine_markers_one(np.stack([q["log_p"] for q in cur]), np.stack([q["effect"] for q in cur]), "any",
                                          None, lib=port_lib)
        o = want["order"]
        np.testing.assert_allclose(got[f]["log.p.value"], want["log_p"][got[f]["order"]], rtol=1e-9, atol=1e-12)
        # p-values that agree to 1e-10 may rank differently only among near-ties: compare Top through its definition
        assert np.array_equal(np.sort(got[f]["Top"]), np.sort(want["Top"][o]))
        assert (np.diff(got[f]["Top"]) >= 0).all()
