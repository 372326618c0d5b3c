"""GPU parity: libscranb200 through the C ABI vs the CPU oracle on the same seeded inputs.
Integer results (U, tie term, min-rank) must be bit-exact; means and variances are bit-exact too
(same operation order, no FMA); everything else within 1e-12 relative."""
import numpy as np
import pytest

from helpers import all_pairs, as_csc, assert_tables_match, make_matrix

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sb():
    import scran_b200
    scran_b200.default_context()
    return scran_b200


# ---------------------------------------------------------------------------------------------
# compute_blocked_stats_none  (tests/testthat/test-model-var.R:14-57, 97-131, 195-212)
# ---------------------------------------------------------------------------------------------

@pytest.mark.parametrize("sparse", [True, False])
def test_blocked_stats_bit_exact(sb, oracle_lib, sparse):
    from oracle import native
    x = make_matrix(300, 257, seed=11)
    rng = np.random.default_rng(3)
    block = rng.integers(0, 5, 257).astype(np.int32)
    block[::19] = native.NA_INT          # NA block ids are skipped
    block[block == 4] = 3
    block[7] = 4                         # a one-cell block; block 5 stays empty
    m = as_csc(x) if sparse else x
    mean, var, det = sb.compute_blocked_stats_none(m, block, 6, want_detected=True)
    rmean, rvar = oracle_lib.blocked_stats(m, block, 6)
    assert np.array_equal(mean, rmean, equal_nan=True)
    assert np.array_equal(var, rvar, equal_nan=True)
    assert np.isnan(mean[:, 5]).all() and np.isnan(var[:, 5]).all() and np.isnan(var[:, 4]).all()
    for b in range(5):
        assert np.array_equal(det[:, b], (x[:, block == b] > 0).sum(1))
    # the reference's own check: equal to rowMeans / rowVars (test-model-var.R:14-30)
    keep = block == 1
    assert np.allclose(mean[:, 1], x[:, keep].mean(1), rtol=1e-12, atol=1e-14)
    assert np.allclose(var[:, 1], x[:, keep].var(1, ddof=1), rtol=1e-10, atol=1e-14)


@pytest.mark.parametrize("sparse", [True, False])
def test_blocked_stats_norm_and_lognorm(sb, oracle_lib, sparse):
    """compute_blocked_stats_norm / _lognorm (src/compute_blocked_stats.cpp:129-185): norm bit-exact;
    lognorm within 1e-12 (device log), including pseudo != 1 where the reference reads dense columns."""
    from oracle import native
    rng = np.random.default_rng(31)
    x = np.where(rng.random((90, 301)) < 0.3, rng.poisson(4.0, (90, 301)).astype(float), 0.0)
    x[5] = 0.0
    x[6] = rng.poisson(30.0, 301) + 1.0
    block = rng.integers(0, 4, 301).astype(np.int32)
    block[::17] = native.NA_INT
    sf = np.exp(rng.normal(0, 0.4, 301))
    m = as_csc(x) if sparse else x
    mean, var = sb.compute_blocked_stats_norm(m, block, 4, sf)
    rmean, rvar = oracle_lib.blocked_stats(m, block, 4, transform=2, sf=sf)
    assert np.array_equal(mean, rmean, equal_nan=True)
    assert np.array_equal(var, rvar, equal_nan=True)
    for pseudo in (1.0, 0.5, 2.0):
        mean, var = sb.compute_blocked_stats_lognorm(m, block, 4, sf, pseudo)
        rmean, rvar = oracle_lib.blocked_stats(m, block, 4, transform=1, sf=sf, pseudo=pseudo)
        np.testing.assert_allclose(mean, rmean, rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(var, rvar, rtol=1e-12, atol=1e-13)


def test_blocked_stats_stored_zeros_follow_sparse_branch(sb, oracle_lib):
    import scipy.sparse as sp
    x = make_matrix(50, 120, seed=5)
    m = as_csc(x)
    # explicit zeros stay stored: the reference's sparse branch runs them through the update
    data = m.data.copy()
    data[::7] = 0.0
    m2 = sp.csc_matrix((data, m.indices, m.indptr), shape=m.shape)
    block = (np.arange(120) % 3).astype(np.int32)
    mean, var = sb.compute_blocked_stats_none(m2, block, 3)
    rmean, rvar = oracle_lib.blocked_stats(m2, block, 3)
    assert np.array_equal(mean, rmean, equal_nan=True)
    assert np.array_equal(var, rvar, equal_nan=True)


# ---------------------------------------------------------------------------------------------
# overlap_exprs_paired / overlap_exprs  (tests/testthat/test-pairwise-wilcox.R:4-60, 160-175)
# ---------------------------------------------------------------------------------------------

@pytest.mark.parametrize("lfc", [0.0, 0.5, -1.0, 1.0, 1.7])
@pytest.mark.parametrize("sparse", [True, False])
def test_overlap_paired_exact(sb, oracle_lib, port_lib, lfc, sparse):
    x = make_matrix(200, 311, seed=21)
    rng = np.random.default_rng(9)
    g = rng.integers(0, 6, 311)
    groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(5)]  # group 5: cells in no group
    groups.append(np.zeros(0, np.int32))                                    # an empty group
    left, right = all_pairs(6)
    m = as_csc(x) if sparse else x
    u, t = sb.overlap_exprs_paired(m, left, right, groups, lfc, want_ties=True)
    ru = oracle_lib.overlap_paired(m, left, right, groups, lfc)
    assert np.array_equal(u, ru)
    pu, pt = port_lib.overlap_paired(m, left, right, groups, lfc, want_ties=True)
    keep = np.array([len(groups[a - 1]) > 0 and len(groups[b - 1]) > 0 for a, b in zip(left, right)])
    assert np.array_equal(t[:, keep], pt[:, keep])


def test_overlap_paired_cells_in_several_groups(sb, oracle_lib, port_lib):
    """The reference takes any lists of cell indices (src/overlap_exprs.cpp:13-32): overlapping groups, a cell listed
    twice by one group (it counts twice), groups in any order of cells."""
    x = make_matrix(150, 200, seed=23)
    rng = np.random.default_rng(10)
    groups = [np.sort(rng.choice(200, n, replace=False)).astype(np.int32) for n in (60, 90, 40, 200)]
    groups[2] = rng.permutation(np.concatenate([groups[2], groups[2][:7]])).astype(np.int32)  # repeats, unsorted
    left, right = all_pairs(4)
    for lfc in (0.0, 0.4):
        u, t = sb.overlap_exprs_paired(as_csc(x), left, right, groups, lfc, want_ties=True)
        assert np.array_equal(u, oracle_lib.overlap_paired(as_csc(x), left, right, groups, lfc))
        pu, pt = port_lib.overlap_paired(as_csc(x), left, right, groups, lfc, want_ties=True)
        assert np.array_equal(t, pt)
        gu, gt = sb.overlap_exprs(as_csc(x), groups, lfc)
        ru, rt = oracle_lib.overlap(as_csc(x), groups, lfc)
        assert np.array_equal(gu, ru, equal_nan=True) and np.array_equal(gt, rt, equal_nan=True)


def test_overlap_paired_matches_scipy(sb):
    import scipy.stats as st
    x = make_matrix(60, 150, seed=2)
    g = np.arange(150) % 3
    groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(3)]
    u = sb.overlap_exprs_paired(as_csc(x), [1, 3], [2, 1], groups)
    for k, (a, b) in enumerate([(0, 1), (2, 0)]):
        ref = [st.mannwhitneyu(x[r, groups[a]], x[r, groups[b]]).statistic for r in range(60)]
        assert np.array_equal(u[:, k], np.array(ref))


@pytest.mark.parametrize("lfc", [0.0, 0.7])
def test_overlap_exprs_exact(sb, oracle_lib, lfc):
    x = make_matrix(120, 203, seed=33)
    g = np.random.default_rng(4).integers(0, 4, 203)
    groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(4)]
    groups.insert(2, np.zeros(0, np.int32))
    U, T = sb.overlap_exprs(as_csc(x), groups, lfc)
    rU, rT = oracle_lib.overlap(as_csc(x), groups, lfc)
    assert np.array_equal(U, rU, equal_nan=True)
    assert np.array_equal(T, rT, equal_nan=True)


def test_overlap_out_of_range_message(sb):
    x = make_matrix(10, 20, seed=1)
    with pytest.raises(sb.ScranB200Error, match="indices in 'groups' out of range"):
        sb.overlap_exprs_paired(x, [1], [2], [np.array([0, 1]), np.array([2, 20])])


def test_overlap_large_units(sb, port_lib):
    """Units beyond the shared-memory sort capacity (global-memory sort) and beyond 65536
    entries (64-bit count table)."""
    rng = np.random.default_rng(77)
    ncells = 150000
    x = np.zeros((5, ncells))
    x[0] = np.where(rng.random(ncells) < 0.6, np.log2(1 + rng.poisson(3, ncells) + rng.random(ncells)), 0)
    x[1] = np.where(rng.random(ncells) < 0.2, rng.poisson(2, ncells), 0)          # ties, > cap
    x[2] = np.where(rng.random(ncells) < 0.03, rng.random(ncells), 0)             # fits shared memory
    x[3] = 1.0 + 1e-13 * rng.integers(0, 1000, ncells)                             # equal high words
    # nearly everything inside one coarse value bucket (two outliers stretch the range): the radix
    # fallback; on a grid of 2^-20 so that with lfc = 2^-10 values of the two roles coincide exactly
    x[4] = np.where(rng.random(ncells) < 0.3, 1.0 + rng.integers(0, 4096, ncells) * 2.0 ** -20, 0)
    x[4, 5] = 1000.0
    x[4, 6] = 1e-3
    g = rng.integers(0, 3, ncells)
    groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(3)]
    left, right = all_pairs(3)
    for lfc in (0.0, 0.25):
        u, t = sb.overlap_exprs_paired(as_csc(x), left, right, groups, lfc, want_ties=True)
        pu, pt = port_lib.overlap_paired(as_csc(x), left, right, groups, lfc, want_ties=True)
        assert np.array_equal(u, pu)
        assert np.array_equal(t, pt)


def test_overlap_many_groups_large_units(sb, port_lib):
    """200 groups (config 5's count): the count table nearly fills shared memory, so the labels are
    covered by narrow tiles, and with 64-bit counts (units beyond 65536 entries) in batches."""
    rng = np.random.default_rng(79)
    ncells, ngroups = 140000, 200
    x = np.zeros((3, ncells))
    x[0] = rng.random(ncells) + 0.1                                                   # every cell: 140,001 entries
    x[1] = np.where(rng.random(ncells) < 0.07, np.log2(1 + rng.poisson(2, ncells) + rng.random(ncells)), 0)
    x[2] = np.where(rng.random(ncells) < 0.6, rng.poisson(2, ncells), 0)              # long tie runs
    g = rng.integers(0, ngroups, ncells)
    groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(ngroups)]
    left, right = all_pairs(ngroups)
    u, t = sb.overlap_exprs_paired(as_csc(x), left, right, groups, 0.0, want_ties=True)
    pu, pt = port_lib.overlap_paired(as_csc(x), left, right, groups, 0.0, want_ties=True)
    assert np.array_equal(u, pu)
    assert np.array_equal(t, pt)


def test_overlap_medium_units(sb, port_lib):
    """Units of 18,000 - 65,535 entries: beyond the shared-memory sort, within 16-bit positions, i.e. the
    path that scatters into bucket order in global scratch and ranks window by window. Continuous values,
    a skewed distribution (large buckets that are not uniform), integer values (large uniform buckets),
    negative values, stored zeros; lfc != 0 doubles every entry (two roles)."""
    import scipy.sparse as sp
    rng = np.random.default_rng(83)
    ncells = 64000
    x = np.zeros((9, ncells))
    x[0] = np.where(rng.random(ncells) < 0.45, np.log2(1 + rng.poisson(1.2, ncells) + rng.random(ncells)), 0)
    x[1] = np.where(rng.random(ncells) < 0.99, rng.random(ncells) + 0.25, 0)           # ~63,400 entries
    x[2] = np.where(rng.random(ncells) < 0.6, rng.poisson(2, ncells), 0)                # long tie runs
    x[3] = np.where(rng.random(ncells) < 0.8, np.exp(8 * rng.random(ncells)), 0)        # heavy skew: a few huge buckets
    x[3, :40] = 1e6
    x[4] = np.where(rng.random(ncells) < 0.5, rng.normal(size=ncells), 0)               # both sides of zero
    x[5] = np.where(rng.random(ncells) < 0.7, np.round(rng.normal(size=ncells), 2), 0)  # thousands of short runs
    x[6] = np.where(rng.random(ncells) < 0.35, 1.0 + 1e-9 * rng.integers(0, 50, ncells), 0)  # 50 values a hair apart
    x[7] = np.where(rng.random(ncells) < 0.4, rng.random(ncells), 0)
    x[8] = np.where(rng.random(ncells) < 0.3, rng.random(ncells) * 100, 0)
    m = as_csc(x)
    data = m.data.copy()
    rows = np.repeat(np.arange(m.shape[1]), np.diff(m.indptr))  # (column of each entry; unused, keeps the intent clear)
    data[(m.indices == 7) & (rng.random(len(data)) < 0.1)] = 0.0  # stored zeros in gene 7
    m = sp.csc_matrix((data, m.indices, m.indptr), shape=m.shape)
    for ngroups in (7, 40):
        g = rng.integers(0, ngroups, ncells)
        groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(ngroups)]
        left, right = all_pairs(ngroups)
        for lfc in (0.0, 0.5):
            u, t = sb.overlap_exprs_paired(m, left, right, groups, lfc, want_ties=True)
            pu, pt = port_lib.overlap_paired(m, left, right, groups, lfc, want_ties=True)
            assert np.array_equal(u, pu), (ngroups, lfc, np.argwhere(u != pu)[:5])
            assert np.array_equal(t, pt), (ngroups, lfc)


def test_overlap_long_tie_runs(sb, port_lib):
    """Integer-valued data: runs of equal values much longer than a warp's block of 32 entries, in
    units small enough for the warp-group walk, with 3 and with 40 groups (one and two labels a lane)."""
    rng = np.random.default_rng(78)
    ncells = 9000
    x = np.zeros((5, ncells))
    x[0] = rng.poisson(1.5, ncells)
    x[1] = np.where(rng.random(ncells) < 0.5, rng.poisson(4, ncells), 0)
    x[2] = np.where(rng.random(ncells) < 0.3, 2.0, 0)                       # one run
    x[3] = np.where(rng.random(ncells) < 0.5, -rng.poisson(1, ncells), rng.poisson(1, ncells))  # both sides of zero
    x[4] = np.where(rng.random(ncells) < 0.6, np.round(rng.normal(size=ncells), 1), 0)  # short and long runs mixed
    for ngroups in (3, 40):
        g = rng.integers(0, ngroups, ncells)
        groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(ngroups)]
        left, right = all_pairs(ngroups)
        for lfc in (0.0, 1.0):
            u, t = sb.overlap_exprs_paired(as_csc(x), left, right, groups, lfc, want_ties=True)
            pu, pt = port_lib.overlap_paired(as_csc(x), left, right, groups, lfc, want_ties=True)
            assert np.array_equal(u, pu)
            assert np.array_equal(t, pt)


# ---------------------------------------------------------------------------------------------
# computeMinRank
# ---------------------------------------------------------------------------------------------

def test_compute_min_rank(sb):
    from oracle import rlevel
    rng = np.random.default_rng(8)
    x = rng.normal(size=(500, 7))
    x[::5, 2] = np.nan
    x[10:40, 3] = 0.25          # ties
    x[3, :] = np.nan            # a row with no value
    got = sb.compute_min_rank(x)
    want = rlevel.compute_min_rank(x)
    assert np.array_equal(got.astype(np.int64), want)


def test_compute_min_rank_many_rows(sb):
    from oracle import rlevel
    rng = np.random.default_rng(18)
    x = np.round(rng.normal(size=(40000, 3)), 2)   # beyond the shared-memory sort, many ties
    assert np.array_equal(sb.compute_min_rank(x).astype(np.int64), rlevel.compute_min_rank(x))


def test_compute_min_rank_degenerate_columns(sb):
    """Columns that stress the bucket ranking: most genes sharing one value (one bucket larger than
    shared memory), infinities and NaN, values clustered far beyond the bucket resolution with two
    outliers stretching the range, a constant column."""
    from oracle import rlevel
    rng = np.random.default_rng(19)
    n = 40000
    x = np.empty((n, 5))
    x[:, 0] = np.where(rng.random(n) < 0.8, 0.5, rng.random(n))            # AUC-like: 80 % exactly 0.5
    x[:, 1] = rng.normal(size=n)
    x[::7, 1] = np.inf
    x[::11, 1] = -np.inf
    x[::13, 1] = np.nan
    x[:, 2] = 1.0 + 1e-13 * rng.permutation(n)                              # distinct, tightly clustered ...
    x[0, 2] = 1e6                                                           # ... and two outliers
    x[1, 2] = -1e6
    x[:, 3] = 3.25
    x[:, 4] = np.round(rng.normal(size=n), 1)
    assert np.array_equal(sb.compute_min_rank(x).astype(np.int64), rlevel.compute_min_rank(x))
    for c in range(5):
        assert np.array_equal(sb.compute_min_rank(x[:, c:c + 1]).astype(np.int64), rlevel.compute_min_rank(x[:, c:c + 1]))


def test_compute_min_rank_long_columns(sb):
    """More rows than 16-bit positions hold: the sorting kernel."""
    from oracle import rlevel
    rng = np.random.default_rng(20)
    x = np.round(rng.normal(size=(70000, 2)), 3)
    assert np.array_equal(sb.compute_min_rank(x).astype(np.int64), rlevel.compute_min_rank(x))


# ---------------------------------------------------------------------------------------------
# scoreMarkers  (tests/testthat/test-score-markers.R)
# ---------------------------------------------------------------------------------------------

def _score_both(sb, x, groups, **kw):
    from oracle import native, rlevel
    got = sb.score_markers(as_csc(x), groups, **kw)
    want = rlevel.score_markers(as_csc(x), groups, lib=native.load("best"), **kw)
    return got, want


@pytest.mark.parametrize("full", [False, True])
def test_score_markers_unblocked(sb, full):
    x = make_matrix(400, 300, seed=101)
    groups = np.random.default_rng(1).integers(0, 5, 300)
    got, want = _score_both(sb, x, groups, full_stats=full)
    assert_tables_match(got, want)


def test_score_markers_dense_equals_sparse(sb):
    x = make_matrix(150, 200, seed=102)
    groups = np.arange(200) % 4
    a = sb.score_markers(x, groups)
    b = sb.score_markers(as_csc(x), groups)
    for g in a:
        for c in a[g]:
            assert np.array_equal(a[g][c], b[g][c], equal_nan=True)


def test_score_markers_character_groups_and_lfc(sb):
    x = make_matrix(250, 260, seed=103)
    labels = np.array(["b", "a", "d", "c"])[np.random.default_rng(2).integers(0, 4, 260)]
    got, want = _score_both(sb, x, list(labels), lfc=0.5, full_stats=True)
    assert_tables_match(got, want)


def test_score_markers_blocked(sb):
    x = make_matrix(300, 400, seed=104)
    rng = np.random.default_rng(3)
    groups = rng.integers(0, 4, 400)
    block = rng.integers(0, 3, 400)
    got, want = _score_both(sb, x, groups, block=block, full_stats=True)
    assert_tables_match(got, want)   # self. / other. included: against the restatement of scuttle::correctGroupSummary (parity unpinned)


def test_score_markers_blocking_pins_of_the_reference_tests(sb):
    """What the reference's own tests pin about blocking (tests/testthat/test-score-markers.R:216-256): the blocked
    effect of a pair is the weighted mean of the per-block effects, weights = the products of the group sizes in the
    block (from the full.* matrices of scoring every block alone), and self.average + (G - 1) other.average is the
    same for every group while self.average is not."""
    rng = np.random.default_rng(31)
    y = rng.poisson(0.5, (10, 100)).astype(np.float64)
    cluster = rng.integers(0, 5, 100)
    block = rng.integers(0, 3, 100)
    out = sb.score_markers(y, cluster, block=block, full_stats=True)
    per = [sb.score_markers(y[:, block == b], cluster[block == b], full_stats=True) for b in range(3)]
    for i in range(5):
        for eff in ("logFC.cohen", "AUC", "logFC.detected"):
            rights = list(out[i]["full." + eff + ".right"])
            for j in rights:
                v = np.stack([p[i]["full." + eff][:, list(p[i]["full." + eff + ".right"]).index(j)] for p in per], 1)
                w = np.array([p[i]["full." + eff + ".weight"][list(p[i]["full." + eff + ".right"]).index(j)] for p in per])
                wm = (~np.isnan(v)) * w[None, :]
                with np.errstate(invalid="ignore"):
                    want = np.nansum(v * wm, 1) / wm.sum(1)
                got = out[i]["full." + eff][:, rights.index(j)]
                np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-12, equal_nan=True)
    summed = [out[i]["self.average"] + out[i]["other.average"] * 4 for i in range(5)]
    for i in range(1, 5):
        np.testing.assert_allclose(summed[i], summed[0], rtol=1e-10, atol=1e-12)
        assert not np.allclose(out[i]["self.average"], out[0]["self.average"])


def test_score_markers_group_specific_batches_stay_finite(sb):
    """tests/testthat/test-score-markers.R:259-291, 296-343: groups that live in one batch only. Pairs that never share
    a batch have NA effects; the self. / other. columns stay finite."""
    rng = np.random.default_rng(32)
    y = rng.poisson(0.5, (10, 100)).astype(np.float64)
    cluster = rng.integers(10, 15, 100)
    block = np.where(cluster % 2 == 1, "odd", "even")
    out = sb.score_markers(y, cluster, block=block, full_stats=True)
    for i in range(10, 15):
        rights = list(out[i]["full.logFC.detected.right"])
        for j in rights:
            col = out[i]["full.logFC.detected"][:, rights.index(j)]
            if j % 2 != i % 2:
                assert np.isnan(col).all()
            else:
                alone = sb.score_markers(y[:, cluster % 2 == i % 2], cluster[cluster % 2 == i % 2], full_stats=True)
                ref = alone[i]["full.logFC.detected"][:, list(alone[i]["full.logFC.detected.right"]).index(j)]
                np.testing.assert_allclose(col, ref, rtol=1e-12, atol=1e-12)
        for c in sb.COLUMNS[:4]:
            assert np.isfinite(out[i][c]).all(), (i, c)
    block2 = rng.integers(1, 3, 100)
    block2[cluster == 10] = 1
    block2[cluster == 14] = 2
    out = sb.score_markers(y, cluster, block=block2, full_stats=True)
    for i in range(10, 15):
        for c in sb.COLUMNS[:4]:
            assert np.isfinite(out[i][c]).all(), (i, c)
    rights = list(out[10]["full.AUC.right"])
    assert np.isnan(out[10]["full.AUC"][:, rights.index(14)]).all()


def test_score_markers_blocked_lfc_missing_combinations(sb):
    """Group-specific batches: some pairs never share a block -> NaN columns
    (test-score-markers.R:259-344)."""
    x = make_matrix(200, 360, seed=105)
    rng = np.random.default_rng(4)
    groups = rng.integers(0, 4, 360)
    block = rng.integers(0, 3, 360)
    block[groups == 3] = 2          # group 3 only in block 2
    block[(groups == 0) & (block == 2)] = 0   # group 0 never in block 2
    got, want = _score_both(sb, x, groups, block=block, lfc=0.3, full_stats=True)
    assert_tables_match(got, want)
    assert np.isnan(got[0]["full.AUC"][:, list(got[0]["full.AUC.right"]).index(3)]).all()


def test_score_markers_pairings(sb):
    x = make_matrix(180, 240, seed=106)
    groups = np.random.default_rng(5).integers(0, 5, 240)
    for pairings in ([1, 3, 4], ([0, 1], [2, 3, 4]), np.array([[0, 1], [2, 1], [4, 0]])):
        got, want = _score_both(sb, x, groups, pairings=pairings, full_stats=True)
        assert_tables_match(got, want)


def test_score_markers_many_groups(sb):
    """More labels than one warp of count lanes (G + 1 > 32)."""
    x = make_matrix(120, 2000, seed=107, rate=1.0)
    groups = np.random.default_rng(6).integers(0, 40, 2000)
    got, want = _score_both(sb, x, groups)
    assert_tables_match(got, want)


def test_score_markers_seventy_groups(sb):
    """More than 64 comparisons per group: the medians come from a bisection on the value keys; also
    even / odd counts of defined comparisons (a group of one cell makes Cohen's d undefined)."""
    x = make_matrix(60, 2100, seed=108, rate=0.8)
    x[3] = 0.0                      # every effect of this gene ties
    groups = np.random.default_rng(7).integers(0, 70, 2100)
    groups[groups == 69] = 68       # 69 groups: 68 comparisons each
    groups[0] = 69                  # ... and one group with a single cell
    got, want = _score_both(sb, x, groups)
    assert_tables_match(got, want)


def test_score_markers_one_cell_groups(sb):
    """A group with a single cell: variance NA, so Cohen's d is NA unless the difference is 0
    (R/scoreMarkers.R:417-426); AUC and logFC.detected stay defined."""
    x = make_matrix(100, 90, seed=108)
    groups = np.arange(90) % 3
    groups[0] = 3
    got, want = _score_both(sb, x, groups, full_stats=True)
    assert_tables_match(got, want)


@pytest.mark.parametrize("case", ["one_cell", "seventy", "blocked_missing", "two_hundred"])
def test_score_markers_both_summary_kernels(sb, case, monkeypatch):
    """The summaries have two kernels (a warp per gene tile and table; a CTA per gene with all its comparison
    values in shared memory, taken for tables of 65 - 256 comparisons). Both are run on the same inputs - NA
    effects, infinite effects, heavy ties, even and odd counts - and must agree with each other (order
    statistics exactly, means to the last bits) and with the oracle."""
    rng = np.random.default_rng(31)
    kw = {}
    if case == "one_cell":
        x = make_matrix(100, 90, seed=108)
        groups = np.arange(90) % 3
        groups[0] = 3
    elif case == "seventy":
        x = make_matrix(60, 2100, seed=108, rate=0.8)
        x[3] = 0.0
        groups = np.random.default_rng(7).integers(0, 70, 2100)
        groups[groups == 69] = 68
        groups[0] = 69
    elif case == "blocked_missing":
        x = make_matrix(80, 600, seed=109)
        x[5] = 1.0                     # zero variance everywhere: infinite or undefined Cohen's d
        groups = rng.integers(0, 6, 600)
        kw["block"] = rng.integers(0, 3, 600)
        groups[kw["block"] == 2] = np.minimum(groups[kw["block"] == 2], 3)  # groups 4, 5 are absent from batch 2
    else:
        x = make_matrix(24, 3000, seed=110, rate=0.6)
        x[2] = 0.0
        groups = rng.integers(0, 200, 3000)   # 199 comparisons per table
    res = {}
    for mode in ("tile", "gene"):
        monkeypatch.setenv("SCRANB200_SUMMARIES", mode)
        res[mode] = sb.score_markers(as_csc(x), groups, **kw)
    monkeypatch.delenv("SCRANB200_SUMMARIES")
    for g in res["tile"]:
        for c in res["tile"][g]:
            a, b = np.asarray(res["tile"][g][c], np.float64), np.asarray(res["gene"][g][c], np.float64)
            if c.startswith("mean."):
                assert np.allclose(a, b, rtol=1e-14, atol=0.0, equal_nan=True), (g, c)
            else:
                assert np.array_equal(a, b, equal_nan=True), (g, c)
    from oracle import native, rlevel
    want = rlevel.score_markers(as_csc(x), groups, lib=native.load("best"), **kw)
    assert_tables_match(res["gene"], want)


# ---------------------------------------------------------------------------------------------
# pairwiseWilcox / pairwiseTTests (tests/testthat/test-pairwise-wilcox.R, test-pairwise-t.R): device
# U / tie term / means / variances under the host p-value layer, against the oracle's restatement
# ---------------------------------------------------------------------------------------------

def _pair_lists_match(got, want, rtol=1e-10):
    assert [(q["first"], q["second"]) for q in got] == [(q["first"], q["second"]) for q in want]
    for a, b in zip(got, want):
        for key in ("effect", "log_p", "log_fdr"):
            x, y = np.asarray(a[key], np.float64), np.asarray(b[key], np.float64)
            ok = np.isclose(x, y, rtol=rtol, atol=1e-12, equal_nan=True) | (x == y)
            assert ok.all(), (a["first"], a["second"], key, np.nanmax(np.abs(x - y)))


@pytest.mark.parametrize("direction", ["any", "up", "down"])
@pytest.mark.parametrize("lfc", [0.0, 0.5])
def test_pairwise_wilcox_layers(sb, port_lib, direction, lfc):
    from oracle import rlevel
    rng = np.random.default_rng(41)
    x = make_matrix(70, 420, seed=141, rate=1.2)
    groups = rng.integers(0, 4, 420)
    block = rng.integers(0, 3, 420)
    for blk in (None, block):
        want = rlevel.pairwise_wilcox(as_csc(x), groups, block=blk, lfc=lfc, direction=direction, lib=port_lib)
        got = sb.pairwise_wilcox(as_csc(x), groups, block=blk, lfc=lfc, direction=direction)
        _pair_lists_match(got, want)


def test_pairwise_wilcox_extreme_tails_and_empty_combinations(sb, port_lib):
    """Perfectly separated groups of thousands of cells (log p-values of -1e3 and below, where exp()
    underflows: asymptotic tail and the Newton inverse), a gene without any variation, and a block
    level in which one group is absent."""
    from oracle import rlevel
    rng = np.random.default_rng(43)
    ncells = 6000
    groups = rng.integers(0, 3, ncells)
    block = rng.integers(0, 2, ncells)
    block[groups == 2] = 0                      # group 2 is absent from block 1
    x = np.zeros((6, ncells))
    x[0] = groups * 10.0 + rng.random(ncells)   # perfectly separated
    x[1] = np.where(groups == 1, rng.random(ncells) + 5, 0.0)
    x[2] = 1.0                                  # no variation: every value ties
    x[3] = rng.poisson(0.3, ncells)
    x[4] = rng.normal(size=ncells)
    for blk in (None, block):
        want = rlevel.pairwise_wilcox(as_csc(x), groups, block=blk, lib=port_lib)
        got = sb.pairwise_wilcox(as_csc(x), groups, block=blk)
        _pair_lists_match(got, want)
        assert min(np.nanmin(q["log_p"]) for q in got) < -1000


@pytest.mark.parametrize("direction", ["any", "up", "down"])
@pytest.mark.parametrize("lfc", [0.0, 0.5])
def test_pairwise_t_tests_layers(sb, port_lib, direction, lfc):
    from oracle import rlevel
    rng = np.random.default_rng(42)
    x = make_matrix(70, 420, seed=142, rate=1.2)
    groups = rng.integers(0, 4, 420)
    block = rng.integers(0, 3, 420)
    for blk in (None, block):
        for std in (False, True):
            got = sb.pairwise_t_tests(as_csc(x), groups, block=blk, lfc=lfc, direction=direction, std_lfc=std)
            want = rlevel.pairwise_ttests(as_csc(x), groups, block=blk, lfc=lfc, direction=direction, std_lfc=std, lib=port_lib)
            _pair_lists_match(got, want)


def test_pairwise_t_tests_equal_welch_t_test(sb):
    """The reference's own check (tests/testthat/test-pairwise-t.R:52-95): without blocking, every pair's
    p-value is stats::t.test's (Welch), here scipy.stats.ttest_ind(equal_var=False); one-sided directions too."""
    import scipy.stats as st
    rng = np.random.default_rng(44)
    x = make_matrix(60, 300, seed=144, rate=1.5, integer_rows=False)
    x[1, ::3] = 2.5                                   # the all-equal row gets some variation
    groups = rng.integers(0, 3, 300)
    for direction, alt in (("any", "two-sided"), ("up", "greater"), ("down", "less")):
        got = sb.pairwise_t_tests(x, groups, direction=direction)
        for q in got:
            a, b = x[:, groups == q["first"]], x[:, groups == q["second"]]
            ref = st.ttest_ind(a, b, axis=1, equal_var=False, alternative=alt)
            ok = np.isfinite(ref.pvalue) & (a.var(1) > 1e-8) & (b.var(1) > 1e-8)  # (the reference floors tiny variances)
            np.testing.assert_allclose(q["log_p"][ok], np.log(ref.pvalue[ok]), rtol=1e-9, atol=1e-12)
            np.testing.assert_allclose(q["effect"], a.mean(1) - b.mean(1), rtol=1e-12, atol=1e-13)


def test_pairwise_t_tests_extreme_tails(sb, port_lib):
    """Thousands of cells per group and well separated means: log p-values far below the smallest double
    (the continued fraction of the incomplete beta function in log space), large Welch degrees of freedom,
    a group with a single cell in one block (that block is not a valid test), a constant gene."""
    from oracle import rlevel
    rng = np.random.default_rng(45)
    ncells = 9000
    groups = rng.integers(0, 3, ncells)
    block = rng.integers(0, 2, ncells)
    one = np.flatnonzero((groups == 2) & (block == 1))
    block[one[1:]] = 0                               # group 2 keeps ONE cell in block 1
    x = np.zeros((6, ncells))
    x[0] = groups * 4.0 + rng.normal(size=ncells)    # |t| ~ 150
    x[1] = groups * 0.05 + rng.normal(size=ncells)
    x[2] = 1.0
    x[3] = rng.poisson(0.3, ncells)
    x[4] = groups * 30.0 + rng.random(ncells)        # |t| in the thousands
    x[5] = np.where(groups == 1, 1e-3 * rng.random(ncells), 0.0)
    for blk in (None, block):
        for lfc in (0.0, 1.0):
            got = sb.pairwise_t_tests(x, groups, block=blk, lfc=lfc)
            want = rlevel.pairwise_ttests(x, groups, block=blk, lfc=lfc, lib=port_lib)
            _pair_lists_match(got, want)
    assert min(np.nanmin(q["log_p"]) for q in got) < -1000
