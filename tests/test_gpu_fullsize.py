"""Parity at BASELINE.json's sizes, where the CPU oracle cannot score a whole matrix in reasonable
time: identities that do not depend on the size, and the oracle on a few
genes over ALL cells of config 3. The matrices are the bench's synthetic ones, generated on the
device (scranb200_synth_csc)."""
import numpy as np
import pytest

from helpers import assert_tables_match

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sb():
    import scran_b200
    return scran_b200


def _labels(ncells, G, B, seed):
    rng = np.random.default_rng(seed)
    groups = rng.integers(0, G, ncells).astype(np.int32)
    block = rng.integers(0, B, ncells).astype(np.int32)
    groups[:G * B] = np.repeat(np.arange(G), B)  # every combination non-empty
    block[:G * B] = np.tile(np.arange(B), G)
    return groups, block


def test_u_of_both_orders_adds_up_at_config2_size(sb):
    """U(a, b; lfc) + U(b, a; -lfc) = n_a n_b for every gene and pair (ties count half on both sides),
    and the tie term does not depend on the order: 33,538 genes x 100,000 cells, 20 groups."""
    ctx = sb.Context()
    ngenes, ncells, G = 33538, 100000, 20
    groups, _ = _labels(ncells, G, 1, 2)
    m = sb.DeviceMatrix.synthetic(ctx, ngenes, 0, ngenes, ncells, groups, G, 0.07, 1002)
    cells = [np.flatnonzero(groups == g).astype(np.int32) for g in range(G)]
    sizes = np.array([len(c) for c in cells], np.float64)
    left, right = [], []
    for a in range(G):
        for b in range(G):
            if a != b:
                left.append(a + 1)
                right.append(b + 1)
    left, right = np.array(left, np.int32), np.array(right, np.int32)
    prod = sizes[left - 1] * sizes[right - 1]
    index = {(a, b): k for k, (a, b) in enumerate(zip(left.tolist(), right.tolist()))}
    mirror = np.array([index[(b, a)] for a, b in zip(left.tolist(), right.tolist())])
    u0, t0 = sb.overlap_exprs_paired(m, left, right, cells, 0.0, want_ties=True)
    assert u0.shape == (ngenes, len(left))
    assert np.array_equal(u0 + u0[:, mirror], np.broadcast_to(prod, u0.shape))
    assert np.array_equal(t0, t0[:, mirror])
    assert np.array_equal(2 * u0, np.round(2 * u0)) and (u0 >= 0).all()
    up = sb.overlap_exprs_paired(m, left, right, cells, 0.5)
    um = sb.overlap_exprs_paired(m, left, right, cells, -0.5)
    assert np.array_equal(up + um[:, mirror], np.broadcast_to(prod, up.shape))
    # an all-zero gene: every value ties, U = n_a n_b / 2
    mean, var = sb.compute_blocked_stats_none(m, np.zeros(ncells, np.int32), 1)
    zero_genes = np.flatnonzero((mean[:, 0] == 0) & (var[:, 0] == 0))
    assert len(zero_genes) > 0
    assert np.array_equal(u0[zero_genes], np.broadcast_to(prod / 2, (len(zero_genes), len(left))))
    m.free()


def test_gene_block_scores_equal_whole_matrix_scores_at_config2_size(sb):
    """Everything but the min-rank is a per-gene quantity (the multi-GPU path relies on it): scoring
    the row block [g0, g1) alone gives bit-identical columns. 33,538 x 100,000, 20 groups, 4 batches."""
    ctx = sb.Context()
    ngenes, ncells, G, B = 33538, 100000, 20, 4
    groups, block = _labels(ncells, G, B, 3)
    g0, g1 = 12000, 13500
    whole = sb.DeviceMatrix.synthetic(ctx, ngenes, 0, ngenes, ncells, groups, G, 0.07, 1002)
    res_whole = sb.score_markers(whole, groups, block=block)
    whole.free()
    part = sb.DeviceMatrix.synthetic(ctx, ngenes, g0, g1, ncells, groups, G, 0.07, 1002)
    res_part = sb.score_markers(part, groups, block=block)
    part.free()
    for g in res_whole:
        for c in res_whole[g]:
            if c.startswith("rank."):
                continue
            a, b = res_whole[g][c][g0:g1], res_part[g][c]
            assert np.array_equal(a, b, equal_nan=True), (g, c)
    # the summaries are consistent with each other on every gene
    for g in res_whole:
        for e in sb.EFFECTS:
            lo, med, hi = (res_whole[g]["min." + e], res_whole[g]["median." + e], res_whole[g]["max." + e])
            ok = np.isnan(lo) | ((lo <= med) & (med <= hi))
            assert ok.all(), (g, e)
        auc_lo, auc_hi = res_whole[g]["min.AUC"], res_whole[g]["max.AUC"]
        assert (np.isnan(auc_lo) | ((auc_lo >= 0) & (auc_hi <= 1))).all()


def test_config3_cells_against_the_oracle_on_a_few_genes(sb):
    """All 1.3M cells of config 3 (50 groups, 8 batches: units beyond 65,536 entries, the shared count
    table, the split of large units), on the genes the CPU oracle can score in seconds, among them an
    always-expressed gene, an integer-valued gene and the gene with negative values."""
    from oracle import native, rlevel
    import scipy.sparse as sp
    ctx = sb.Context()
    ngenes, ncells, G, B = 33538, 1300000, 50, 8
    groups, block = _labels(ncells, G, B, 4)
    g0, g1 = 2, 14  # gene 5: integer-valued, 7: in every cell, 13: negative values
    m = sb.DeviceMatrix.synthetic(ctx, ngenes, g0, g1, ncells, groups, G, 0.07, 1003)
    p, i, x = m.download()
    host = sp.csc_matrix((x, i, p), shape=(g1 - g0, ncells))
    got = sb.score_markers(m, groups, block=block)
    m.free()
    want = rlevel.score_markers(host, groups, block=block, lib=native.load("best"))
    assert_tables_match(got, want, skip_selfother=False)
