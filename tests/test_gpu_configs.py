"""GPU: the workloads BASELINE.json names, each against the CPU oracle at its own size - whole where the
oracle can score the matrix in seconds (config 1), on gene ranges over ALL cells otherwise (the per-gene
cost is what grows with the cell count; ranges are chosen to hold the special genes of the synthetic
matrices: an integer-valued gene, one expressed in every cell, one with negative values).

  config 1  scoreMarkers, 10,000 x 3,000, 10 groups, no block              whole matrix
  config 2  pairwiseWilcox / scoreMarkers, 33,538 x 100,000, 20 x 4         64 genes x all cells
  config 4  pairwiseTTests(lfc = 0.5) + blocked stats on the 1.3M-cell matrix   16 genes x all cells
  config 5  scoreMarkers, 20,000 x 500,000, 200 groups                      8 genes x all cells
(config 3: tests/test_gpu_fullsize.py.)"""
import numpy as np
import pytest
import scipy.sparse as sp

from helpers import assert_tables_match

pytestmark = pytest.mark.gpu


def _labels(ncells, G, B, seed):
    rng = np.random.default_rng(seed)
    groups = rng.integers(0, G, ncells).astype(np.int32)
    block = rng.integers(0, B, ncells).astype(np.int32)
    groups[:G * B] = np.repeat(np.arange(G), B)
    block[:G * B] = np.tile(np.arange(B), G)
    return groups, block


def _rows(sb, ctx, ngenes, ranges, ncells, groups, G, density, seed):
    """The genes of `ranges` of the synthetic matrix as ONE device matrix plus its host copy."""
    ps, is_, xs, nrow = [], [], [], 0
    mats = []
    for g0, g1 in ranges:
        m = sb.DeviceMatrix.synthetic(ctx, ngenes, g0, g1, ncells, groups, G, density, seed)
        p, i, x = m.download()
        m.free()
        mats.append(sp.csc_matrix((x, i, p), shape=(g1 - g0, ncells)))
    host = sp.vstack(mats).tocsc()
    host.sort_indices()
    return sb.DeviceMatrix(host, ctx), host


def test_config1_whole_matrix(oracle_lib):
    import scran_b200 as sb
    from oracle import rlevel
    ctx = sb.default_context()
    ngenes, ncells, G = 10000, 3000, 10
    groups, _ = _labels(ncells, G, 1, 1)
    m = sb.DeviceMatrix.synthetic(ctx, ngenes, 0, ngenes, ncells, groups, G, 0.63, 1001)
    p, i, x = m.download()
    host = sp.csc_matrix((x, i, p), shape=(ngenes, ncells))
    got = sb.score_markers(m, groups)
    m.free()
    want = rlevel.score_markers(host, groups, lib=oracle_lib)
    assert_tables_match(got, want)


def test_config2_gene_ranges(oracle_lib, port_lib):
    import scran_b200 as sb
    from oracle import rlevel
    ctx = sb.default_context()
    ngenes, ncells, G, B = 33538, 100000, 20, 4
    groups, block = _labels(ncells, G, B, 2)
    m, host = _rows(sb, ctx, ngenes, [(0, 24), (1000, 1016), (20000, 20024)], ncells, groups, G, 0.07, 1002)
    got = sb.score_markers(m, groups, block=block)
    want = rlevel.score_markers(host, groups, block=block, lib=oracle_lib)
    assert_tables_match(got, want)
    # pairwiseWilcox with its p-value layer (findMarkers(test.type = "wilcox") is this plus combineMarkers)
    from test_gpu_parity import _pair_lists_match
    gotw = sb.pairwise_wilcox(m, groups, block=block)
    wantw = rlevel.pairwise_wilcox(host, groups, block=block, lib=port_lib)
    _pair_lists_match(gotw, wantw)
    m.free()


def test_config4_t_tests_and_blocked_stats(oracle_lib, port_lib):
    import scran_b200 as sb
    from oracle import rlevel
    from test_gpu_parity import _pair_lists_match
    ctx = sb.default_context()
    ngenes, ncells, G, B = 33538, 1300000, 50, 8
    groups, block = _labels(ncells, G, B, 4)
    m, host = _rows(sb, ctx, ngenes, [(2, 14), (30000, 30004)], ncells, groups, G, 0.07, 1003)
    # compute_blocked_stats_none with the t-tests' 400 block-major ids (R/pairwiseTTests.R:242) and modelGeneVar's 8
    ids400 = (groups + block * G).astype(np.int32)
    for ids, n in ((ids400, G * B), (block, B)):
        mean, var = sb.compute_blocked_stats_none(m, ids, n)
        rmean, rvar = oracle_lib.blocked_stats(host, ids, n)
        assert np.array_equal(mean, rmean, equal_nan=True)
        assert np.array_equal(var, rvar, equal_nan=True)
    got = sb.pairwise_t_tests(m, groups, block=block, lfc=0.5)
    want = rlevel.pairwise_ttests(host, groups, block=block, lfc=0.5, lib=port_lib)
    _pair_lists_match(got, want)
    m.free()


def test_config5_two_hundred_groups(oracle_lib):
    import scran_b200 as sb
    from oracle import rlevel
    ctx = sb.default_context()
    ngenes, ncells, G = 20000, 500000, 200
    groups, _ = _labels(ncells, G, 1, 5)
    m, host = _rows(sb, ctx, ngenes, [(4, 9), (13, 14), (10000, 10002)], ncells, groups, G, 0.07, 1005)
    got = sb.score_markers(m, groups)
    m.free()
    want = rlevel.score_markers(host, groups, lib=oracle_lib)
    assert_tables_match(got, want)
