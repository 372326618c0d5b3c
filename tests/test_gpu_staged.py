"""GPU: the staged entry points that the gene-sharded multi-GPU path is built from
(scranb200_score_effects_dev / _score_minrank_dev / _score_finish_dev), driven on ONE GPU exactly the
way `world` ranks would drive them: every "rank" scores its own row block into the rows
[row0, row0 + n) of shared [.., ngenes] buffers (ld = ngenes, row0 != 0), the comparison columns
are ranked in `world` separate calls over [q0, q1) into one min-rank table, and the table is
written back once. The result must be bit-identical to the fused single-GPU call and agree with
the oracle (ranks exactly, the rest to 1e-12).

This mirrors the reference's own decomposition, R/scoreMarkers.R:234-256 (row blocks through
rowBlockApply, rbind, then the min-rank over ALL genes), and its multi-worker == serial test,
tests/testthat/test-pairwise-wilcox.R:287-296.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import as_csc, assert_tables_match, make_matrix

pytestmark = pytest.mark.gpu


def _cut(n, world):
    per = (n + world - 1) // world
    return [(min(n, r * per), min(n, (r + 1) * per)) for r in range(world)]


def _staged(sb, torch, x, groups, block, world, lfc=0.0, uneven=False):
    from scran_b200._lib import check, lib, ptr
    L = lib()
    ctx = sb.default_context()
    ngenes = x.shape[0]
    plan = sb.ScorePlan(list(groups), None if block is None else list(block), None, lfc, ctx=ctx)
    K, Q = plan.K, plan.Q
    dev = torch.device("cuda", ctx.device)
    eff = torch.full((3, Q, ngenes), float("nan"), dtype=torch.float64, device=dev)
    out = torch.full((K, 19, ngenes), float("nan"), dtype=torch.float64, device=dev)
    minrank = torch.empty((K, 3, ngenes), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    granges = _cut(ngenes, world)
    if uneven:  # shards of very different sizes, one of them empty
        cuts = [0, ngenes // 7, ngenes // 7, ngenes]
        granges = list(zip(cuts[:-1], cuts[1:]))
    for g0, g1 in granges:
        if g1 == g0:
            continue
        m = sb.DeviceMatrix(as_csc(x[g0:g1]), ctx)
        check(L.scranb200_score_effects_dev(ctx.handle, m.handle, plan.handle, ngenes, g0, ptr(eff), ptr(out)))
        m.free()
    check(L.scranb200_fill_i32(ctx.handle, ptr(minrank), minrank.numel(), np.iinfo(np.int32).max))
    for q0, q1 in _cut(Q, len(granges)):
        check(L.scranb200_score_minrank_dev(ctx.handle, plan.handle, ngenes, ngenes, ptr(eff), q0, q1, ptr(minrank)))
    check(L.scranb200_score_finish_dev(ctx.handle, plan.handle, ngenes, ngenes, ptr(minrank), ptr(out)))
    torch.cuda.synchronize()
    res = out.cpu().numpy()
    tables = {}
    for k in range(K):
        g = plan.universe[plan.out_groups[k]]
        tables[g] = {name: res[k, c] for c, name in enumerate(sb.COLUMNS)}
    plan.free()
    return tables


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("blocked", [False, True])
def test_staged_row_blocks_equal_the_fused_call_and_the_oracle(world, blocked, oracle_lib):
    import torch

    import scran_b200 as sb
    from oracle import rlevel
    x = make_matrix(421, 380, seed=41 + world)
    rng = np.random.default_rng(world)
    groups = rng.integers(0, 5, 380)
    block = rng.integers(0, 3, 380) if blocked else None
    fused = sb.score_markers(as_csc(x), groups, block=block)
    staged = _staged(sb, torch, x, groups, block, world)
    for g in fused:
        for c in sb.COLUMNS:
            a, b = np.asarray(fused[g][c], np.float64), staged[g][c]
            if c.startswith("rank."):
                b = np.where(np.isfinite(b), b, np.iinfo(np.int32).max)
            assert np.array_equal(a, b, equal_nan=True), (g, c)
    want = rlevel.score_markers(as_csc(x), groups, block=block, lib=oracle_lib)
    assert_tables_match(fused, want)


def test_staged_uneven_and_empty_shards_with_lfc():
    import torch

    import scran_b200 as sb
    x = make_matrix(300, 260, seed=77)
    rng = np.random.default_rng(8)
    groups = rng.integers(0, 4, 260)
    block = rng.integers(0, 2, 260)
    fused = sb.score_markers(as_csc(x), groups, block=block, lfc=0.5)
    staged = _staged(sb, torch, x, groups, block, 3, lfc=0.5, uneven=True)
    for g in fused:
        for c in sb.COLUMNS:
            a, b = np.asarray(fused[g][c], np.float64), staged[g][c]
            if c.startswith("rank."):
                b = np.where(np.isfinite(b), b, np.iinfo(np.int32).max)
            assert np.array_equal(a, b, equal_nan=True), (g, c)
