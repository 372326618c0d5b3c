"""GPU (needs two or more devices; skipped otherwise): the in-library gene-sharded path,
scranb200_ctx_comm_init_all + scranb200_score_markers_sharded, with the contexts of ONE process and
one host thread per context, against the fused single-GPU call: every column bit-identical,
min-ranks included. Uneven row ranges, a rank without rows, result on the root only and on every
rank. (The reference's multi-worker == serial test: tests/testthat/test-pairwise-wilcox.R:287-296.)"""
import ctypes as C
import threading

import numpy as np
import pytest

from helpers import as_csc, make_matrix

pytestmark = pytest.mark.gpu


def _device_count():
    import torch
    return torch.cuda.device_count()


def _run_sharded(sb, x, groups, block, row_starts, root, lfc=0.0):
    from scran_b200._lib import check, lib, ptr
    L = lib()
    W = len(row_starts) - 1
    ctxs = [sb.Context(d) for d in range(W)]
    arr = (C.c_void_p * W)(*[c.handle for c in ctxs])
    check(L.scranb200_ctx_comm_init_all(arr, W))
    starts = np.ascontiguousarray(row_starts, np.int64)
    ngenes = x.shape[0]
    outs, errs = [None] * W, [None] * W

    def work(r):
        try:
            ctx = ctxs[r]
            plan = sb.ScorePlan(list(groups), None if block is None else list(block), None, lfc, ctx=ctx)
            g0, g1 = int(starts[r]), int(starts[r + 1])
            m = sb.DeviceMatrix(as_csc(x[g0:g1]), ctx)
            want_out = root < 0 or root == r
            out = np.full((plan.K, 19, ngenes), -1.0) if want_out else None
            rc = L.scranb200_score_markers_sharded(ctx.handle, m.handle, plan.handle, ngenes, ptr(starts), root, ptr(out), None)
            if rc != 0:
                errs[r] = L.scranb200_last_error().decode()
            outs[r] = out
            m.free()
            plan.free()
        except Exception as e:  # noqa: BLE001
            errs[r] = repr(e)

    threads = [threading.Thread(target=work, args=(r,)) for r in range(W)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert not any(t.is_alive() for t in threads), "a rank is stuck in the exchange"
    assert errs == [None] * W, errs
    for c in ctxs:
        c.close()
    return outs


@pytest.mark.parametrize("root", [0, -1])
def test_sharded_equals_fused_two_gpus(root):
    if _device_count() < 2:
        pytest.skip("needs two GPUs")
    import scran_b200 as sb
    from scran_b200._lib import check, lib, ptr
    x = make_matrix(523, 410, seed=91)
    rng = np.random.default_rng(12)
    groups = rng.integers(0, 6, 410)
    block = rng.integers(0, 3, 410)
    ctx = sb.default_context()
    plan = sb.ScorePlan(list(groups), list(block), None, 0.0, ctx=ctx)
    m = sb.DeviceMatrix(as_csc(x), ctx)
    want = np.zeros((plan.K, 19, 523))
    check(lib().scranb200_score_markers(ctx.handle, m.handle, plan.handle, ptr(want), None))
    m.free()
    plan.free()
    outs = _run_sharded(sb, x, groups, block, [0, 140, 523], root)
    assert np.array_equal(outs[0], want, equal_nan=True)
    if root < 0:
        assert np.array_equal(outs[1], want, equal_nan=True)
    else:
        assert outs[1] is None


def test_sharded_with_an_empty_rank_and_lfc():
    n = _device_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    import scran_b200 as sb
    from scran_b200._lib import check, lib, ptr
    x = make_matrix(300, 260, seed=5)
    rng = np.random.default_rng(2)
    groups = rng.integers(0, 4, 260)
    ctx = sb.default_context()
    plan = sb.ScorePlan(list(groups), None, None, 0.5, ctx=ctx)
    m = sb.DeviceMatrix(as_csc(x), ctx)
    want = np.zeros((plan.K, 19, 300))
    check(lib().scranb200_score_markers(ctx.handle, m.handle, plan.handle, ptr(want), None))
    m.free()
    plan.free()
    starts = [0, 300, 300] if n == 2 else [0, 120, 120, 300][: n + 1] if n == 3 else [0, 120, 120, 300]
    if len(starts) - 1 > n:
        starts = [0, 300, 300]
    outs = _run_sharded(sb, x, groups, None, starts, 0, lfc=0.5)
    assert np.array_equal(outs[0], want, equal_nan=True)
