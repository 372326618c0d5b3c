"""GPU: a matrix uploaded in uneven column chunks (scranb200_mat_begin / _append_columns / _finish, SURVEY.md
section 8 f4) gives bit-identical results to the same matrix uploaded as one object, at config 2's size
(33,538 x 100,000, 20 groups, 4 batches; 7 chunks, one of them empty, int32 and int64 column pointers, a
capacity hint that is too small so that the device arrays have to grow). And summaryMarkerStats."""
import numpy as np
import pytest

from helpers import as_csc, make_matrix

pytestmark = pytest.mark.gpu


def test_seven_uneven_chunks_equal_one_upload_at_config2_size():
    import scipy.sparse as sp

    import scran_b200 as sb
    ctx = sb.default_context()
    ngenes, ncells, G, B = 33538, 100000, 20, 4
    rng = np.random.default_rng(6)
    groups = rng.integers(0, G, ncells).astype(np.int32)
    block = rng.integers(0, B, ncells).astype(np.int32)
    whole = sb.DeviceMatrix.synthetic(ctx, ngenes, 0, ngenes, ncells, groups, G, 0.07, 1002)
    p, i, x = whole.download()
    want = sb.score_markers(whole, groups, block=block)
    whole.free()
    cuts = [0, 1, 17000, 17000, 40123, 40124, 99000, ncells]   # uneven, one empty chunk, single-column chunks
    chunks = []
    for k, (a, b) in enumerate(zip(cuts[:-1], cuts[1:])):
        pp = p[a:b + 1] - p[a]
        if k % 2 == 0:
            pp = pp.astype(np.int32)                            # R's own @p
        chunks.append((pp, i[p[a]:p[b]], x[p[a]:p[b]]))
    m = sb.DeviceMatrix.from_column_chunks(ngenes, ncells, chunks, nnz_hint=len(x) // 3, ctx=ctx)
    assert m.nnz == len(x)
    got = sb.score_markers(m, groups, block=block)
    m.free()
    for g in want:
        for c in want[g]:
            assert np.array_equal(np.asarray(got[g][c]), np.asarray(want[g][c]), equal_nan=True), (g, c)


def test_chunk_builder_refuses_bad_input():
    import ctypes as C

    import scran_b200 as sb
    from scran_b200._lib import ScranB200Error, lib
    x = as_csc(make_matrix(50, 40, seed=3))
    with pytest.raises(ScranB200Error, match="fewer columns"):
        sb.DeviceMatrix.from_column_chunks(50, 40, [x[:, :30]])
    with pytest.raises(ScranB200Error, match="more columns"):
        sb.DeviceMatrix.from_column_chunks(50, 40, [x[:, :30], x[:, 20:]])
    bad = x[:, :10].tocsc()
    with pytest.raises(ScranB200Error, match="start at 0"):
        sb.DeviceMatrix.from_column_chunks(50, 10, [(bad.indptr + 1, bad.indices, bad.data)])
    lib().scranb200_mat_abort(None)


def test_summary_marker_stats(oracle_lib):
    """R/summaryMarkerStats.R:62-95: per-group means / detected proportions and their mean over the other groups."""
    import scran_b200 as sb
    x = make_matrix(200, 300, seed=61)
    rng = np.random.default_rng(4)
    groups = rng.integers(0, 5, 300)
    got = sb.summary_marker_stats(as_csc(x), groups)
    means = np.stack([x[:, groups == g].mean(1) for g in range(5)], 1)
    det = np.stack([(x[:, groups == g] > 0).mean(1) for g in range(5)], 1)
    for g in range(5):
        others = [h for h in range(5) if h != g]
        np.testing.assert_allclose(got[g]["self.average"], means[:, g], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(got[g]["other.average"], means[:, others].mean(1), rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(got[g]["self.detected"], det[:, g], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(got[g]["other.detected"], det[:, others].mean(1), rtol=1e-12, atol=1e-14)
    full = sb.score_markers(as_csc(x), groups)
    for g in range(5):
        for c in sb.COLUMNS[:4]:
            assert np.array_equal(got[g][c], full[g][c])
