"""Host-side mirror of the reference's interface for the marker-scoring path.

Same names, argument meaning and error behaviour as the reference functions they stand for, so
that the parity tests read like the reference's own tests:

    score_markers          scoreMarkers()            R/scoreMarkers.R:173-266
    pairwise_wilcox        pairwiseWilcox()          R/pairwiseWilcox.R:150-256
    pairwise_t_tests       pairwiseTTests()          R/pairwiseTTests.R:179-309 (block= branch)
    compute_blocked_stats_none                       R/RcppExports.R:24-26
    overlap_exprs / overlap_exprs_paired             R/RcppExports.R:48-54
    compute_min_rank       computeMinRank()          R/computeMinRank.R:45-50

All numerics run in libscranb200.so on the GPU through the C ABI, the p-value layers of
pairwiseWilcox / pairwiseTTests included; this module only converts
arguments (what the R closures do before .Call) and reshapes results.
"""
import ctypes as C
from collections import OrderedDict

import numpy as np

from . import _lib
from ._lib import NA_INT, NCOLS, ScranB200Error, check, lib, ptr

EFFECTS = ("logFC.cohen", "AUC", "logFC.detected")
COLUMNS = ["self.average", "other.average", "self.detected", "other.detected"] + [
    "%s.%s" % (s, e) for e in EFFECTS for s in ("mean", "min", "median", "max", "rank")]


class Context:
    """One GPU. Created lazily by the module-level helpers; explicit for multi-GPU callers."""

    def __init__(self, device=0):
        h = C.c_void_p()
        check(lib().scranb200_ctx_create(int(device), C.byref(h)))
        self.handle = h
        self.device = int(device)

    def close(self):
        if self.handle:
            lib().scranb200_ctx_destroy(self.handle)
            self.handle = None

    def sync(self):
        check(lib().scranb200_ctx_sync(self.handle))

    @property
    def launches(self):
        return int(lib().scranb200_ctx_launches(self.handle))

    def last_timing(self):
        out = np.zeros(8, np.float64)
        n = lib().scranb200_ctx_last_timing(self.handle, ptr(out), 8)
        names = ["transpose", "stats", "sort", "count", "effects", "summaries", "minrank", "total"]
        return OrderedDict(zip(names[:n], out[:n].tolist()))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default_ctx = None


def default_context():
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context(0)
    return _default_ctx


class DeviceMatrix:
    """A genes x cells matrix resident in HBM (CSC). Accepts scipy.sparse, dense ndarray, or raw
    CSC arrays; keeps nothing on the host."""

    def __init__(self, x, ctx=None):
        self.ctx = ctx or default_context()
        h = C.c_void_p()
        try:
            import scipy.sparse as sp
        except ImportError:  # pragma: no cover
            sp = None
        if isinstance(x, tuple):  # (nrow, ncol, p, i, x) host arrays
            nrow, ncol, p, i, v = x
            p = np.ascontiguousarray(p, np.int64)
            i = np.ascontiguousarray(i, np.int32)
            v = np.ascontiguousarray(v, np.float64)
            check(lib().scranb200_upload_csc(self.ctx.handle, nrow, ncol, ptr(p), ptr(i), ptr(v), C.byref(h)))
        elif sp is not None and sp.issparse(x):
            x = x.tocsc()
            if not x.has_sorted_indices:
                x = x.sorted_indices()
            nrow, ncol = x.shape
            i = np.ascontiguousarray(x.indices, np.int32)
            v = np.ascontiguousarray(x.data, np.float64)
            if x.indptr.dtype == np.int32:  # R's own @p
                p = np.ascontiguousarray(x.indptr)
                check(lib().scranb200_upload_csc32(self.ctx.handle, nrow, ncol, ptr(p), ptr(i), ptr(v), C.byref(h)))
            else:
                p = np.ascontiguousarray(x.indptr, np.int64)
                check(lib().scranb200_upload_csc(self.ctx.handle, nrow, ncol, ptr(p), ptr(i), ptr(v), C.byref(h)))
        else:
            x = np.asarray(x, dtype=np.float64)
            if x.ndim != 2:
                raise ValueError("matrix expected")
            nrow, ncol = x.shape
            cm = np.ascontiguousarray(x.T)  # column-major storage
            check(lib().scranb200_upload_dense(self.ctx.handle, nrow, ncol, ptr(cm), C.byref(h)))
        self.handle = h
        self.shape = (int(nrow), int(ncol))

    @classmethod
    def from_column_chunks(cls, nrow, ncol, chunks, nnz_hint=0, ctx=None):
        """A matrix that arrives as CSC pieces of consecutive columns (scranb200_mat_begin / _append_columns /
        _finish): chunks = iterable of scipy.sparse matrices or (p, i, x) triples, p starting at 0 (int32 p, R's
        own, is taken as it is)."""
        self = cls.__new__(cls)
        self.ctx = ctx or default_context()
        b = C.c_void_p()
        check(lib().scranb200_mat_begin(self.ctx.handle, int(nrow), int(ncol), int(nnz_hint), C.byref(b)))
        try:
            for ch in chunks:
                if not isinstance(ch, tuple):
                    ch = ch.tocsc()
                    if not ch.has_sorted_indices:
                        ch = ch.sorted_indices()
                    ch = (ch.indptr, ch.indices, ch.data)
                p, i, v = ch
                i = np.ascontiguousarray(i, np.int32)
                v = np.ascontiguousarray(v, np.float64)
                if np.asarray(p).dtype == np.int32:
                    p = np.ascontiguousarray(p)
                    check(lib().scranb200_mat_append_columns32(b, len(p) - 1, ptr(p), ptr(i), ptr(v)))
                else:
                    p = np.ascontiguousarray(p, np.int64)
                    check(lib().scranb200_mat_append_columns(b, len(p) - 1, ptr(p), ptr(i), ptr(v)))
            h = C.c_void_p()
            check(lib().scranb200_mat_finish(b, C.byref(h)))
        except Exception:
            lib().scranb200_mat_abort(b)
            raise
        self.handle = h
        self.shape = (int(nrow), int(ncol))
        return self

    @classmethod
    def wrap_device(cls, ctx, nrow, ncol, nnz, d_p, d_i, d_x):
        self = cls.__new__(cls)
        self.ctx = ctx
        h = C.c_void_p()
        check(lib().scranb200_wrap_device_csc(ctx.handle, nrow, ncol, nnz, ptr(d_p), ptr(d_i), ptr(d_x), C.byref(h)))
        self.handle = h
        self.shape = (int(nrow), int(ncol))
        self._keep = (d_p, d_i, d_x)
        return self

    @classmethod
    def synthetic(cls, ctx, ngenes_total, g0, g1, ncells, groups, ngroups, density, seed):
        self = cls.__new__(cls)
        self.ctx = ctx
        h = C.c_void_p()
        groups = np.ascontiguousarray(groups, np.int32)
        check(lib().scranb200_synth_csc(ctx.handle, ngenes_total, g0, g1, ncells, ptr(groups), ngroups, float(density),
                                        int(seed), C.byref(h)))
        self.handle = h
        self.shape = (int(g1 - g0), int(ncells))
        return self

    @property
    def nnz(self):
        return int(lib().scranb200_mat_nnz(self.handle))

    def download(self):
        p = np.zeros(self.shape[1] + 1, np.int64)
        i = np.zeros(max(1, self.nnz), np.int32)
        x = np.zeros(max(1, self.nnz), np.float64)
        check(lib().scranb200_download_csc(self.ctx.handle, self.handle, ptr(p), ptr(i), ptr(x)))
        return p, i[:self.nnz], x[:self.nnz]

    def free(self):
        if getattr(self, "handle", None):
            lib().scranb200_mat_free(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def _as_matrix(x, ctx=None):
    return x if isinstance(x, DeviceMatrix) else DeviceMatrix(x, ctx)


# ---------------------------------------------------------------------------------------------
# The three .Call-level entry points
# ---------------------------------------------------------------------------------------------

def compute_blocked_stats_none(x, block, nblocks, want_detected=False):
    """R: compute_blocked_stats_none(mat, block, nblocks) -> list(mean, var); block 0-based,
    NA_INT = skip. Returns [ngenes, nblocks] arrays (and the detected counts on request)."""
    m = _as_matrix(x)
    block = np.ascontiguousarray(block, np.int32)
    if block.shape[0] != m.shape[1]:
        raise ValueError("length of 'block' should be the same as 'ncol(x)'")
    mean = np.zeros((nblocks, m.shape[0]), np.float64)
    var = np.zeros((nblocks, m.shape[0]), np.float64)
    det = np.zeros((nblocks, m.shape[0]), np.float64) if want_detected else None
    check(lib().scranb200_blocked_stats(m.ctx.handle, m.handle, ptr(block), int(nblocks), ptr(mean), ptr(var), ptr(det)))
    if want_detected:
        return mean.T, var.T, det.T
    return mean.T, var.T


def _blocked_stats_transformed(x, block, nblocks, transform, sf, pseudo):
    m = _as_matrix(x)
    block = np.ascontiguousarray(block, np.int32)
    sf = np.ascontiguousarray(sf, np.float64)
    if block.shape[0] != m.shape[1]:
        raise ValueError("length of 'block' should be the same as 'ncol(x)'")
    if sf.shape[0] != m.shape[1]:
        raise ValueError("length of 'sf' should be the same as 'ncol(x)'")
    mean = np.zeros((nblocks, m.shape[0]), np.float64)
    var = np.zeros((nblocks, m.shape[0]), np.float64)
    check(lib().scranb200_blocked_stats_transformed(m.ctx.handle, m.handle, ptr(block), int(nblocks), int(transform),
                                                    ptr(sf), float(pseudo), ptr(mean), ptr(var)))
    return mean.T, var.T


def compute_blocked_stats_lognorm(x, block, nblocks, sf, pseudo):
    """R: compute_blocked_stats_lognorm(mat, block, nblocks, sf, pseudo) (R/RcppExports.R:16-18):
    mean / variance per block of log2(x / sf + pseudo)."""
    return _blocked_stats_transformed(x, block, nblocks, 1, sf, pseudo)


def compute_blocked_stats_norm(x, block, nblocks, sf):
    """R: compute_blocked_stats_norm(mat, block, nblocks, sf) (R/RcppExports.R:20-22): mean /
    variance per block of x / sf."""
    return _blocked_stats_transformed(x, block, nblocks, 2, sf, 1.0)


def _flatten_groups(groups):
    gptr = np.zeros(len(groups) + 1, np.int64)
    for k, g in enumerate(groups):
        gptr[k + 1] = gptr[k] + len(g)
    cells = [np.asarray(g, np.int32).ravel() for g in groups]
    cells = np.concatenate(cells) if cells else np.zeros(0, np.int32)
    return gptr, np.ascontiguousarray(np.concatenate([cells, np.zeros(1, np.int32)]), np.int32)


def overlap_exprs_paired(x, left, right, groups, lfc=0.0, want_ties=False):
    """R: overlap_exprs_paired(exprs, left, right, groups, lfc); left/right 1-based, groups a list
    of 0-based cell index vectors. Returns U as [ngenes, npairs] (R's matrix transposed)."""
    m = _as_matrix(x)
    left = np.ascontiguousarray(left, np.int32)
    right = np.ascontiguousarray(right, np.int32)
    gptr, cells = _flatten_groups(groups)
    u = np.zeros((m.shape[0], max(1, len(left))), np.float64)
    t = np.zeros_like(u) if want_ties else None
    lp = left if left.size else np.zeros(1, np.int32)
    rp = right if right.size else np.zeros(1, np.int32)
    check(lib().scranb200_overlap_exprs_paired(m.ctx.handle, m.handle, ptr(lp), ptr(rp), len(left), len(groups),
                                               ptr(gptr), ptr(cells), float(lfc), ptr(u), ptr(t)))
    u = u[:, :len(left)]
    if want_ties:
        return u, t[:, :len(left)]
    return u


def overlap_exprs(x, groups, lfc=0.0):
    """R: overlap_exprs(exprs, groups, lfc) -> (U, ties), each [ngroups, ngroups, ngenes] with
    entry (i, j) = column j of the reference's matrix i (NaN where the reference has none)."""
    m = _as_matrix(x)
    gptr, cells = _flatten_groups(groups)
    G = len(groups)
    u = np.zeros((G, G, m.shape[0]), np.float64)
    t = np.zeros((G, G, m.shape[0]), np.float64)
    check(lib().scranb200_overlap_exprs(m.ctx.handle, m.handle, G, ptr(gptr), ptr(cells), float(lfc), ptr(u), ptr(t)))
    return u, t


def compute_min_rank(x):
    """R: computeMinRank(x) with the defaults. x: [nrow, ncol]; NaN = NA."""
    ctx = default_context()
    x = np.asarray(x, np.float64)
    cm = np.ascontiguousarray(x.T)
    out = np.zeros(x.shape[0], np.int32)
    check(lib().scranb200_compute_min_rank(ctx.handle, ptr(cm), x.shape[0], x.shape[1], ptr(out)))
    return out


# ---------------------------------------------------------------------------------------------
# scoreMarkers
# ---------------------------------------------------------------------------------------------

def expand_pairings(pairings, universe):
    """.expand_pairings (R/expandPairings.R:1-58): -> (id1, id2) 0-based indices into universe."""
    universe = list(universe)
    pos = {v: k for k, v in enumerate(universe)}

    def subset(request, clean=True):
        out = list(range(len(universe))) if request is None else [pos.get(v) for v in request]
        if clean:
            seen = []
            for v in out:
                if v is not None and v not in seen:
                    seen.append(v)
            out = seen
        return out

    def clean_expand(xs, ys):
        return [(a, b) for b in ys for a in xs if a != b]

    if isinstance(pairings, np.ndarray) and pairings.ndim == 2:
        if pairings.shape[1] != 2:
            raise ValueError("'pairings' should be a numeric/character matrix with 2 columns")
        s1 = subset(list(pairings[:, 0]), clean=False)
        s2 = subset(list(pairings[:, 1]), clean=False)
        pairs = [(a, b) for a, b in zip(s1, s2) if a is not None and b is not None]
    elif isinstance(pairings, (list, tuple)) and len(pairings) == 2 and all(
            p is None or isinstance(p, (list, tuple, np.ndarray)) for p in pairings):
        conv = [subset(None if p is None else list(p)) for p in pairings]
        pairs = clean_expand(conv[0], conv[1])
    elif isinstance(pairings, (list, tuple)) and len(pairings) != 2 and any(
            isinstance(p, (list, tuple, np.ndarray)) for p in pairings):
        raise ValueError("'pairings' as a list should have length 2")
    else:
        avail = subset(None if pairings is None else list(pairings))
        pairs = clean_expand(avail, avail)
    return [a for a, _ in pairs], [b for _, b in pairs]


class ScorePlan:
    """The bookkeeping of .scoreMarkers before the matrix is touched, held by the C++ host code."""

    def __init__(self, groups, block=None, pairings=None, lfc=0.0, ctx=None):
        self.ctx = ctx or default_context()
        groups = list(groups)
        self.universe = sorted(set(groups))
        code = {g: k for k, g in enumerate(self.universe)}
        gcodes = np.array([code[g] for g in groups], np.int32)
        bcodes = None
        if block is not None:
            block = list(block)
            if len(block) != len(groups):
                raise ValueError("length of 'block' does not equal 'ncol(x)'")
            blevels = sorted(set(block))
            bmap = {b: k for k, b in enumerate(blevels)}
            bcodes = np.array([bmap[b] for b in block], np.int32)
        id1, id2 = expand_pairings(pairings, self.universe)
        dl = np.ascontiguousarray(id1, np.int32)
        dr = np.ascontiguousarray(id2, np.int32)
        if dl.size == 0:
            dl = dr = np.zeros(1, np.int32)
            nd = 0
        else:
            nd = len(id1)
        h = C.c_void_p()
        self._keep = (gcodes, bcodes, dl, dr)
        if pairings is None:
            check(lib().scranb200_plan_create(self.ctx.handle, len(groups), ptr(gcodes), len(self.universe), ptr(bcodes),
                                              None, None, 0, float(lfc), C.byref(h)))
        else:
            if nd == 0:
                raise ValueError("no valid comparisons in 'pairings'")
            check(lib().scranb200_plan_create(self.ctx.handle, len(groups), ptr(gcodes), len(self.universe), ptr(bcodes),
                                              ptr(dl), ptr(dr), nd, float(lfc), C.byref(h)))
        self.handle = h
        self.ncells = len(groups)
        info = np.zeros(5, np.int64)
        check(lib().scranb200_plan_info(h, ptr(info), 5))
        self.K, self.ndesired, self.Q, self.npairs, self.C = [int(v) for v in info]
        og = np.zeros(max(1, self.K), np.int32)
        rp = np.zeros(self.K + 1, np.int64)
        rg = np.zeros(max(1, self.ndesired), np.int32)
        w = np.zeros(max(1, self.ndesired), np.float64)
        check(lib().scranb200_plan_tables(h, ptr(og), ptr(rp), ptr(rg), ptr(w)))
        self.out_groups = og[:self.K]
        self.right_ptr = rp
        self.right_group = rg[:self.ndesired]
        self.weight = w[:self.ndesired]

    def free(self):
        if getattr(self, "handle", None):
            lib().scranb200_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def score_markers(x, groups, block=None, pairings=None, lfc=0.0, full_stats=False, plan=None):
    """scoreMarkers(x, groups, block=, pairings=, lfc=, full.stats=). Returns an OrderedDict
    group -> OrderedDict(column -> array) with the reference's column order; rank.* are integer
    min-ranks; with full_stats the full.<effect> matrices ([ngenes, ncomparisons]) follow their
    effect's rank column, as in the reference's DataFrames."""
    m = _as_matrix(x)
    if len(groups) != m.shape[1]:
        raise ValueError("length of 'groups' does not equal 'ncol(x)'")
    own = plan is None
    plan = plan or ScorePlan(groups, block, pairings, lfc, ctx=m.ctx)
    ngenes = m.shape[0]
    out = np.zeros((max(1, plan.K), NCOLS, ngenes), np.float64)
    full = np.zeros((3, max(1, plan.ndesired), ngenes), np.float64) if full_stats else None
    check(lib().scranb200_score_markers(m.ctx.handle, m.handle, plan.handle, ptr(out), ptr(full)))
    res = OrderedDict()
    for k in range(plan.K):
        g = plan.universe[plan.out_groups[k]]
        cols = OrderedDict()
        for c, name in enumerate(COLUMNS):
            col = out[k, c]
            if name.startswith("rank."):
                col = np.where(np.isfinite(col), col, np.iinfo(np.int32).max).astype(np.int64)
            cols[name] = col
            if full_stats and name.startswith("rank."):
                e = EFFECTS.index(name[5:])
                lo, hi = plan.right_ptr[k], plan.right_ptr[k + 1]
                cols["full." + name[5:]] = full[e, lo:hi].T
                cols["full." + name[5:] + ".weight"] = plan.weight[lo:hi]
                cols["full." + name[5:] + ".right"] = [plan.universe[r] for r in plan.right_group[lo:hi]]
        res[g] = cols
    if own:
        plan.free()
    return res


def summary_marker_stats(x, groups, block=None):
    """summaryMarkerStats(x, groups) (R/summaryMarkerStats.R:62-95): OrderedDict group -> OrderedDict with
    self.average, other.average, self.detected, other.detected, from the blocked statistics on the device."""
    m = _as_matrix(x)
    if len(groups) != m.shape[1]:
        raise ValueError("length of 'groups' does not equal 'ncol(x)'")
    plan = ScorePlan(groups, block, None, 0.0, ctx=m.ctx)
    out = np.zeros((max(1, plan.K), 4, m.shape[0]), np.float64)
    check(lib().scranb200_summary_marker_stats(m.ctx.handle, m.handle, plan.handle, ptr(out)))
    res = OrderedDict()
    for k in range(plan.K):
        res[plan.universe[plan.out_groups[k]]] = OrderedDict((name, out[k, c]) for c, name in enumerate(COLUMNS[:4]))
    plan.free()
    return res


# ---------------------------------------------------------------------------------------------
# pairwiseWilcox / pairwiseTTests: statistics and p-value layers on the GPU
# ---------------------------------------------------------------------------------------------

def _setup_groups(groups, ncells):
    """.setup_groups (R/utils_markers.R:1-29) on codes: returns levels and 0-based codes (-1 = NA)."""
    groups = list(groups)
    if len(groups) != ncells:
        raise ValueError("length of 'groups' does not equal 'ncol(x)'")
    levels = sorted(set(g for g in groups if g is not None))
    if len(levels) < 2:
        raise ValueError("need at least two unique levels in 'groups'")
    code = {g: k for k, g in enumerate(levels)}
    return levels, np.array([code[g] if g is not None else -1 for g in groups], np.int64)


def _block_codes(block, ncells, keep):
    """Block levels in sorted order -> int32 codes; cells with keep == False do not define a level
    (R/pairwiseTTests.R:238: block[is.na(groups)] <- NA)."""
    if block is None:
        return None, 1
    block = np.asarray(list(block))
    if len(block) != ncells:
        raise ValueError("length of 'block' does not equal 'ncol(x)'")
    blevels = sorted(set(block[keep].tolist()))
    codes = np.searchsorted(np.array(blevels), block) if blevels else np.zeros(ncells, np.int64)
    codes = np.where(keep, np.minimum(codes, max(0, len(blevels) - 1)), 0)
    return np.ascontiguousarray(codes, np.int32), max(1, len(blevels))


def _pair_list(levels, eff, lp, lf):
    import warnings
    out, k = [], 0
    for a in levels:
        for b in levels:
            if a == b:
                continue
            if np.all(np.isnan(lp[k])) and levels.index(b) < levels.index(a):
                warnings.warn("no within-block comparison between %s and %s" % (a, b))
            out.append(dict(first=a, second=b, effect=eff[k].copy(), log_p=lp[k].copy(), log_fdr=lf[k].copy()))
            k += 1
    return out


def pairwise_wilcox(x, groups, block=None, lfc=0.0, direction="any"):
    """pairwiseWilcox(x, groups, block=, lfc=, direction=, log.p=TRUE): list of dicts
    (first, second, effect=AUC, log_p, log_fdr) in the reference's pair order. The statistics and the
    p-value layer (normal approximation, Stouffer over the block levels, log BH) run on the device."""
    if direction not in ("any", "up", "down"):
        raise ValueError("'direction' should be one of 'any', 'up', 'down'")
    m = _as_matrix(x)
    levels, codes = _setup_groups(groups, m.shape[1])
    G = len(levels)
    bcodes, nb = _block_codes(block, m.shape[1], np.ones(m.shape[1], bool))
    gcodes = np.ascontiguousarray(codes, np.int32)
    nout = G * (G - 1)
    eff = np.zeros((max(1, nout), m.shape[0]), np.float64)
    lp = np.zeros_like(eff)
    lf = np.zeros_like(eff)
    check(lib().scranb200_pairwise_wilcox(m.ctx.handle, m.handle, ptr(gcodes), G, ptr(bcodes), nb, float(lfc),
                                          ("any", "up", "down").index(direction), ptr(eff), ptr(lp), ptr(lf)))
    return _pair_list(levels, eff, lp, lf)


def pairwise_t_tests(x, groups, block=None, lfc=0.0, direction="any", std_lfc=False):
    """pairwiseTTests(x, groups, block=, lfc=, direction=, std.lfc=, log.p=TRUE) for the block= branch:
    list of dicts (first, second, effect=logFC, log_p, log_fdr) in the reference's pair order. Blocked
    means / variances, Welch's t-test, the log tails of the t distribution, Stouffer over the block
    levels and log BH all run on the device (scranb200_pairwise_t_tests)."""
    if direction not in ("any", "up", "down"):
        raise ValueError("'direction' should be one of 'any', 'up', 'down'")
    m = _as_matrix(x)
    levels, codes = _setup_groups(groups, m.shape[1])
    G = len(levels)
    bcodes, nb = _block_codes(block, m.shape[1], codes >= 0)
    gcodes = np.ascontiguousarray(codes, np.int32)
    nout = G * (G - 1)
    eff = np.zeros((max(1, nout), m.shape[0]), np.float64)
    lp = np.zeros_like(eff)
    lf = np.zeros_like(eff)
    check(lib().scranb200_pairwise_t_tests(m.ctx.handle, m.handle, ptr(gcodes), G, ptr(bcodes), nb, float(lfc),
                                           ("any", "up", "down").index(direction), int(bool(std_lfc)), ptr(eff), ptr(lp),
                                           ptr(lf)))
    return _pair_list(levels, eff, lp, lf)


# ---------------------------------------------------------------------------------------------
# combineMarkers / findMarkers
# ---------------------------------------------------------------------------------------------

def combine_markers(de_lists, pval_type="any", min_prop=None, effect_field="effect", sorted_rows=True, ctx=None):
    """combineMarkers(de.lists, pairs, pval.type=, min.prop=, log.p.in=TRUE, log.p.out=TRUE, sorted=)
    (R/combineMarkers.R:182-300) on the list pairwise_wilcox / pairwise_t_tests return (every element carries
    its own `first` / `second`, the reference's `pairs`). Returns an OrderedDict first -> OrderedDict with
    "Top" (pval_type "any" only), "log.p.value", "log.FDR", "summary.<effect>", one "<effect>.<second>" column
    per contrast, and "order" (the row order the reference sorts by; applied to every column if sorted_rows)."""
    if pval_type not in ("any", "some", "all"):
        raise ValueError("'pval.type' should be one of 'any', 'some', 'all'")
    if min_prop is None:
        min_prop = 0.0 if pval_type == "any" else 0.5
    ctx = ctx or default_context()
    firsts = []
    for q in de_lists:
        if q["first"] not in firsts:
            firsts.append(q["first"])
    out = OrderedDict()
    for f in firsts:
        cur = [q for q in de_lists if q["first"] == f]
        logp = np.ascontiguousarray(np.stack([q["log_p"] for q in cur], 0), np.float64)
        eff = np.ascontiguousarray(np.stack([q[effect_field] for q in cur], 0), np.float64)
        ncon, ngenes = logp.shape
        top = np.zeros(ngenes, np.int32)
        p = np.zeros(ngenes, np.float64)
        fdr = np.zeros(ngenes, np.float64)
        summ = np.zeros(ngenes, np.float64)
        order = np.zeros(ngenes, np.int32)
        check(lib().scranb200_combine_markers(ctx.handle, ngenes, ncon, ptr(logp), ptr(eff), ("any", "some", "all").index(pval_type),
                                              float(min_prop), ptr(top), ptr(p), ptr(fdr), ptr(summ), ptr(order)))
        cols = OrderedDict()
        if pval_type == "any":
            cols["Top"] = top
        cols["log.p.value"] = p
        cols["log.FDR"] = fdr
        cols["summary." + effect_field] = summ
        for k, q in enumerate(cur):
            cols["%s.%s" % (effect_field, q["second"])] = eff[k]
        if sorted_rows:
            for name in list(cols):
                cols[name] = cols[name][order]
        cols["order"] = order
        out[f] = cols
    return out


def find_markers_wilcox(x, groups, block=None, lfc=0.0, direction="any", pval_type="any", min_prop=None):
    """findMarkers(x, groups, test.type="wilcox", block=, lfc=, direction=, pval.type=, min.prop=, log.p=TRUE)
    (R/findMarkers.R:123-156): pairwiseWilcox followed by combineMarkers, both on the device."""
    return combine_markers(pairwise_wilcox(x, groups, block=block, lfc=lfc, direction=direction), pval_type, min_prop)
