"""Host-side mirror of the reference's interface for the marker-scoring path.

Same names, argument meaning and error behaviour as the reference functions they stand for, so
that the parity tests read like the reference's own tests:

    score_markers          scoreMarkers()            R/scoreMarkers.R:173-266
    pairwise_wilcox        pairwiseWilcox()          R/pairwiseWilcox.R:150-256
    pairwise_t_tests       pairwiseTTests()          R/pairwiseTTests.R:179-309 (block= branch)
    compute_blocked_stats_none                       R/RcppExports.R:24-26
    overlap_exprs / overlap_exprs_paired             R/RcppExports.R:48-54
    compute_min_rank       computeMinRank()          R/computeMinRank.R:45-50

All numerics run in libscranb200.so on the GPU through the C ABI; this module only converts
arguments (what the R closures do before .Call) and reshapes results. The p-value layers of
pairwiseWilcox / pairwiseTTests stay on the host exactly as they stay in R.
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


# ---------------------------------------------------------------------------------------------
# pairwiseWilcox / pairwiseTTests: native statistics on the GPU, p-value layer on the host
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


def _stouffer_log(logps, weights):
    from scipy.special import log_ndtr, ndtri_exp
    z = np.stack([ndtri_exp(lp) for lp in logps], axis=0)
    w = np.asarray(weights, np.float64)[:, None]
    return log_ndtr((z * w).sum(axis=0) / np.sqrt((w ** 2).sum()))


def log_bh(logp):
    """.logBH (R/utils_markers.R:184-190)."""
    logp = np.asarray(logp, np.float64)
    o = np.argsort(logp, kind="stable")
    n = len(o)
    rep = logp[o] + np.log(n / np.arange(1, n + 1))
    rep = np.minimum.accumulate(rep[::-1])[::-1]
    out = np.empty(n)
    out[o] = rep
    return out


def _choose(left, right, direction):
    if direction == "up":
        return right
    if direction == "down":
        return left
    return np.minimum(0, np.minimum(left, right) + np.log(2))


def _blocked_template(levels, nblocks, direction, statfun):
    """.pairwise_blocked_template (R/utils_markers.R:52-130)."""
    import warnings
    res = {}
    for i, host in enumerate(levels):
        for target in levels[:i]:
            outs = [statfun(b, i, levels.index(target)) for b in range(nblocks)]
            valid = [k for k, o in enumerate(outs) if o["valid"]]
            if valid:
                w = np.array([outs[k]["weight"] for k in valid], np.float64)
                cl = _stouffer_log([outs[k]["left"] for k in valid], w)
                cr = _stouffer_log([outs[k]["right"] for k in valid], w)
                hp, tp = _choose(cl, cr, direction), _choose(cr, cl, direction)
                fwd = np.stack([outs[k]["forward"] for k in valid], 1)
                rvs = np.stack([outs[k]["reverse"] for k in valid], 1)
                fe = (fwd * w[None, :]).sum(1) / w.sum()
                re_ = (rvs * w[None, :]).sum(1) / w.sum()
            else:
                n = len(outs[0]["left"])
                hp = tp = fe = re_ = np.full(n, np.nan)
                warnings.warn("no within-block comparison between %s and %s" % (host, target))
            res[(host, target)] = (fe, hp)
            res[(target, host)] = (re_, tp)
    ordered = []
    for a in levels:
        for b in levels:
            if (a, b) in res:
                eff, lp = res[(a, b)]
                ordered.append(dict(first=a, second=b, effect=eff, log_p=lp,
                                    log_fdr=log_bh(lp) if not np.all(np.isnan(lp)) else lp))
    return ordered


def pairwise_wilcox(x, groups, block=None, lfc=0.0, direction="any", host_layer=False):
    """pairwiseWilcox(x, groups, block=, lfc=, direction=, log.p=TRUE): list of dicts
    (first, second, effect=AUC, log_p, log_fdr) in the reference's pair order. The statistics and the
    p-value layer (normal approximation, Stouffer over the block levels, log BH) run on the device;
    host_layer=True keeps the p-value layer in numpy / scipy on top of the device U and tie terms."""
    if host_layer:
        return _pairwise_wilcox_host_layer(x, groups, block, lfc, direction)
    if direction not in ("any", "up", "down"):
        raise ValueError("'direction' should be one of 'any', 'up', 'down'")
    m = _as_matrix(x)
    levels, codes = _setup_groups(groups, m.shape[1])
    G = len(levels)
    if block is None:
        bcodes, nb = None, 1
    else:
        block = np.asarray(list(block))
        if len(block) != m.shape[1]:
            raise ValueError("length of 'block' does not equal 'ncol(x)'")
        blevels = sorted(set(block.tolist()))
        bcodes = np.ascontiguousarray(np.searchsorted(np.array(blevels), block), np.int32)
        nb = len(blevels)
    gcodes = np.ascontiguousarray(codes, np.int32)
    nout = G * (G - 1)
    eff = np.zeros((max(1, nout), m.shape[0]), np.float64)
    lp = np.zeros_like(eff)
    lf = np.zeros_like(eff)
    check(lib().scranb200_pairwise_wilcox(m.ctx.handle, m.handle, ptr(gcodes), G, ptr(bcodes), nb, float(lfc),
                                          ("any", "up", "down").index(direction), ptr(eff), ptr(lp), ptr(lf)))
    out, k = [], 0
    for a in levels:
        for b in levels:
            if a == b:
                continue
            if np.all(np.isnan(lp[k])):
                import warnings
                if levels.index(b) < levels.index(a):
                    warnings.warn("no within-block comparison between %s and %s" % (a, b))
            out.append(dict(first=a, second=b, effect=eff[k].copy(), log_p=lp[k].copy(), log_fdr=lf[k].copy()))
            k += 1
    return out


def _pairwise_wilcox_host_layer(x, groups, block=None, lfc=0.0, direction="any"):
    from scipy.special import log_ndtr
    m = _as_matrix(x)
    levels, codes = _setup_groups(groups, m.shape[1])
    G = len(levels)
    if block is None:
        blocks = [np.arange(m.shape[1])]
    else:
        block = np.asarray(list(block))
        if len(block) != m.shape[1]:
            raise ValueError("length of 'block' does not equal 'ncol(x)'")
        blocks = [np.flatnonzero(block == b) for b in sorted(set(block.tolist()))]
    # R/pairwiseWilcox.R:221-243 makes one native call per block level; here all block levels go through
    # ONE paired call (two with lfc != 0: the mirrored entries use -lfc, src/overlap_exprs.cpp:179-187),
    # every (block, group) combination being a group of its own, and the results are laid out as the
    # per-block overlap_exprs() matrices: entry (i, j) present for j < i (all i != j with lfc != 0).
    all_n, combos, left, right, where = [], [], [], [], []
    for b, chosen in enumerate(blocks):
        cg = codes[chosen]
        by_group = [chosen[cg == g].astype(np.int32) for g in range(G)]
        all_n.append(np.array([len(v) for v in by_group], np.float64))
        base = len(combos)
        combos.extend(by_group)
        for i in range(G):
            if len(by_group[i]) == 0:
                continue  # src/overlap_exprs.cpp:175,177: empty groups are skipped, their entries stay 0
            for j in range(i):
                if len(by_group[j]) == 0:
                    continue
                left.append(base + i + 1)
                right.append(base + j + 1)
                where.append((b, i, j))
    ngenes = m.shape[0]
    all_u = [np.full((G, G, ngenes), np.nan) for _ in blocks]
    all_t = [np.full((G, G, ngenes), np.nan) for _ in blocks]
    for b in range(len(blocks)):
        for i in range(G):
            for j in range(G):
                if j < i or (lfc != 0 and i != j):
                    all_u[b][i, j] = 0.0
                    all_t[b][i, j] = 0.0
    if left:
        for sign in ((1.0,) if lfc == 0 else (1.0, -1.0)):
            U, T = overlap_exprs_paired(m, np.array(left, np.int32), np.array(right, np.int32), combos, sign * lfc,
                                        want_ties=True)
            for k, (b, i, j) in enumerate(where):
                a, c = (i, j) if sign > 0 else (j, i)
                all_u[b][a, c] = U[:, k]
                all_t[b][a, c] = T[:, k]

    def sigma(nh, nt, ties):  # .get_sigma, R/pairwiseWilcox.R:347-350
        s2 = (nh * nt / 12) * ((nh + nt + 1) - ties / ((nh + nt) * (nh + nt - 1)))
        with np.errstate(invalid="ignore"):
            return np.maximum(np.sqrt(s2), 1e-8)

    def statfun(b, host, target):  # R/pairwiseWilcox.R:263-345
        hn, tn = all_n[b][host], all_n[b][target]
        prod = hn * tn
        with np.errstate(invalid="ignore", divide="ignore"):
            if lfc == 0:
                eff = all_u[b][host, target]
                auc = eff / prod
                out = dict(forward=auc, reverse=1 - auc)
                z = eff - prod / 2
                sig = sigma(hn, tn, all_t[b][host, target])
                out["left"] = log_ndtr((z + 0.5) / sig)
                out["right"] = log_ndtr(-(z - 0.5) / sig)
            else:
                minus, added = all_u[b][host, target], all_u[b][target, host]
                if direction == "any":
                    auc = (minus / 2 + added / 2) / prod
                    out = dict(forward=auc, reverse=1 - auc)
                elif direction == "up":
                    out = dict(forward=minus / prod, reverse=1 - added / prod)
                else:
                    out = dict(forward=added / prod, reverse=1 - minus / prod)
                msig = sigma(hn, tn, all_t[b][host, target])
                asig = sigma(hn, tn, all_t[b][target, host])
                out["left"] = log_ndtr(((added - prod / 2) + 0.5) / asig)
                out["right"] = log_ndtr(-((minus - prod / 2) - 0.5) / msig)
        out["weight"] = prod
        out["valid"] = prod > 1.5
        return out

    return _blocked_template(levels, len(blocks), direction, statfun)


def pairwise_t_tests(x, groups, block=None, lfc=0.0, direction="any", std_lfc=False):
    """pairwiseTTests(x, groups, block=, lfc=, direction=, std.lfc=) for the block= branch."""
    from scipy.stats import t as tdist
    m = _as_matrix(x)
    levels, codes = _setup_groups(groups, m.shape[1])
    G = len(levels)
    if block is None:
        bcode = np.zeros(len(codes), np.int64)
        nb = 1
    else:
        block = list(block)
        if len(block) != m.shape[1]:
            raise ValueError("length of 'block' does not equal 'ncol(x)'")
        blv = sorted(set(b for b, g in zip(block, codes) if g >= 0))
        bcode = np.array([blv.index(b) if g >= 0 else 0 for b, g in zip(block, codes)])
        nb = len(blv)
    ids = np.where(codes >= 0, codes + bcode * G, NA_INT).astype(np.int32)  # R/pairwiseTTests.R:242
    means, varis = compute_blocked_stats_none(m, ids, nb * G)
    ncells = np.bincount(ids[ids >= 0], minlength=nb * G).astype(np.float64)
    thr = abs(lfc)

    def statfun(b, host, target):  # R/pairwiseTTests.R:270-302, 311-332, 439-461
        hn, tn = ncells[b * G + host], ncells[b * G + target]
        hv, tv = varis[:, b * G + host], varis[:, b * G + target]
        hs2, ts2 = np.maximum(hv, 1e-8), np.maximum(tv, 1e-8)
        hdf, tdf = max(0.0, hn - 1), max(0.0, tn - 1)
        n = means.shape[0]
        if hdf > 0 and tdf > 0:
            herr, terr = hs2 / hn, ts2 / tn
            err = herr + terr
            df = err ** 2 / (herr ** 2 / hdf + terr ** 2 / tdf)
        else:
            err = df = np.full(n, np.nan)
        cur = means[:, b * G + host] - means[:, b * G + target]
        with np.errstate(invalid="ignore", divide="ignore"):
            if thr == 0:
                tt = cur / np.sqrt(err)
                left, right = tdist.logcdf(tt, df), tdist.logsf(tt, df)
            else:
                left = tdist.logcdf((cur + thr) / np.sqrt(err), df)
                right = tdist.logsf((cur - thr) / np.sqrt(err), df)
            eff = cur.copy()
            if std_lfc:
                pooled = ((hn - 1) * hv + (tn - 1) * tv) / (tn + hn - 2)
                zero = eff == 0
                eff = eff / np.sqrt(pooled)
                eff[zero] = 0
            wgt = 1 / (1 / hn + 1 / tn) if hn > 0 and tn > 0 else 0.0
        return dict(forward=eff, reverse=-eff, left=left, right=right, weight=wgt, valid=bool(np.all(~np.isnan(df))))

    return _blocked_template(levels, nb, direction, statfun)
