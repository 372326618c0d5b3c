"""TEST INFRASTRUCTURE ONLY: numpy restatement of the R-level arithmetic above the two native
kernels, i.e. what the reference does in R between the user call and the returned DataFrames.

Follows, with the operation order kept:
  * scoreMarkers           /root/reference/R/scoreMarkers.R:173-266 (driver), :272-345 (setup),
                           :351-445 (effects), :463-582 (averaging, collation)
  * computeMinRank         /root/reference/R/computeMinRank.R:45-50
  * pairings               /root/reference/R/expandPairings.R:1-58
  * pairwiseWilcox layer   /root/reference/R/pairwiseWilcox.R:192-350
  * pairwiseTTests layer   /root/reference/R/pairwiseTTests.R:226-332, :439-461
  * block combination      /root/reference/R/utils_markers.R:52-120, :132-147, :184-190
The native parts (blocked mean/variance, U and tie term) come from oracle.native (the
reference's own object code when built, else the C port).

Third-party arithmetic that is NOT under /root/reference and is restated from documented
behaviour: matrixStats::rowWeightedMeans (weights standardised to sum
to one, product, long-double row sum), base rowMeans (long-double sum / n), matrixStats
rowMedians/rowMins/rowMaxs/colRanks(ties="min"), scuttle::summarizeAssayByGroup ("> 0" counts),
scuttle::correctGroupSummary (identity when unblocked; blocked case: weighted least squares on
~0+group+block, PARITY UNPINNED), metapod Stouffer combination, stats::pnorm/pt via scipy.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
"""
import itertools
from collections import OrderedDict

import numpy as np

from . import native

LD = np.longdouble  # R accumulates row sums in long double on x86-64


# ----------------------------------------------------------------------------------------------
# Setup bookkeeping
# ----------------------------------------------------------------------------------------------

def expand_pairings(pairings, universe):
    """R/expandPairings.R:1-58 -> (id1, id2), 0-based indices into universe."""
    universe = list(universe)
    pos = {v: k for k, v in enumerate(universe)}

    def subset(request, clean=True):
        if request is None:
            out = list(range(len(universe)))
        else:
            out = [pos.get(v, None) for v in request]
        if clean:
            seen = []
            for v in out:
                if v is not None and v not in seen:
                    seen.append(v)
            out = seen
        return out

    def clean_expand(xs, ys):  # expand.grid: first factor varies fastest
        return [(a, b) for b in ys for a in xs if a != b]

    if isinstance(pairings, np.ndarray) and pairings.ndim == 2:
        if pairings.shape[1] != 2:
            raise ValueError("'pairings' should be a numeric/character matrix with 2 columns")
        s1 = subset(list(pairings[:, 0]), clean=False)
        s2 = subset(list(pairings[:, 1]), clean=False)
        pairs = [(a, b) for a, b in zip(s1, s2) if a is not None and b is not None]
    elif isinstance(pairings, (list, tuple)) and len(pairings) == 2 and all(
            isinstance(p, (list, tuple, np.ndarray)) or p is None for p in pairings):
        conv = [subset(None if p is None else list(p)) for p in pairings]
        pairs = clean_expand(conv[0], conv[1])
    else:
        avail = subset(None if pairings is None else list(pairings))
        pairs = clean_expand(avail, avail)
    return [a for a, _ in pairs], [b for _, b in pairs]


def _uniquify(rows):
    """R/scoreMarkers.R:272-290: unique tuples in sorted order, and each row's index into them."""
    uniq = sorted(set(rows))
    lookup = {r: k for k, r in enumerate(uniq)}
    return uniq, np.array([lookup[r] for r in rows], dtype=np.int64)


class ScoreSetup:
    """Everything .scoreMarkers computes before touching the matrix (R/scoreMarkers.R:191-217)."""

    def __init__(self, groups, block=None, pairings=None, lfc=0.0):
        groups = list(groups)
        self.universe = sorted(set(groups))
        id1, id2 = expand_pairings(pairings, self.universe)
        self.desired = [(self.universe[a], self.universe[b]) for a, b in zip(id1, id2)]
        used = set(a for a, _ in self.desired) | set(b for _, b in self.desired)
        self.keep = np.array([g in used for g in groups], dtype=bool)  # :196-202
        groups = [g for g, k in zip(groups, self.keep) if k]
        if block is not None:
            block = [b for b, k in zip(list(block), self.keep) if k]
        self.has_block = block is not None

        rows = [(g,) for g in groups] if block is None else list(zip(groups, block))
        self.combos, self.comb_id = _uniquify(rows)  # sorted by (group, block): :277-287
        C = len(self.combos)
        self.ncells = np.bincount(self.comb_id, minlength=C).astype(np.float64)  # :208
        self.collapse = (lfc == 0)

        # .reindex_comparisons_for_combinations (:304-339)
        if block is None:
            by_block = [list(range(C))]
        else:
            levels = sorted(set(c[1] for c in self.combos))
            by_block = [[k for k in range(C) if self.combos[k][1] == lv] for lv in levels]
        left, right = [], []
        for indices in by_block:
            grp = {self.combos[k][0]: k for k in indices}
            seen = OrderedDict()
            for (dl, dr) in self.desired:
                if dl in grp and dr in grp:
                    a, b = grp[dl], grp[dr]
                    key = (max(a, b), min(a, b)) if self.collapse else (a, b)
                    seen.setdefault(key, None)
            left += [k[0] for k in seen]
            right += [k[1] for k in seen]
        self.left = np.array(left, dtype=np.int64)    # 0-based combination indices
        self.right = np.array(right, dtype=np.int64)

        # .group_by_used_combinations (:341-345)
        usedc = set(left) | set(right)
        self.involved = [np.flatnonzero(self.comb_id == k).astype(np.int32) if k in usedc else np.zeros(0, np.int32)
                         for k in range(C)]

        # .identify_effects_to_average (:463-476)
        gp = [(self.combos[a][0], self.combos[b][0]) for a, b in zip(left, right)]
        self.averaged, aid = _uniquify(gp)
        self.to_average = [np.flatnonzero(aid == k) for k in range(len(self.averaged))]

        # .cross_reference_to_desired (:495-521)
        apos = {p: k for k, p in enumerate(self.averaged)}
        fpos = {(p[1], p[0]): k for k, p in enumerate(self.averaged)} if self.collapse else {}
        self.desired_indices = OrderedDict()
        for lg in sorted(set(a for a, _ in self.desired)):
            chosen = [(a, b) for (a, b) in self.desired if a == lg]
            self.desired_indices[lg] = dict(
                right=[b for _, b in chosen],
                direct=[apos.get(p, None) for p in chosen],
                flipped=[fpos.get(p, None) for p in chosen],
            )


# ----------------------------------------------------------------------------------------------
# Small numeric helpers that mimic R's accumulation
# ----------------------------------------------------------------------------------------------

def row_weighted_means(e, w):
    """matrixStats::rowWeightedMeans(e, w, na.rm=TRUE) as used at R/scoreMarkers.R:488."""
    e = np.asarray(e, dtype=np.float64)
    w = np.asarray(w, dtype=np.float64)
    valid = ~np.isnan(e)
    W = np.where(valid, w[None, :], 0.0)
    wS = W.astype(LD).sum(axis=1).astype(np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        Wn = W / wS[:, None]
        prod = Wn * np.where(valid, e, 0.0)
    return prod.astype(LD).sum(axis=1).astype(np.float64)


def row_means_narm(m):
    valid = ~np.isnan(m)
    s = np.where(valid, m, 0.0).astype(LD).sum(axis=1)
    n = valid.sum(axis=1)
    with np.errstate(invalid="ignore", divide="ignore"):
        return (s / n.astype(LD)).astype(np.float64)


def row_medians_narm(m):
    out = np.full(m.shape[0], np.nan)
    for r in range(m.shape[0]):
        v = np.sort(m[r][~np.isnan(m[r])])
        n = v.size
        if n == 0:
            continue
        out[r] = v[n // 2] if n % 2 else (v[n // 2 - 1] + v[n // 2]) / 2
    return out


def row_min_narm(m):
    return np.where(np.isnan(m), np.inf, m).min(axis=1)


def row_max_narm(m):
    return np.where(np.isnan(m), -np.inf, m).max(axis=1)


def compute_min_rank(x):
    """R/computeMinRank.R:45-50 with the defaults (ties.method='min', decreasing=TRUE):
    per column rank = 1 + #(non-NA entries strictly greater); NA stays NA; per row the minimum
    over columns ignoring NA. Rows with no non-NA entry get int32 max (R: Inf with a warning)."""
    x = np.asarray(x, dtype=np.float64)
    n, m = x.shape
    best = np.full(n, np.iinfo(np.int32).max, dtype=np.int64)
    for j in range(m):
        col = x[:, j]
        ok = ~np.isnan(col)
        v = np.sort(col[ok])
        rank = 1 + (v.size - np.searchsorted(v, col[ok], side="right"))
        best[ok] = np.minimum(best[ok], rank)
    return best


# ----------------------------------------------------------------------------------------------
# scoreMarkers
# ----------------------------------------------------------------------------------------------

def cohen_d(means, varis, left, right, lfc):
    """R/scoreMarkers.R:417-426."""
    delta = means[:, left] - means[:, right] - lfc
    pooled = (varis[:, left] + varis[:, right]) / 2
    with np.errstate(invalid="ignore", divide="ignore"):
        d = delta / np.sqrt(pooled)
    d[delta == 0] = 0
    return d


def lfc_detected(ndet, left, right, nl, nr):
    """R/scoreMarkers.R:428-440."""
    mn = np.minimum(nl, nr)
    lp = 1 * nl / mn
    rp = 1 * nr / mn
    lprop = (ndet[:, left] + lp[None, :]) / (nl + lp * 2)[None, :]
    rprop = (ndet[:, right] + rp[None, :]) / (nr + rp * 2)[None, :]
    return np.log2(lprop / rprop)


def _num_detected(m, comb_id, C):
    """scuttle::summarizeAssayByGroup(statistics='num.detected'): count of values > 0."""
    import scipy.sparse as sp
    if sp.issparse(m):
        pos = (m > 0).astype(np.float64).tocsc()
    else:
        pos = (np.asarray(m) > 0).astype(np.float64)
    ind = np.zeros((len(comb_id), C))
    ind[np.arange(len(comb_id)), comb_id] = 1
    return np.asarray(pos @ ind)


def correct_group_summary(mat, combos, has_block, weights, transform):
    """scuttle::correctGroupSummary as called at R/scoreMarkers.R:395-399. Unblocked (or every
    group seen in one block only): the per-group value itself. Blocked: PARITY UNPINNED (the
    package source is not available here) - weighted least squares on ~0 + group + block with
    weights = cells per combination; the group coefficients are returned (first block level is
    the baseline), after the logit round trip with offset 0.01 for proportions."""
    groups = sorted(set(c[0] for c in combos))
    x = mat.copy()
    if transform == "logit":
        off = 0.01
        x = (x + off) / (1 + 2 * off)
        x = np.log(x / (1 - x))
    gi = np.array([groups.index(c[0]) for c in combos])
    blocks = sorted(set(c[1] for c in combos)) if has_block else [None]
    if len(blocks) <= 1:
        D = np.zeros((len(combos), len(groups)))
        D[np.arange(len(combos)), gi] = 1
    else:
        bi = np.array([blocks.index(c[1]) for c in combos])
        D = np.zeros((len(combos), len(groups) + len(blocks) - 1))
        D[np.arange(len(combos)), gi] = 1
        for k in range(len(combos)):
            if bi[k] > 0:
                D[k, len(groups) + bi[k] - 1] = 1
    if len(blocks) <= 1 and len(set(gi)) == len(gi):
        avg = x[:, np.argsort(gi)]  # one point per group: the fit is the point itself
    else:
        sw = np.sqrt(weights)
        coef, *_ = np.linalg.lstsq(D * sw[:, None], (x * sw[None, :]).T, rcond=None)
        avg = coef[:len(groups)].T
    if transform == "logit":
        avg = np.exp(avg)
        avg = avg / (avg + 1)
        avg = avg * (1 + 2 * 0.01) - 0.01
        avg = np.clip(avg, 0, 1)
    return groups, avg


def score_markers(x, groups, block=None, pairings=None, lfc=0.0, full_stats=False, lib=None):
    """The whole of scoreMarkers on one matrix. x: scipy.sparse or dense [ngenes, ncells].
    Returns OrderedDict group -> OrderedDict column name -> array, columns in the reference's
    order (R/scoreMarkers.R:370, 392, 401-411, 562-576); rank.* are the min-ranks."""
    import scipy.sparse as sp
    lib = lib or native.load("best")
    st = ScoreSetup(groups, block, pairings, lfc)
    x = x.tocsc()[:, np.flatnonzero(st.keep)] if sp.issparse(x) else np.asarray(x, dtype=np.float64)[:, st.keep]
    ngenes = x.shape[0]
    C = len(st.combos)
    nl = st.ncells[st.left]
    nr = st.ncells[st.right]

    means, varis = lib.blocked_stats(x, st.comb_id.astype(np.int32), C)  # :358
    effects = OrderedDict()
    effects["logFC.cohen"] = cohen_d(means, varis, st.left, st.right, lfc)  # :359
    ndet = _num_detected(x, st.comb_id, C)  # :361-363
    U = lib.overlap_paired(x, st.left + 1, st.right + 1, st.involved, lfc)  # :366, 442-445
    effects["AUC"] = U / (nl * nr)[None, :]
    effects["logFC.detected"] = lfc_detected(ndet, st.left, st.right, nl, nr)  # :364
    weights = nl * nr  # :369

    out = OrderedDict((g, OrderedDict()) for g in st.desired_indices)
    rank_inputs = {}
    for name, eff in effects.items():
        rev = (lambda v: 1 - v) if name == "AUC" else (lambda v: -v)  # :373-377
        averaged = []
        cw = []
        for idx in st.to_average:  # :479-492
            cw.append(weights[idx].sum())
            averaged.append(row_weighted_means(eff[:, idx], weights[idx]))
        for g, cur in st.desired_indices.items():  # :525-582
            cols = []
            w = []
            for d, f in zip(cur["direct"], cur["flipped"]):
                if d is not None:
                    cols.append(averaged[d])
                    w.append(cw[d])
                elif f is not None:
                    cols.append(rev(averaged[f]))
                    w.append(cw[f])
                else:
                    cols.append(np.full(ngenes, np.nan))
                    w.append(0.0)
            mat = np.stack(cols, axis=1)
            o = out[g]
            o["mean." + name] = row_means_narm(mat)
            o["min." + name] = row_min_narm(mat)
            o["median." + name] = row_medians_narm(mat)
            o["max." + name] = row_max_narm(mat)
            o["rank." + name] = compute_min_rank(mat)  # :251-256
            if full_stats:
                o["full." + name] = mat
                o["full." + name + ".weight"] = np.array(w)
                o["full." + name + ".right"] = list(cur["right"])
            rank_inputs[(g, name)] = mat

    # self/other average and detected (:395-411); prepended in the order detected, then average,
    # each as (self, other), so the final order is self.average, other.average, self.detected, ...
    prop = ndet / st.ncells[None, :]
    glist, det = correct_group_summary(prop, st.combos, st.has_block, st.ncells, "logit")
    _, ave = correct_group_summary(means, st.combos, st.has_block, st.ncells, "raw")
    final = OrderedDict()
    for g, cur in st.desired_indices.items():
        cols = OrderedDict()
        others = [glist.index(r) for r in cur["right"] if r in glist]
        me = glist.index(g)
        for label, mat in (("average", ave), ("detected", det)):
            cols["self." + label] = mat[:, me]
            cols["other." + label] = row_means_narm(mat[:, sorted(set(others))]) if others else np.full(ngenes, np.nan)
        cols.update(out[g])
        final[g] = cols
    return final


# ----------------------------------------------------------------------------------------------
# pairwiseWilcox / pairwiseTTests statistic layers
# ----------------------------------------------------------------------------------------------

def _stouffer_log(logps, weights):
    """metapod::combineParallelPValues(method='stouffer', log.p=TRUE): z = qnorm(p), weighted sum
    over sqrt(sum w^2), back to a log lower-tail p-value."""
    from scipy.special import log_ndtr, ndtri_exp
    z = np.stack([ndtri_exp(lp) for lp in logps], axis=0)
    w = np.asarray(weights, dtype=np.float64)[:, None]
    comb = (z * w).sum(axis=0) / np.sqrt((w ** 2).sum())
    return log_ndtr(comb)


def log_bh(logp):
    """R/utils_markers.R:184-190."""
    logp = np.asarray(logp, dtype=np.float64)
    o = np.argsort(logp, kind="stable")
    n = len(o)
    rep = logp[o] + np.log(n / np.arange(1, n + 1))
    rep = np.minimum.accumulate(rep[::-1])[::-1]
    out = np.empty(n)
    out[o] = rep
    return out


def _choose(left, right, direction):
    """R/utils_markers.R:132-147."""
    if direction == "up":
        return right
    if direction == "down":
        return left
    return np.minimum(0, np.minimum(left, right) + np.log(2))


def _blocked_template(levels, nblocks, direction, statfun):
    """R/utils_markers.R:52-120. -> list of (first, second, effect, log.p, log.FDR) in the
    reference's output order (:122-130)."""
    res = {}
    for i, host in enumerate(levels):
        for target in levels[:i]:
            outs = [statfun(b, host, target) for b in range(nblocks)]
            valid = [k for k, o in enumerate(outs) if o["valid"]]
            if valid:
                w = np.array([outs[k]["weight"] for k in valid], dtype=np.float64)
                cl = _stouffer_log([outs[k]["left"] for k in valid], w)
                cr = _stouffer_log([outs[k]["right"] for k in valid], w)
                hp = _choose(cl, cr, direction)
                tp = _choose(cr, cl, direction)
                fwd = np.stack([outs[k]["forward"] for k in valid], 1)
                rvs = np.stack([outs[k]["reverse"] for k in valid], 1)
                fe = (fwd * w[None, :]).sum(1) / w.sum()  # metapod::averageParallelStats
                re_ = (rvs * w[None, :]).sum(1) / w.sum()
            else:
                n = len(outs[0]["left"])
                hp = tp = fe = re_ = np.full(n, np.nan)
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


def get_sigma(nh, nt, ties):
    """R/pairwiseWilcox.R:347-350."""
    s2 = (nh * nt / 12) * ((nh + nt + 1) - ties / ((nh + nt) * (nh + nt - 1)))
    with np.errstate(invalid="ignore"):
        return np.maximum(np.sqrt(s2), 1e-8)


def pairwise_wilcox(x, groups, block=None, lfc=0.0, direction="any", lib=None):
    """R/pairwiseWilcox.R:192-345 on group codes 0..G-1 (negative = NA). x [ngenes, ncells]."""
    from scipy.special import log_ndtr
    lib = lib or native.load("best")
    groups = np.asarray(groups)
    G = int(groups.max()) + 1
    levels = list(range(G))
    blocks = [np.arange(len(groups))] if block is None else [np.flatnonzero(np.asarray(block) == b)
                                                              for b in sorted(set(block))]
    all_n, all_u, all_t = [], [], []
    for chosen in blocks:
        cg = groups[chosen]
        by_group = [chosen[cg == g].astype(np.int32) for g in levels]
        all_n.append(np.array([len(v) for v in by_group], dtype=np.float64))
        U, T = lib.overlap(x, by_group, lfc)
        all_u.append(U)
        all_t.append(T)

    def statfun(b, host, target):
        hn, tn = all_n[b][host], all_n[b][target]
        prod = hn * tn
        with np.errstate(invalid="ignore", divide="ignore"):
            if lfc == 0:
                eff = all_u[b][host, target]
                auc = eff / prod
                out = dict(forward=auc, reverse=1 - auc)
                z = eff - prod / 2
                sig = get_sigma(hn, tn, all_t[b][host, target])
                out["left"] = log_ndtr((z + 0.5) / sig)
                out["right"] = log_ndtr(-(z - 0.5) / sig)
            else:
                minus = all_u[b][host, target]
                added = all_u[b][target, host]
                if direction == "any":
                    auc = (minus / 2 + added / 2) / prod
                    out = dict(forward=auc, reverse=1 - auc)
                elif direction == "up":
                    out = dict(forward=minus / prod, reverse=1 - added / prod)
                else:
                    out = dict(forward=added / prod, reverse=1 - minus / prod)
                msig = get_sigma(hn, tn, all_t[b][host, target])
                asig = get_sigma(hn, tn, all_t[b][target, host])
                out["left"] = log_ndtr(((added - prod / 2) + 0.5) / asig)
                out["right"] = log_ndtr(-((minus - prod / 2) - 0.5) / msig)
        out["weight"] = prod
        out["valid"] = prod > 1.5
        return out

    return _blocked_template(levels, len(blocks), direction, statfun)


def log_pt(t, df, lower=True):
    """stats::pt(t, df, lower.tail=lower, log.p=TRUE). scipy's t.logcdf / logsf are log(cdf): exact enough
    until the tail underflows (log p < ~ -700); beyond that the tail is the regularised incomplete beta
    function in 50-digit arithmetic (mpmath), elementwise - only a handful of entries ever get there."""
    from scipy.stats import t as tdist
    t = np.asarray(t, np.float64)
    df = np.broadcast_to(np.asarray(df, np.float64), t.shape)
    with np.errstate(invalid="ignore", divide="ignore"):
        out = np.array(tdist.logcdf(t, df) if lower else tdist.logsf(t, df), np.float64)
    deep = np.isfinite(t) & np.isfinite(df) & ((out < -650) | np.isneginf(out))
    if deep.any():
        import mpmath as mp
        with mp.workdps(50):
            for k in np.flatnonzero(deep.ravel()):
                tt, dd = mp.mpf(float(t.ravel()[k])), mp.mpf(float(df.ravel()[k]))
                a, b, x = dd / 2, mp.mpf(1) / 2, dd / (dd + tt * tt)
                try:
                    tail = mp.betainc(a, b, 0, x, regularized=True) / 2
                except ValueError:  # the hypergeometric series gives up on a few (t, df): Lentz's continued fraction
                    c, d = mp.mpf(1), 1 / (1 - (a + b) * x / (a + 1))
                    h = d
                    for m in range(1, 100000):
                        aa = m * (b - m) * x / ((a - 1 + 2 * m) * (a + 2 * m))
                        d, c = 1 / (1 + aa * d), 1 + aa / c
                        h *= d * c
                        aa = -(a + m) * (a + b + m) * x / ((a + 2 * m) * (a + 1 + 2 * m))
                        d, c = 1 / (1 + aa * d), 1 + aa / c
                        h *= d * c
                        if abs(d * c - 1) < mp.mpf(10) ** -45:
                            break
                    tail = mp.exp(a * mp.log(x) + b * mp.log(1 - x) - (mp.loggamma(a) + mp.loggamma(b) - mp.loggamma(a + b))) * h / a / 2
                out.ravel()[k] = float(mp.log(tail))  # the requested tail is the far one, else log p would be ~ 0
    return out


def pairwise_ttests(x, groups, block=None, lfc=0.0, direction="any", std_lfc=False, lib=None):
    """R/pairwiseTTests.R:226-332, 439-461 on group codes 0..G-1 (negative = NA)."""
    lib = lib or native.load("best")
    groups = np.asarray(groups)
    G = int(groups.max()) + 1
    levels = list(range(G))
    if block is None:
        bcode = np.zeros(len(groups), dtype=np.int64)
        nb = 1
    else:
        blv = sorted(set(b for b, g in zip(block, groups) if g >= 0))
        bcode = np.array([blv.index(b) if g >= 0 else 0 for b, g in zip(block, groups)])
        nb = len(blv)
    ids = np.where(groups >= 0, groups + bcode * G, native.NA_INT).astype(np.int32)  # :242
    means, varis = lib.blocked_stats(x, ids, nb * G)
    ncells = np.bincount(ids[ids >= 0], minlength=nb * G).astype(np.float64)
    thr = abs(lfc)

    def statfun(b, host, target):
        hn, tn = ncells[b * G + host], ncells[b * G + target]
        hs2 = np.maximum(varis[:, b * G + host], 1e-8)
        ts2 = np.maximum(varis[:, b * G + target], 1e-8)
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
                left = log_pt(tt, df, True)
                right = log_pt(tt, df, False)
            else:
                left = log_pt((cur + thr) / np.sqrt(err), df, True)
                right = log_pt((cur - thr) / np.sqrt(err), df, False)
            eff = cur.copy()
            if std_lfc:
                pooled = ((hn - 1) * varis[:, b * G + host] + (tn - 1) * varis[:, b * G + target]) / (tn + hn - 2)
                zero = eff == 0
                eff = eff / np.sqrt(pooled)
                eff[zero] = 0
            wgt = 1 / (1 / hn + 1 / tn) if hn > 0 and tn > 0 else 0.0
        return dict(forward=eff, reverse=-eff, left=left, right=right, weight=wgt,
                    valid=bool(np.all(~np.isnan(df))))

    return _blocked_template(levels, nb, direction, statfun)


# ----------------------------------------------------------------------------------------------
# combineMarkers (R/combineMarkers.R:182-330) on log p-values
# ----------------------------------------------------------------------------------------------

def combine_parallel_log_p(logp, method, min_prop):
    """metapod::combineParallelPValues(list, method=, log.p=TRUE, min.prop=) restated from what the reference's
    tests pin (tests/testthat/test-combine-markers.R:19-137): "simes" = min over the contrasts of the BH-adjusted
    p-values; "berger" = the largest; "holm-min" = the ceiling(n min.prop)-th smallest Holm-adjusted one, n = the
    gene's number of non-NA p-values. logp: [ncon, ngenes], NaN = NA. All NA -> NA."""
    logp = np.asarray(logp, np.float64)
    ncon, ngenes = logp.shape
    out = np.full(ngenes, np.nan)
    for g in range(ngenes):
        p = np.sort(logp[~np.isnan(logp[:, g]), g])
        n = len(p)
        if n == 0:
            continue
        k = np.arange(1, n + 1)
        if method == "simes":
            out[g] = min(0.0, np.min(p + np.log(n / k)))
        elif method == "berger":
            out[g] = p[-1]
        elif method == "holm-min":
            adj = np.maximum.accumulate(np.minimum(0.0, p + np.log(n - k + 1)))
            jump = max(int(np.ceil(n * min_prop)), 1) - 1
            out[g] = adj[jump]
        else:
            raise ValueError(method)
    return out


def rank_first(v):
    """rank(v, ties.method="first", na.last="keep") -> int32 with NA_INT for NA."""
    v = np.asarray(v, np.float64)
    out = np.full(len(v), native.NA_INT, np.int32)
    ok = np.flatnonzero(~np.isnan(v))
    o = ok[np.argsort(v[ok], kind="stable")]
    out[o] = np.arange(1, len(o) + 1)
    return out


def order_na_last(v):
    """order(v): ascending, ties by position, NA last."""
    v = np.asarray(v, np.float64)
    ok = np.flatnonzero(~np.isnan(v))
    na = np.flatnonzero(np.isnan(v))
    return np.concatenate([ok[np.argsort(v[ok], kind="stable")], na])


def combine_markers_one(logp, effects, pval_type="any", min_prop=None, lib=None):
    """.combine_markers_internal (R/combineMarkers.R:232-300) for the contrasts of one first group, log.p.in =
    log.p.out = TRUE, sorted = TRUE. logp / effects: [ncon, ngenes]. Returns dict(Top (any only), log_p, log_fdr,
    summary, order): everything in the ORIGINAL gene order plus the ordering the reference applies last. Top and
    the "some" summary effect come from the reference's own native helpers (lib)."""
    lib = lib or native.load("best")
    logp = np.asarray(logp, np.float64)
    method = {"any": "simes", "some": "holm-min", "all": "berger"}[pval_type]
    if min_prop is None:
        min_prop = 0.0 if pval_type == "any" else 0.5
    pval = combine_parallel_log_p(logp, method, min_prop)
    res = {}
    if pval_type == "any":
        ranked = np.stack([rank_first(logp[c]) for c in range(logp.shape[0])], 0)
        top = lib.top_from_ranks(ranked, min_prop)
        res["Top"] = top
        order = order_na_last(np.where(top == native.NA_INT, np.nan, top.astype(np.float64)))
    else:
        order = order_na_last(pval)
    res["log_p"] = pval
    res["log_fdr"] = log_bh(pval) if not np.any(np.isnan(pval)) else np.full(len(pval), np.nan)
    if effects is not None:
        effects = np.asarray(effects, np.float64)
        ngenes = logp.shape[1]
        chosen = np.full(ngenes, np.nan)
        if pval_type == "all":      # .choose_effect_size (:302-330): the effect of the largest p-value
            worst = np.full(ngenes, -np.inf)
            for c in range(logp.shape[0]):
                worse = (logp[c] > worst) & ~np.isnan(logp[c])
                chosen[worse] = effects[c][worse]
                worst[worse] = logp[c][worse]
        elif pval_type == "any":
            best = np.full(ngenes, np.inf)
            for c in range(logp.shape[0]):
                better = (logp[c] < best) & ~np.isnan(logp[c])
                chosen[better] = effects[c][better]
                best[better] = logp[c][better]
        else:
            chosen = lib.choose_middle(logp, effects, min_prop)
        res["summary"] = chosen
    res["order"] = order
    return res
