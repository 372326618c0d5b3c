"""TEST INFRASTRUCTURE ONLY: ctypes bindings for the two CPU checkers.

``port`` = oracle/liboracle_port.so, our plain-C restatement (oracle/port.c).
``ref``  = oracle/_ref/liboracle_ref.so, the reference's own src/overlap_exprs.cpp and
src/compute_blocked_stats.cpp compiled unmodified against oracle/stubs (only present when
it was built in a container that has /root/reference; it then travels with the repo).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
NA_INT = np.iinfo(np.int32).min

_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def build(with_ref=True):
    """Compile the checkers (gcc/g++ only). The ref target is a no-op without /root/reference."""
    subprocess.check_call(["make", "-s", "-C", HERE, "port"])
    if with_ref:
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


class Matrix:
    """A CSC or dense column-major double matrix in the form both checkers take."""

    def __init__(self, x):
        try:
            import scipy.sparse as sp
        except ImportError:  # pragma: no cover
            sp = None
        if sp is not None and sp.issparse(x):
            x = x.tocsc()
            if not x.has_sorted_indices:
                x = x.sorted_indices()
            self.sparse = 1
            self.nrow, self.ncol = x.shape
            self.p = np.ascontiguousarray(x.indptr, dtype=np.int64)
            self.i = np.ascontiguousarray(x.indices, dtype=np.int32)
            self.x = np.ascontiguousarray(x.data, dtype=np.float64)
        else:
            x = np.asarray(x, dtype=np.float64)
            self.sparse = 0
            self.nrow, self.ncol = x.shape
            self.p = np.zeros(1, np.int64)
            self.i = np.zeros(1, np.int32)
            self.x = np.ascontiguousarray(x.T).ravel()  # column-major storage

    def args(self):
        return (self.nrow, self.ncol, self.p, self.i, self.x, self.sparse)


def _flatten_groups(groups):
    gptr = np.zeros(len(groups) + 1, np.int64)
    for k, g in enumerate(groups):
        gptr[k + 1] = gptr[k] + len(g)
    cells = np.concatenate([np.asarray(g, np.int32) for g in groups]) if len(groups) else np.zeros(0, np.int32)
    if cells.size == 0:
        cells = np.zeros(1, np.int32)
    return gptr, np.ascontiguousarray(cells, dtype=np.int32)


class Checker:
    """One of the two CPU libraries behind a common Python surface."""

    def __init__(self, kind):
        self.kind = kind
        if kind == "port":
            path = os.path.join(HERE, "liboracle_port.so")
            prefix = "oracle_port_"
        elif kind == "ref":
            path = os.path.join(HERE, "_ref", "liboracle_ref.so")
            prefix = "oracle_ref_"
        else:
            raise ValueError(kind)
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.prefix = prefix
        getattr(self.lib, prefix + "last_error").restype = C.c_char_p
        mat = [C.c_int64, C.c_int64, _i64p, _i32p, _f64p, C.c_int]
        f = getattr(self.lib, prefix + "blocked_stats")
        f.restype = C.c_int
        f.argtypes = mat + [_i32p, C.c_int, C.c_int, C.c_void_p, C.c_double, _f64p, _f64p] + (
            [C.c_void_p] if kind == "port" else [])
        f = getattr(self.lib, prefix + "overlap_paired")
        f.restype = C.c_int
        f.argtypes = mat + [_i32p, _i32p, C.c_int64, C.c_int, _i64p, _i32p, C.c_double, _f64p] + (
            [C.c_void_p, C.c_int64, C.c_int64] if kind == "port" else [])
        f = getattr(self.lib, prefix + "overlap")
        f.restype = C.c_int
        f.argtypes = mat + [C.c_int, _i64p, _i32p, C.c_double, _f64p, _f64p]
        f = getattr(self.lib, prefix + "top_from_ranks")
        f.restype = C.c_int
        f.argtypes = [C.c_int, C.c_int64, _i32p, C.c_double, _i32p]
        f = getattr(self.lib, prefix + "choose_middle")
        f.restype = C.c_int
        f.argtypes = [C.c_int, C.c_int64, _f64p, _f64p, C.c_double, _f64p]

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(getattr(self.lib, self.prefix + "last_error")().decode())

    def blocked_stats(self, m, block, nblocks, transform=0, sf=None, pseudo=1.0, want_nz=False):
        """-> (mean, var[, nz]) as [nrow, nblocks] arrays. block: int per cell, NA_INT = skip."""
        if not isinstance(m, Matrix):
            m = Matrix(m)
        block = np.ascontiguousarray(block, dtype=np.int32)
        mean = np.zeros((nblocks, m.nrow), np.float64)
        var = np.zeros((nblocks, m.nrow), np.float64)
        sfp = None
        if sf is not None:
            sf = np.ascontiguousarray(sf, dtype=np.float64)
            sfp = sf.ctypes.data_as(C.c_void_p)
        extra = []
        nz = None
        if self.kind == "port":
            if want_nz:
                nz = np.zeros((nblocks, m.nrow), np.float64)
                extra = [nz.ctypes.data_as(C.c_void_p)]
            else:
                extra = [None]
        elif want_nz:
            raise ValueError("the reference does not return its nnzero matrix")
        self._check(getattr(self.lib, self.prefix + "blocked_stats")(*m.args(), block, nblocks, transform, sfp, pseudo,
                                                            mean.reshape(-1), var.reshape(-1), *extra))
        if want_nz:
            return mean.T, var.T, nz.T
        return mean.T, var.T

    def overlap_paired(self, m, left, right, groups, lfc=0.0, want_ties=False, gene_range=None):
        """left/right 1-based group indices; groups: list of 0-based cell index arrays.
        -> U as [nrow, npairs] (and ties, port only)."""
        if not isinstance(m, Matrix):
            m = Matrix(m)
        left = np.ascontiguousarray(left, dtype=np.int32)
        right = np.ascontiguousarray(right, dtype=np.int32)
        gptr, cells = _flatten_groups(groups)
        out = np.zeros((m.nrow, len(left)), np.float64)
        extra = []
        ties = None
        if self.kind == "port":
            g0, g1 = gene_range if gene_range is not None else (0, 0)
            if want_ties:
                ties = np.zeros((m.nrow, len(left)), np.float64)
                extra = [ties.ctypes.data_as(C.c_void_p), g0, g1]
            else:
                extra = [None, g0, g1]
        elif want_ties or gene_range is not None:
            raise ValueError("not available from the reference entry point")
        lp = left if left.size else np.zeros(1, np.int32)
        rp = right if right.size else np.zeros(1, np.int32)
        op = out.reshape(-1) if out.size else np.zeros(1, np.float64)
        self._check(getattr(self.lib, self.prefix + "overlap_paired")(*m.args(), lp, rp, len(left), len(groups), gptr, cells,
                                                             float(lfc), op, *extra))
        return (out, ties) if want_ties else out

    def overlap(self, m, groups, lfc=0.0):
        """-> (U, T) as [ngroups, ngroups, nrow]; entry (i, j) = column j of the reference's
        matrix i; NaN where the reference has no such column."""
        if not isinstance(m, Matrix):
            m = Matrix(m)
        gptr, cells = _flatten_groups(groups)
        G = len(groups)
        U = np.zeros((G, G, m.nrow), np.float64)
        T = np.zeros((G, G, m.nrow), np.float64)
        self._check(getattr(self.lib, self.prefix + "overlap")(*m.args(), G, gptr, cells, float(lfc), U.reshape(-1),
                                                      T.reshape(-1)))
        return U, T


    def top_from_ranks(self, ranks, prop):
        """compute_Top_statistic_from_ranks(Ranks, prop): ranks [ncon, ngenes] int (NA_INT = NA) -> int[ngenes]."""
        ranks = np.ascontiguousarray(ranks, dtype=np.int32)
        out = np.zeros(max(1, ranks.shape[1]), np.int32)
        self._check(getattr(self.lib, self.prefix + "top_from_ranks")(ranks.shape[0], ranks.shape[1],
                                                              ranks.reshape(-1) if ranks.size else np.zeros(1, np.int32),
                                                              float(prop), out))
        return out[:ranks.shape[1]]

    def choose_middle(self, pvals, effects, prop):
        """choose_middle_effect_size(Pvals, Effects, prop): [ncon, ngenes] doubles (NaN = NA) -> double[ngenes]."""
        pvals = np.ascontiguousarray(pvals, dtype=np.float64)
        effects = np.ascontiguousarray(effects, dtype=np.float64)
        out = np.zeros(max(1, pvals.shape[1]), np.float64)
        self._check(getattr(self.lib, self.prefix + "choose_middle")(pvals.shape[0], pvals.shape[1], pvals.reshape(-1),
                                                             effects.reshape(-1), float(prop), out))
        return out[:pvals.shape[1]]


def load(kind="best"):
    """'port', 'ref', or 'best' (ref when present, else port)."""
    if kind == "best":
        try:
            return Checker("ref")
        except (FileNotFoundError, OSError):
            return Checker("port")
    return Checker(kind)


def have_ref():
    return os.path.exists(os.path.join(HERE, "_ref", "liboracle_ref.so"))
