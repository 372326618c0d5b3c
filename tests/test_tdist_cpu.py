"""CPU: the log tails of Student's t that the device t-test layer evaluates (scran_b200/csrc/tdist.cuh,
standing in for stats::pt(log.p=TRUE) at R/pairwiseTTests.R:439-461). The header is plain C++ as well
as CUDA, so it is compiled here with g++ into a scratch library - test scaffolding, nothing the product
loads - and checked against the regularised incomplete beta function in 50-digit arithmetic (mpmath) and
against scipy's t distribution, from the centre to log p < -1e5, for fractional and very large degrees of
freedom."""
import ctypes
import os
import subprocess
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def log_pt():
    d = tempfile.mkdtemp(prefix="tdist_")
    src = os.path.join(d, "w.cpp")
    with open(src, "w") as f:
        f.write('#include "%s"\n' % os.path.join(ROOT, "scran_b200", "csrc", "tdist.cuh"))
        f.write('extern "C" void log_pt_many(const double* t, const double* df, int n, int lower, double* out) {\n'
                '  for (int i = 0; i < n; ++i) out[i] = sb::log_pt(t[i], df[i], lower != 0);\n}\n')
    so = os.path.join(d, "w.so")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC", src, "-o", so])
    lib = ctypes.CDLL(so)

    def f(t, df, lower=True):
        t = np.ascontiguousarray(np.broadcast_to(np.asarray(t, np.float64), np.broadcast(t, df).shape))
        df = np.ascontiguousarray(np.broadcast_to(np.asarray(df, np.float64), t.shape))
        out = np.zeros(t.shape)
        lib.log_pt_many(t.ctypes.data_as(ctypes.c_void_p), df.ctypes.data_as(ctypes.c_void_p), t.size, int(lower),
                        out.ctypes.data_as(ctypes.c_void_p))
        return out
    return f


def _ref(t, df, lower):
    import mpmath as mp
    with mp.workdps(50):
        t, df = mp.mpf(t), mp.mpf(df)
        try:
            tail = mp.betainc(df / 2, mp.mpf(1) / 2, 0, df / (df + t * t), regularized=True) / 2
        except ValueError:
            # mpmath's hypergeometric series gives up on a few (t, df): evaluate the continued fraction of the
            # incomplete beta function in 50-digit arithmetic instead (Lentz; same identity, independent precision)
            a, b, x = df / 2, mp.mpf(1) / 2, df / (df + t * t)
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
        small = (t <= 0) == lower
        return float(mp.log(tail)) if small else float(mp.log1p(-tail))


def test_log_pt_against_incomplete_beta(log_pt):
    ts = [0, 1e-8, 0.01, 0.5, 1, 1.5, 1.7, 1.8, 2, 3, 5, 10, 37, 80, 300, 2000, 1e5]
    dfs = [0.3, 1, 2.5, 7.77, 30, 123.4, 1000.5, 6500, 52000.3]
    worst = 0.0
    for df in dfs:
        for t in ts:
            for sign in (1, -1):
                for lower in (True, False):
                    got = float(log_pt(sign * t, df, lower).ravel()[0])
                    want = _ref(sign * t, df, lower)
                    err = abs(got - want) / max(abs(want), 1e-300)
                    worst = max(worst, err)
                    assert err < 1e-11, (t, df, lower, got, want)
    assert worst < 1e-11


def test_log_pt_against_scipy_and_limits(log_pt):
    from scipy.stats import t as tdist
    rng = np.random.default_rng(3)
    t = rng.normal(0, 3, 4000)
    df = np.exp(rng.uniform(np.log(0.5), np.log(1e5), 4000))
    np.testing.assert_allclose(log_pt(t, df, True), tdist.logcdf(t, df), rtol=2e-11, atol=1e-15)
    np.testing.assert_allclose(log_pt(t, df, False), tdist.logsf(t, df), rtol=2e-11, atol=1e-15)
    assert np.isnan(log_pt(np.nan, 5.0)) and np.isnan(log_pt(1.0, np.nan)) and np.isnan(log_pt(1.0, 0.0))
    assert log_pt(-np.inf, 4.0, True) == -np.inf and log_pt(np.inf, 4.0, True) == 0.0
    assert log_pt(np.inf, 4.0, False) == -np.inf
    # the normal limit
    from scipy.special import log_ndtr
    np.testing.assert_allclose(log_pt(np.array([-40.0, -3.0, 0.0, 2.0]), np.inf, True),
                               log_ndtr(np.array([-40.0, -3.0, 0.0, 2.0])), rtol=1e-12)
