"""GPU: scranb200_compute_min_rank_static, the min-rank entry point that never allocates device
memory, against the regular scranb200_compute_min_rank on the same inputs.

The static variant is for users of the library who cannot let it call cudaMalloc while it runs:
an application that has handed the whole GPU to its own memory pool, or a latency-sensitive
caller that cannot afford the device-wide stall of an allocation. Its scratch is one fixed array
inside the library that every call reuses, so this test makes two calls in a row with different
shapes (the second would see the first one's leftovers if the entry point did not reset what it
reads) and checks both, integer for integer, against the allocating entry point. The device
buffers here belong to the caller (torch), as they would in such an application.

This file sorts last on purpose: a run with -x has already been through every other test.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(sb, torch, x):
    from scran_b200._lib import check, lib, ptr
    ctx = sb.default_context()
    want = sb.compute_min_rank(x)                       # regular entry point (host buffers, context scratch)
    d_x = torch.from_numpy(np.ascontiguousarray(x.T)).cuda()   # column-major, owned by the caller
    d_out = torch.full((x.shape[0],), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    check(lib().scranb200_compute_min_rank_static(ctx.handle, ptr(d_x), x.shape[0], x.shape[1], ptr(d_out)))
    return d_out.cpu().numpy(), want


def test_static_min_rank_matches_regular_entry_point_twice():
    import torch

    import scran_b200 as sb
    rng = np.random.default_rng(31)
    # first call: columns longer than the shared-memory sort -> the static array holds the sort buffers
    a = np.round(rng.normal(size=(30000, 5)), 2)        # many ties
    a[::7, 1] = np.nan
    a[11, :] = np.nan                                   # a row without any value
    got, want = _both(sb, torch, a)
    assert np.array_equal(got, want)
    assert got[11] == np.iinfo(np.int32).max
    from oracle import rlevel
    assert np.array_equal(got.astype(np.int64), rlevel.compute_min_rank(a))   # and against the CPU restatement
    # second call right after, smaller and with other contents: must not see the first call's state
    b = rng.normal(size=(700, 9))
    b[5:60, 3] = 0.5
    got, want = _both(sb, torch, b)
    assert np.array_equal(got, want)
    # and the tall case once more, to cross the two shapes in the other direction
    got, want = _both(sb, torch, a[:20011])
    assert np.array_equal(got, want)


def test_static_min_rank_rejects_what_does_not_fit():
    import torch

    import scran_b200 as sb
    from scran_b200._lib import lib, ptr
    ctx = sb.default_context()
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    o = torch.zeros(8, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    assert lib().scranb200_compute_min_rank_static(ctx.handle, ptr(d), 1, 5000, ptr(o)) != 0
    assert "4096 columns" in lib().scranb200_last_error().decode()
    assert lib().scranb200_compute_min_rank_static(ctx.handle, ptr(d), 800000, 1, ptr(o)) != 0
    assert "too tall" in lib().scranb200_last_error().decode()
