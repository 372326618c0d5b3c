"""GPU: malformed inputs are refused with a message instead of being read out of bounds, and the
handles stay safe to free in any order."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _upload(sb, nrow, ncol, p, i, x, p32=False):
    from scran_b200._lib import lib, ptr
    ctx = sb.default_context()
    h = C.c_void_p()
    p = np.ascontiguousarray(p, np.int32 if p32 else np.int64)
    i = np.ascontiguousarray(i, np.int32)
    x = np.ascontiguousarray(x, np.float64)
    fn = lib().scranb200_upload_csc32 if p32 else lib().scranb200_upload_csc
    rc = fn(ctx.handle, nrow, ncol, ptr(p), ptr(i), ptr(x), C.byref(h))
    msg = lib().scranb200_last_error().decode()
    if rc == 0:
        lib().scranb200_mat_free(h)
    return rc, msg


@pytest.mark.parametrize("p32", [False, True])
def test_column_pointers_are_validated(p32):
    import scran_b200 as sb
    i = np.array([0, 2, 1, 3, 0], np.int32)
    x = np.ones(5)
    assert _upload(sb, 4, 3, [0, 2, 4, 5], i, x, p32)[0] == 0
    rc, msg = _upload(sb, 4, 3, [1, 2, 4, 5], i, x, p32)
    assert rc != 0 and "start at 0" in msg
    rc, msg = _upload(sb, 4, 3, [0, 4, 2, 5], i, x, p32)
    assert rc != 0 and "non-decreasing" in msg
    rc, msg = _upload(sb, 2, 3, [0, 3, 4, 5], i, x, p32)
    assert rc != 0 and "more stored entries than the matrix has rows" in msg


def test_wrapped_device_arrays_are_validated():
    import torch

    import scran_b200 as sb
    from scran_b200._lib import lib, ptr
    ctx = sb.default_context()
    d_i = torch.tensor([0, 2, 1, 3, 0], dtype=torch.int32, device="cuda")
    d_x = torch.ones(5, dtype=torch.float64, device="cuda")
    for p, nnz, ok in [([0, 2, 4, 5], 5, True), ([0, 2, 4, 5], 4, False), ([0, 4, 2, 5], 5, False), ([1, 2, 4, 5], 5, False)]:
        d_p = torch.tensor(p, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        h = C.c_void_p()
        rc = lib().scranb200_wrap_device_csc(ctx.handle, 4, 3, nnz, ptr(d_p), ptr(d_i), ptr(d_x), C.byref(h))
        assert (rc == 0) == ok, (p, nnz)
        if rc == 0:
            lib().scranb200_mat_free(h)
        else:
            assert "column pointers" in lib().scranb200_last_error().decode()


def test_rows_out_of_order_are_refused():
    import scran_b200 as sb
    from scran_b200._lib import ScranB200Error
    m = sb.DeviceMatrix((4, 2, np.array([0, 2, 3]), np.array([2, 1, 0], np.int32), np.ones(3)))
    with pytest.raises(ScranB200Error, match="strictly increasing"):
        sb.compute_blocked_stats_none(m, np.zeros(2, np.int32), 1)
    m.free()


def test_context_may_be_destroyed_before_its_matrices_and_plans():
    import scran_b200 as sb
    from scran_b200._lib import ScranB200Error
    ctx = sb.Context(0)
    m = sb.DeviceMatrix(np.eye(5), ctx)
    plan = sb.ScorePlan([0, 1, 0, 1, 0], ctx=ctx)
    ctx.close()
    m.free()       # only deletes the handle: the context has already released the arrays
    plan.free()
    ctx2 = sb.Context(0)
    m2 = sb.DeviceMatrix(np.eye(5), ctx2)
    other = sb.default_context()
    from scran_b200._lib import lib, ptr
    mean = np.zeros((1, 5))
    var = np.zeros((1, 5))
    rc = lib().scranb200_blocked_stats(other.handle, m2.handle, ptr(np.zeros(5, np.int32)), 1, ptr(mean), ptr(var), None)
    assert rc != 0 and "another context" in lib().scranb200_last_error().decode()
    m2.free()
    ctx2.close()
