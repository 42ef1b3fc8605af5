"""GPU: cbp_singletons_noalloc, the phase-1 entry point for callers that may not allocate device
memory (a worker thread inside an application whose pool allocator already owns all free HBM, or
per-node propagation where a cudaMalloc would stall every other thread's stream).

It must give what the allocating entry point cbp_singletons gives on the same input: the same two
counts and the same bounds.  Integer bounds are compared exactly and continuous ones within the
1e-9 relative / 1e-12 absolute tolerance the other parity tests use (tests/helpers.py).  The call
is made twice in a row on different inputs, because all calls share one fixed working array inside
the library and the second call starts from whatever the first one left there.  The file name keeps
this test last in collection order; it passes in a fresh process as well as after the other tests
(nothing here depends on state they leave behind, and the library clears what it needs itself).
"""
import numpy as np
import pytest
import torch

from cbc_b200 import capi
from cbc_b200 import instances as I
from helpers import assert_same_bounds

pytestmark = pytest.mark.gpu


def _device_copy(inst, lb, ub):
    """Everything the entry point reads or writes, as torch tensors on cuda:0 (the caller owns them)."""
    d = {
        "rowStart": torch.tensor(np.ascontiguousarray(inst.rowStart, np.int64), device="cuda"),
        "col": torch.tensor(np.ascontiguousarray(inst.col, np.int32), device="cuda"),
        "val": torch.tensor(np.ascontiguousarray(inst.val, np.float64), device="cuda"),
        "sense": torch.tensor(np.ascontiguousarray(inst.sense, np.uint8), device="cuda"),
        "rhs": torch.tensor(np.ascontiguousarray(inst.rhs, np.float64), device="cuda"),
        "range": torch.tensor(np.ascontiguousarray(inst.range, np.float64), device="cuda"),
        "colType": torch.tensor(np.ascontiguousarray(inst.colType, np.int8), device="cuda"),
        "lb": torch.tensor(np.ascontiguousarray(lb, np.float64), device="cuda"),
        "ub": torch.tensor(np.ascontiguousarray(ub, np.float64), device="cuda"),
    }
    torch.cuda.synchronize()
    return d


def _call(inst, d):
    return capi.singletons_noalloc(inst.m, inst.n, d["rowStart"].data_ptr(), d["col"].data_ptr(), d["val"].data_ptr(),
                                   d["sense"].data_ptr(), d["rhs"].data_ptr(), d["range"].data_ptr(),
                                   d["colType"].data_ptr(), inst.inf, d["lb"].data_ptr(), d["ub"].data_ptr())


def _instances():
    from test_gpu_parity import _with_singleton_rows

    a = _with_singleton_rows(I.mixed_bigm(m=3000, n=3000, seed=91), 900, 92)
    b = _with_singleton_rows(I.knapsack_assignment(m=1500, n=4000, seed=93), 700, 94)
    return a, b


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(gpu_ctx, oracle):
    torch.cuda.set_device(0)
    for inst in _instances():  # two calls in a row, second one on a different problem
        mat = capi.Matrix(gpu_ctx, inst)
        nt_ref, nf_ref, lb_ref, ub_ref = mat.singletons(inst.lb, inst.ub)
        mat.close()
        assert nt_ref > 50 and nf_ref > 5
        # ... and both are what the CPU oracle gives
        nt_o, nf_o, lb_o, ub_o = oracle.singletons(oracle.Problem(inst), inst.lb, inst.ub)
        assert (nt_ref, nf_ref) == (nt_o, nf_o)
        assert_same_bounds(inst.colType, lb_ref, ub_ref, lb_o, ub_o, "singletons vs oracle")
        d = _device_copy(inst, inst.lb, inst.ub)
        rc, nt, nf = _call(inst, d)
        assert rc == 0
        assert (nt, nf) == (nt_ref, nf_ref)
        assert_same_bounds(inst.colType, d["lb"].cpu().numpy(), d["ub"].cpu().numpy(), lb_ref, ub_ref, "noalloc")
        # calling it again on its own output changes nothing (the marks of the first call are gone)
        rc, nt2, nf2 = _call(inst, d)
        assert rc == 0 and nt2 == 0 and nf2 == 0
        assert_same_bounds(inst.colType, d["lb"].cpu().numpy(), d["ub"].cpu().numpy(), lb_ref, ub_ref, "noalloc again")


def test_noalloc_infeasible_leaves_bounds_and_oversize_is_refused(gpu_ctx):
    torch.cuda.set_device(0)
    inst, _ = _instances()
    j = int(np.flatnonzero(inst.colType == 1)[0])
    inst.sense = inst.sense.copy()
    inst.rhs = inst.rhs.copy()
    inst.col = inst.col.copy()
    inst.val = inst.val.copy()
    inst.col[-2], inst.val[-2], inst.sense[-1], inst.rhs[-1] = j, 1.0, ord("L"), inst.lb[j] - 1.0  # x_j <= lb_j - 1
    mat = capi.Matrix(gpu_ctx, inst)
    assert mat.singletons(inst.lb, inst.ub)[0] == -1
    mat.close()
    d = _device_copy(inst, inst.lb, inst.ub)
    rc, nt, nf = _call(inst, d)
    assert rc == 0 and nt == -1
    assert np.array_equal(d["lb"].cpu().numpy(), inst.lb) and np.array_equal(d["ub"].cpu().numpy(), inst.ub)
    # more columns than the fixed working memory holds: refused with CBP_ERR_CAPACITY, nothing is touched
    rc, _, _ = capi.singletons_noalloc(inst.m, capi.NOALLOC_MAX_COLS + 1, d["rowStart"].data_ptr(),
                                       d["col"].data_ptr(), d["val"].data_ptr(), d["sense"].data_ptr(),
                                       d["rhs"].data_ptr(), d["range"].data_ptr(), d["colType"].data_ptr(), inst.inf,
                                       d["lb"].data_ptr(), d["ub"].data_ptr())
    assert rc == capi.ERR_CAPACITY
    assert np.array_equal(d["lb"].cpu().numpy(), inst.lb)
    # a singleton row that names a column outside the problem: bad argument, bounds left as they came in
    bad_col = d["col"].clone()
    bad_col[-2] = inst.n + 3
    rc, _, _ = capi.singletons_noalloc(inst.m, inst.n, d["rowStart"].data_ptr(), bad_col.data_ptr(),
                                       d["val"].data_ptr(), d["sense"].data_ptr(), d["rhs"].data_ptr(),
                                       d["range"].data_ptr(), d["colType"].data_ptr(), inst.inf, d["lb"].data_ptr(),
                                       d["ub"].data_ptr())
    assert rc == -1
    assert np.array_equal(d["lb"].cpu().numpy(), inst.lb) and np.array_equal(d["ub"].cpu().numpy(), inst.ub)
