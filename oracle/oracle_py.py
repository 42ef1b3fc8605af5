"""ctypes wrapper for the CPU oracle (oracle/libcbp_oracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  Nothing in cbc_b200/ imports it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libcbp_oracle.so")

STOP_NAMES = {0: "NotRun", 1: "ReachedFixpoint", 2: "HitMaxRounds", 3: "HitTimeLimit", 4: "InfeasibleDetected"}


class OrcProblem(C.Structure):
    _fields_ = [
        ("m", C.c_int),
        ("n", C.c_int),
        ("rowStart", C.c_void_p),
        ("rowLen", C.c_void_p),
        ("col", C.c_void_p),
        ("val", C.c_void_p),
        ("sense", C.c_void_p),
        ("rhs", C.c_void_p),
        ("range", C.c_void_p),
        ("colType", C.c_void_p),
        ("inf", C.c_double),
    ]


class OrcResult(C.Structure):
    _fields_ = [
        ("feasible", C.c_int),
        ("rounds", C.c_int),
        ("nChanged", C.c_int),
        ("infeasibleRow", C.c_int),
        ("infeasibleCol", C.c_int),
        ("stopReason", C.c_int),
        ("seconds", C.c_double),
        ("minMargin", C.c_double),
    ]


ORC_RESULT_DTYPE = np.dtype(
    [
        ("feasible", "i4"),
        ("rounds", "i4"),
        ("nChanged", "i4"),
        ("infeasibleRow", "i4"),
        ("infeasibleCol", "i4"),
        ("stopReason", "i4"),
        ("seconds", "f8"),
        ("minMargin", "f8"),
    ],
    align=True,
)


def build():
    """Compile the oracle in place (gcc, see oracle/Makefile)."""
    subprocess.check_call(["make", "-s", "-C", _HERE])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_propagate.restype = C.c_int
        _lib.orc_propagate_gs.restype = C.c_int
        _lib.orc_singletons.restype = C.c_int
        _lib.orc_round.restype = C.c_int
        _lib.orc_propagate_batch.restype = C.c_int
        _lib.orc_propagate_batch_delta.restype = C.c_int
        _lib.orc_max_threads.restype = C.c_int
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


class Problem:
    """Holds numpy copies of one instance in the oracle's struct layout."""

    def __init__(self, inst):
        self.m, self.n = int(inst.m), int(inst.n)
        self.rowStart = np.ascontiguousarray(inst.rowStart, dtype=np.int64)
        self.col = np.ascontiguousarray(inst.col, dtype=np.int32)
        self.val = np.ascontiguousarray(inst.val, dtype=np.float64)
        self.sense = np.ascontiguousarray(inst.sense, dtype=np.uint8)
        self.rhs = np.ascontiguousarray(inst.rhs, dtype=np.float64)
        self.range = np.ascontiguousarray(inst.range, dtype=np.float64)
        self.colType = np.ascontiguousarray(inst.colType, dtype=np.int8)
        self.inf = float(inst.inf)
        self.s = OrcProblem(
            self.m, self.n, _p(self.rowStart), C.c_void_p(0), _p(self.col), _p(self.val),
            _p(self.sense), _p(self.rhs), _p(self.range), _p(self.colType), self.inf,
        )

    @property
    def nnz(self):
        return int(self.rowStart[-1])


def _res(r):
    return {
        "feasible": bool(r.feasible), "rounds": r.rounds, "nChanged": r.nChanged,
        "infeasibleRow": r.infeasibleRow, "infeasibleCol": r.infeasibleCol,
        "stopReason": r.stopReason, "seconds": r.seconds, "minMargin": r.minMargin,
    }


def row_activity(prob, lb, ub):
    m = prob.m
    mn, mx = np.zeros(m), np.zeros(m)
    im, ix = np.zeros(m, np.int32), np.zeros(m, np.int32)
    lb = np.ascontiguousarray(lb, np.float64)
    ub = np.ascontiguousarray(ub, np.float64)
    lib().orc_row_activity(C.byref(prob.s), _p(lb), _p(ub), _p(mn), _p(mx), _p(im), _p(ix))
    return mn, mx, im, ix


def singletons(prob, lb, ub, tol=1e-7):
    lb = np.array(lb, np.float64)
    ub = np.array(ub, np.float64)
    nf = C.c_int(0)
    nt = lib().orc_singletons(C.byref(prob.s), _p(lb), _p(ub), C.c_double(tol), C.byref(nf))
    return nt, nf.value, lb, ub


def round_rows(prob, lb, ub, tol=1e-7):
    """Row phase of one round: (newLb, newUb, first violated row or -1)."""
    lb = np.ascontiguousarray(lb, np.float64)
    ub = np.ascontiguousarray(ub, np.float64)
    nlb, nub = np.empty_like(lb), np.empty_like(ub)
    changed = np.zeros(prob.n, np.uint8)
    lib().orc_round_rows.restype = C.c_int
    bad = lib().orc_round_rows(C.byref(prob.s), _p(lb), _p(ub), C.c_double(tol), _p(nlb), _p(nub), _p(changed), None)
    return nlb, nub, bad


def propagate(prob, lb, ub, tol=1e-7, max_rounds=0, time_limit=1e100, gauss_seidel=False, round_cap=4096):
    lb = np.array(lb, np.float64)
    ub = np.array(ub, np.float64)
    r = OrcResult()
    rc_arr = np.full(round_cap, -1, np.int32)
    if gauss_seidel:
        rc = lib().orc_propagate_gs(C.byref(prob.s), _p(lb), _p(ub), C.c_double(tol), C.c_int(max_rounds), C.byref(r))
    else:
        rc = lib().orc_propagate(
            C.byref(prob.s), _p(lb), _p(ub), C.c_double(tol), C.c_int(max_rounds),
            C.c_double(time_limit), C.byref(r), _p(rc_arr), C.c_int(round_cap),
        )
    assert rc == 0
    out = _res(r)
    out["roundChanged"] = [int(x) for x in rc_arr[: r.rounds] if x >= 0]
    return out, lb, ub


def propagate_batch(prob, lb, ub, tol=1e-7, max_rounds=0, time_limit=1e100, threads=1):
    """lb/ub: (B, n) arrays; returns (structured results, lb, ub)."""
    lb = np.array(lb, np.float64, order="C")
    ub = np.array(ub, np.float64, order="C")
    B = lb.shape[0]
    res = np.zeros(B, ORC_RESULT_DTYPE)
    rc = lib().orc_propagate_batch(
        C.byref(prob.s), C.c_int(B), _p(lb), _p(ub), C.c_double(tol), C.c_int(max_rounds),
        C.c_double(time_limit), _p(res), C.c_int(threads),
    )
    assert rc == 0
    return res, lb, ub


def propagate_batch_delta(prob, base_lb, base_ub, delta_start, delta_var, delta_bound, tol=1e-7,
                          max_rounds=0, time_limit=1e100, threads=1, want_bounds=True):
    base_lb = np.ascontiguousarray(base_lb, np.float64)
    base_ub = np.ascontiguousarray(base_ub, np.float64)
    delta_start = np.ascontiguousarray(delta_start, np.int64)
    delta_var = np.ascontiguousarray(delta_var, np.uint32)
    delta_bound = np.ascontiguousarray(delta_bound, np.float64)
    B = len(delta_start) - 1
    res = np.zeros(B, ORC_RESULT_DTYPE)
    lbo = np.empty((B, prob.n)) if want_bounds else None
    ubo = np.empty((B, prob.n)) if want_bounds else None
    rc = lib().orc_propagate_batch_delta(
        C.byref(prob.s), _p(base_lb), _p(base_ub), C.c_int(B), _p(delta_start), _p(delta_var),
        _p(delta_bound), C.c_double(tol), C.c_int(max_rounds), C.c_double(time_limit), _p(res),
        _p(lbo), _p(ubo), C.c_int(threads),
    )
    if rc == -2:
        raise ValueError("a bound diff names a column outside the matrix")
    assert rc == 0
    return res, lbo, ubo


def max_threads():
    return lib().orc_max_threads()
