"""Shared comparison helpers for the parity tests."""
import json
import os

import numpy as np

RTOL, ATOL = 1e-9, 1e-12  # north_star tolerance for continuous bounds

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_cases():
    with open(os.path.join(GOLDEN, "micro_instances.json")) as f:
        return json.load(f)["cases"]


def assert_same_bounds(colType, lb_a, ub_a, lb_b, ub_b, what=""):
    """Integer columns exact, continuous within RTOL/ATOL, fixed sets equal."""
    colType = np.asarray(colType)
    isint = colType != 0
    for name, a, b in (("lb", lb_a, lb_b), ("ub", ub_a, ub_b)):
        a, b = np.asarray(a), np.asarray(b)
        bad = np.flatnonzero(isint & (a != b))
        assert bad.size == 0, f"{what}: integer {name} differs at columns {bad[:8]}: {a[bad[:8]]} vs {b[bad[:8]]}"
        cont = ~isint
        diff = np.abs(a[cont] - b[cont])
        lim = ATOL + RTOL * np.maximum(np.abs(a[cont]), np.abs(b[cont]))
        same = (a[cont] == b[cont]) | (diff <= lim)
        assert same.all(), f"{what}: continuous {name} off by {diff[~same].max()} at {np.flatnonzero(cont)[~same][:8]}"
    fa = np.asarray(lb_a) == np.asarray(ub_a)
    fb = np.asarray(lb_b) == np.asarray(ub_b)
    # fixed continuous columns may differ in the last bits only if both are fixed
    assert np.array_equal(fa[isint], fb[isint]), f"{what}: set of fixed integer columns differs"


def assert_same_result(ra, rb, what=""):
    for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
        assert int(ra[k]) == int(rb[k]), f"{what}: {k}: {ra[k]} vs {rb[k]}"
