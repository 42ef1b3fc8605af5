"""GPU: seeded random instances of odd shapes against the CPU oracle — all five row senses, negative and
fractional coefficients, infinite and crossing-prone bounds, duplicate columns in a row, explicit zeros, empty
rows, rows of every length class (1, short, > 32, > 256, > 1024), continuous / binary / general-integer columns.
Single-vector fixpoint, truncated runs, phase 1, and the batch path on bound vectors derived from each instance."""
import numpy as np
import pytest

from cbc_b200 import capi
from cbc_b200 import instances as I
from helpers import assert_same_bounds, assert_same_result

pytestmark = pytest.mark.gpu


def random_instance(seed):
    g = np.random.default_rng(seed)
    n = int(g.integers(5, 400))
    m = int(g.integers(1, 300))
    kinds = g.integers(0, 6, size=m)
    lens = np.where(kinds == 0, 1, np.where(kinds == 1, g.integers(0, 4, size=m), g.integers(2, 33, size=m)))
    if seed % 3 == 0:  # some rows of the longer classes
        lens[g.integers(0, m, size=min(m, 3))] = g.choice([40, 300, 1100, 2500], size=min(m, 3))
    lens = np.minimum(lens, 4 * n)
    rowStart = np.zeros(m + 1, np.int64)
    np.cumsum(lens, out=rowStart[1:])
    nnz = int(rowStart[-1])
    col = g.integers(0, n, size=nnz).astype(np.int32)  # duplicates inside a row are allowed
    style = g.integers(0, 3)
    if style == 0:
        val = g.integers(-9, 10, size=nnz).astype(float)  # includes explicit zeros
    elif style == 1:
        val = np.round(g.normal(0, 5, size=nnz), 3)
    else:
        val = g.choice([1.0, 1.0, -1.0, 1000.0, -250.0, 0.5, 0.0], size=nnz)
    colType = g.choice([0, 1, 2], size=n, p=[0.3, 0.4, 0.3]).astype(np.int8)
    lb = np.where(colType == 1, 0.0, np.round(g.normal(-2, 4, size=n)))
    ub = np.where(colType == 1, 1.0, lb + g.integers(0, 30, size=n))
    x = np.where(colType == 0, g.uniform(0, 1, size=n) * (ub - lb) + lb, lb + np.floor(g.random(n) * (ub - lb + 1)))
    x = np.clip(x, lb, ub)
    lb[(g.random(n) < 0.08) & (colType != 1)] = -I.COIN_INF
    ub[(g.random(n) < 0.08) & (colType != 1)] = I.COIN_INF
    frac = (g.random(n) < 0.1) & (colType == 2) & (ub < I.COIN_INF)
    ub[frac] += 0.4  # integer columns with fractional bounds: R11
    act = np.bincount(np.repeat(np.arange(m), lens), weights=val * x[col], minlength=m)
    sense = g.choice(list(b"LGERN"), size=m, p=[0.35, 0.25, 0.15, 0.15, 0.1]).astype(np.uint8)
    slack = np.where(g.random(m) < 0.4, 0.0, g.integers(0, 8, size=m).astype(float))
    rhs = np.where(sense == ord("G"), act - slack, np.where(sense == ord("E"), act, act + slack))
    if seed % 4 == 1:  # make some instances infeasible
        k = int(g.integers(0, m))
        sense[k], rhs[k] = ord("L"), -1.0e6
    rng = slack + g.integers(0, 6, size=m).astype(float)  # 'R' rows: [rhs - range, rhs] contains the planted activity
    return I.Instance(f"fuzz{seed}", m, n, rowStart, col, val, sense, rhs, rng, colType, lb, ub, I.COIN_INF, None)


@pytest.mark.parametrize("seed", range(40))
def test_fuzz_single_vector(oracle, gpu_ctx, seed):
    inst = random_instance(seed)
    mat = capi.Matrix(gpu_ctx, inst)
    prob = oracle.Problem(inst)
    for kw in ({}, {"max_rounds": 1}, {"max_rounds": 3}):
        rg, lg, ug = mat.propagate(inst.lb, inst.ub, **kw)
        ro, lo, uo = oracle.propagate(prob, inst.lb, inst.ub, **kw)
        assert_same_result(rg, ro, f"seed {seed} {kw}")
        assert rg["roundChanged"] == ro["roundChanged"]
        assert_same_bounds(inst.colType, lg, ug, lo, uo, f"seed {seed} {kw}")
    nt, nf, lg, ug = mat.singletons(inst.lb, inst.ub)
    ont, onf, lo, uo = oracle.singletons(prob, inst.lb, inst.ub)
    assert (nt, nf) == (ont, onf)
    assert_same_bounds(inst.colType, lg, ug, lo, uo, f"seed {seed} singletons")
    mn, mx, im, ix = mat.row_activity(inst.lb, inst.ub)
    omn, omx, oim, oix = oracle.row_activity(prob, inst.lb, inst.ub)
    live = (inst.sense != ord("N"))
    assert np.array_equal(im[live], oim[live]) and np.array_equal(ix[live], oix[live])
    mat.close()


@pytest.mark.parametrize("seed", range(40, 60))
def test_fuzz_batch(oracle, gpu_ctx, seed):
    inst = random_instance(seed)
    g = np.random.default_rng(1000 + seed)
    B = int(g.integers(1, 70))
    # random bound diffs on random columns (values inside or slightly outside the base range)
    per = g.integers(0, 4, size=B)
    start = np.zeros(B + 1, np.int64)
    np.cumsum(per, out=start[1:])
    tot = int(start[-1])
    cols = g.integers(0, inst.n, size=tot)
    up = g.random(tot) < 0.5
    base = np.where(up, np.where(inst.ub[cols] >= inst.inf, 10.0, inst.ub[cols]), np.where(inst.lb[cols] <= -inst.inf, -10.0, inst.lb[cols]))
    vals = base + np.where(up, -1.0, 1.0) * g.integers(0, 4, size=tot) + np.where(g.random(tot) < 0.2, 0.5, 0.0)
    var = np.where(up, cols.astype(np.uint32) | I.UPPER_FLAG, cols.astype(np.uint32)).astype(np.uint32)
    mat = capi.Matrix(gpu_ctx, inst)
    for kw in ({}, {"max_rounds": 2}):
        res, us, uc, ul, uu = mat.propagate_batch_delta(inst.lb, inst.ub, start, var, vals, **kw)
        Lg, Ug = mat.batch_fetch_dense()
        ores, Lo, Uo = oracle.propagate_batch_delta(oracle.Problem(inst), inst.lb, inst.ub, start, var, vals, **kw)
        for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
            bad = np.flatnonzero(res[k] != ores[k])
            assert bad.size == 0, f"seed {seed} {kw}: {k} vectors {bad[:6]}: {res[k][bad[:6]]} vs {ores[k][bad[:6]]}"
        for b in range(B):
            assert_same_bounds(inst.colType, Lg[b], Ug[b], Lo[b], Uo[b], f"seed {seed} vector {b}")
        L0, U0 = I.expand_deltas(inst.lb, inst.ub, start, var, vals)
        res2, L2, U2 = mat.propagate_batch(L0, U0, **kw)
        for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
            assert np.array_equal(res2[k], res[k]), (seed, k)
        assert np.array_equal(L2, Lg) and np.array_equal(U2, Ug)
    mat.close()
