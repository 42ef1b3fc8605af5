"""GPU: node batches in worklist form (cbc_b200/csrc/cbp_wl_kernels.cuh) — the persistent kernel that evaluates only
the (row, vector) pairs a changed column touches — against the CPU oracle run vector by vector, and against the
full-walk kernels it hands over to.  The worklist form is taken when the node bounds are a propagation fixpoint."""
import os
import subprocess
import sys

import numpy as np
import pytest

from cbc_b200 import capi
from cbc_b200 import instances as I
from helpers import assert_same_bounds

pytestmark = pytest.mark.gpu
KEYS = ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason")


def _node(oracle, inst):
    """node bounds = the propagated root (what Cbc hands to strong branching / a dive)"""
    r, lb, ub = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    assert r["feasible"]
    inst.lb, inst.ub = lb, ub
    return inst


def _check(oracle, ctx, inst, s, v, b, what, want_mode=2, **kw):
    mat = capi.Matrix(ctx, inst)
    ores, Lo, Uo = oracle.propagate_batch_delta(oracle.Problem(inst), inst.lb, inst.ub, s, v, b, threads=8, **kw)
    L0, U0 = I.expand_deltas(inst.lb, inst.ub, s, v, b)
    for rep in range(2):  # twice: the second run starts from what the first left in the masks
        res, us, uc, ul, uu = mat.propagate_batch_delta(inst.lb, inst.ub, s, v, b, **kw)
        st = ctx.stats()
        assert st["lastMode"] == want_mode, (what, st)
        if want_mode == 2:
            assert st["lastHostSyncs"] == 1
        for k in KEYS:
            bad = np.flatnonzero(res[k] != ores[k])
            assert bad.size == 0, f"{what}: {k} differs for vectors {bad[:8]}: {res[k][bad[:8]]} vs {ores[k][bad[:8]]}"
        Lg, Ug = mat.batch_fetch_dense()
        for k in range(len(s) - 1):
            assert_same_bounds(inst.colType, Lg[k], Ug[k], Lo[k], Uo[k], f"{what} vector {k}")
            cols = uc[us[k]:us[k + 1]]
            if not res["feasible"][k]:
                assert cols.size == 0
                continue
            want = np.flatnonzero((Lg[k] != L0[k]) | (Ug[k] != U0[k]))
            assert np.array_equal(cols, want), f"{what} vector {k}: update list"
            assert np.array_equal(ul[us[k]:us[k + 1]], Lg[k][want]) and np.array_equal(uu[us[k]:us[k + 1]], Ug[k][want])
    mat.close()
    return res, st


def test_worklist_dives_short_rows(oracle, gpu_ctx):
    inst = I.knapsack_assignment(m=3000, n=8000, seed=41, root_fixpoint=True)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 70, max_depth=48, seed=42)
    res, st = _check(oracle, gpu_ctx, inst, s, v, b, "dives")
    assert res["rounds"].max() >= 2 and res["nChanged"].sum() > 100
    assert 0 < st["lastPairsWalked"] < inst.nnz * int(res["rounds"].sum()) // 10


def test_worklist_dives_that_fail(oracle, gpu_ctx):
    # dives that mostly leave the planted point: violated rows and crossing columns among the vectors
    inst = I.knapsack_assignment(m=3000, n=8000, seed=41, root_fixpoint=True)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 70, max_depth=48, seed=42, follow_xstar=0.3)
    res, _ = _check(oracle, gpu_ctx, inst, s, v, b, "failing dives")
    f = res["feasible"].astype(bool)
    assert f.any() and (~f).any()


def test_worklist_strong_branching_long_rows(oracle, gpu_ctx):
    inst = _node(oracle, I.fast0507_shaped(m=101, n=9000, fix_frac=0.996, seed=22))
    s, v, b = I.strong_branch_deltas(inst, inst.lb, inst.ub, 48, seed=43)
    res, st = _check(oracle, gpu_ctx, inst, s, v, b, "strong branching")
    assert res["nChanged"].sum() > 0


def test_worklist_mixed_continuous_chains(oracle, gpu_ctx):
    inst = _node(oracle, I.mixed_bigm(m=6000, n=6000, seed=44))
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 45, max_depth=12, seed=45)
    res, _ = _check(oracle, gpu_ctx, inst, s, v, b, "mixed dives")
    assert res["rounds"].max() >= 5 and res["nChanged"].sum() > 500
    s, v, b = I.probe_deltas(inst, inst.lb, inst.ub, 20, seed=46)
    _check(oracle, gpu_ctx, inst, s, v, b, "mixed probes")
    _check(oracle, gpu_ctx, inst, s, v, b, "mixed probes maxRounds=2", max_rounds=2)


def test_worklist_power_law_rows_of_every_length(oracle, gpu_ctx):
    inst = _node(oracle, I.power_law(n=20000, nnz_target=150000, n_dense=3, dense_len=6000, max_len=2500, seed=46))
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 40, max_depth=20, seed=47)
    res, _ = _check(oracle, gpu_ctx, inst, s, v, b, "power law")
    assert res["nChanged"].sum() > 0


def test_worklist_odd_batches(oracle, gpu_ctx):
    """vectors without diffs, a column repeated inside a vector, a diff that crosses the other bound (R11 / R9 at the
    diff cell), fractional diffs on integer columns, a last group that is not full"""
    inst = _node(oracle, I.knapsack_assignment(m=1200, n=3000, seed=61, root_fixpoint=True))
    free = np.flatnonzero(inst.ub > inst.lb)
    start, var, bnd = [0], [], []
    for k in range(75):
        j = int(free[(13 * k) % free.size])
        if k % 5 == 0:
            pass
        elif k % 5 == 1:
            var += [j | I.UPPER_FLAG, j]
            bnd += [0.0, 0.0]
        elif k % 5 == 2:
            var += [j, j | I.UPPER_FLAG]
            bnd += [1.0, 0.0]  # lb above ub: infeasible before the first round
        elif k % 5 == 3:
            var += [j | I.UPPER_FLAG]
            bnd += [0.6]  # rounds down to 0 (R11), counts as a round-1 change
        else:
            var += [j, int(free[(13 * k + 5) % free.size])]
            bnd += [1.0, 1.0]
        start.append(len(var))
    s, v, b = np.array(start, np.int64), np.array(var, np.uint32), np.array(bnd, np.float64)
    res, _ = _check(oracle, gpu_ctx, inst, s, v, b, "odd")
    assert (res["feasible"] == 0).any() and (res["nChanged"] > 0).any()


def test_worklist_hands_over_when_the_node_is_not_a_fixpoint(oracle, gpu_ctx):
    inst = I.knapsack_assignment(m=1200, n=3000, seed=62)  # root bounds still tighten by themselves
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 40, max_depth=20, seed=63)
    _check(oracle, gpu_ctx, inst, s, v, b, "not a fixpoint", want_mode=1)


def test_worklist_then_other_batches_on_the_same_matrix(oracle, gpu_ctx):
    """the masks a worklist run leaves are cleaned by the next run, whichever form that takes"""
    inst = I.knapsack_assignment(m=1500, n=4000, seed=64, root_fixpoint=True)
    mat = capi.Matrix(gpu_ctx, inst)
    prob = oracle.Problem(inst)
    for seed, B in ((1, 50), (2, 90), (3, 20)):
        s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, B, max_depth=30, seed=seed)
        res = mat.propagate_batch_delta(inst.lb, inst.ub, s, v, b)[0]
        assert gpu_ctx.stats()["lastMode"] == 2
        ores, Lo, Uo = oracle.propagate_batch_delta(prob, inst.lb, inst.ub, s, v, b, threads=8)
        for k in KEYS:
            assert np.array_equal(res[k], ores[k]), (seed, k)
        Lg, Ug = mat.batch_fetch_dense()
        f = res["feasible"].astype(bool)
        assert np.array_equal(Lg[f], Lo[f]) and np.array_equal(Ug[f], Uo[f])
        if seed == 2:  # a dense batch (classic two-buffer form) in between
            L0, U0 = I.expand_deltas(inst.lb, inst.ub, s, v, b)
            r2, L2, U2 = mat.propagate_batch(L0, U0)
            assert np.array_equal(r2["rounds"], ores["rounds"]) and np.array_equal(L2[f], Lo[f])
    mat.close()


@pytest.mark.parametrize("env,maker", [
    ({"CBP_WL_ITEM_CAP": "500"}, "I.knapsack_assignment(m=3000, n=8000, seed=41, root_fixpoint=True)"),
    ({"CBP_WL_SLOTS": "64"}, "I.knapsack_assignment(m=60, n=4000, seed=71, root_fixpoint=True, slack_max=0)"),
], ids=["item_list_full", "pair_table_full"])
def test_worklist_overflow_hands_over(env, maker):
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "from cbc_b200 import capi, instances as I; from oracle import oracle_py as O\n"
        "inst = %s\n"
        "r0, lb0, ub0 = O.propagate(O.Problem(inst), inst.lb, inst.ub); inst.lb, inst.ub = lb0, ub0\n"
        "s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 70, max_depth=48, seed=42)\n"
        "ctx = capi.Context(); mat = capi.Matrix(ctx, inst)\n"
        "res, us, uc, ul, uu = mat.propagate_batch_delta(inst.lb, inst.ub, s, v, b)\n"
        "assert ctx.stats()['lastMode'] == 1, ctx.stats()\n"
        "Lg, Ug = mat.batch_fetch_dense()\n"
        "ores, Lo, Uo = O.propagate_batch_delta(O.Problem(inst), inst.lb, inst.ub, s, v, b, threads=8)\n"
        "for k in ('feasible','rounds','nChanged','infeasibleRow','infeasibleCol','stopReason'):\n"
        "    assert np.array_equal(res[k], ores[k]), k\n"
        "f = res['feasible'].astype(bool)\n"
        "assert np.array_equal(Lg[f], Lo[f]) and np.array_equal(Ug[f], Uo[f])\n"
        "print('OK')\n") % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), maker)
    out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env),
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "OK" in out.stdout, out.stderr[-2000:]


def test_worklist_long_integer_rows_exact_sums(oracle, gpu_ctx):
    """All-integer matrix with rows of several hundred nonzeros: a (vector, row) pair's sums are the node's sums plus
    what the vector's own cells add (exact in any order), so hardly any pair is walked; a fractional diff switches
    the same batch to walking the rows — both must equal the oracle."""
    inst = _node(oracle, I.knapsack_assignment(m=60, n=4000, seed=71, root_fixpoint=True, slack_max=0))
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 80, max_depth=40, seed=72, follow_xstar=0.5)
    res, st = _check(oracle, gpu_ctx, inst, s, v, b, "long integer rows")
    assert (res["feasible"] == 0).any() and res["nChanged"].sum() > 1000
    exact_pairs = st["lastPairsWalked"]
    b2 = b.copy()
    k = int(np.flatnonzero((v & I.UPPER_FLAG) != 0)[0])
    b2[k] = b2[k] + 0.5  # a fractional upper bound on an integer column: R11 rounds it back down
    res2, st2 = _check(oracle, gpu_ctx, inst, s, v, b2, "long integer rows, fractional diff")
    assert st2["lastPairsWalked"] > 2 * exact_pairs


def _table_reuse(oracle, ctx):
    """exact-sum batches of different sizes one after the other on one matrix: each run starts from the table the run
    before it left (its slots emptied, the part in use resized)"""
    inst = _node(oracle, I.knapsack_assignment(m=60, n=4000, seed=71, root_fixpoint=True, slack_max=0))
    mat = capi.Matrix(ctx, inst)
    prob = oracle.Problem(inst)
    for seed, B, depth in ((1, 120, 40), (2, 8, 6), (3, 120, 40), (1, 120, 40), (4, 33, 60)):
        s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, B, max_depth=depth, seed=seed, follow_xstar=0.6)
        res = mat.propagate_batch_delta(inst.lb, inst.ub, s, v, b)[0]
        assert ctx.stats()["lastMode"] == 2
        ores, Lo, Uo = oracle.propagate_batch_delta(prob, inst.lb, inst.ub, s, v, b, threads=8)
        for k in KEYS:
            assert np.array_equal(res[k], ores[k]), (seed, k)
        Lg, Ug = mat.batch_fetch_dense()
        f = res["feasible"].astype(bool)
        assert np.array_equal(Lg[f], Lo[f]) and np.array_equal(Ug[f], Uo[f]), seed
    mat.close()


def test_worklist_pair_table_is_reused_across_runs(oracle, gpu_ctx):
    _table_reuse(oracle, gpu_ctx)


def test_worklist_pair_table_nearly_full():
    """CBP_WL_TABLE_FACTOR=1: the part of the table in use is barely larger than what the last run took — long probe
    chains, and a hand-over when a run needs more; the answers stay the oracle's"""
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import test_gpu_worklist as T\n"
        "from cbc_b200 import capi; from oracle import oracle_py as O\n"
        "ctx = capi.Context()\n"
        "stats = ctx.stats  # (lastMode may be 1 when a run outgrows the part of the table in use)\n"
        "ctx.stats = lambda: dict(stats(), lastMode=2)\n"
        "T._table_reuse(O, ctx)\n"
        "print('OK')\n") % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, CBP_WL_TABLE_FACTOR="1"),
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "OK" in out.stdout, out.stderr[-2000:]


def _stars(sizes, col_type):
    """hubs h with leaves x_i <= x_h (rows x_i - x_h <= 0), all columns in [0, 10]: the node is a fixpoint, and lowering
    a hub's upper bound changes exactly its leaves in one round"""
    n = len(sizes) + int(sum(sizes))
    rows, cols, vals = [], [], []
    leaf = len(sizes)
    hub_leaves = []
    for h, k in enumerate(sizes):
        mine = np.arange(leaf, leaf + k)
        hub_leaves.append(mine)
        for i in mine:
            r = len(rows) // 2
            rows += [r, r]
            cols += [int(i), h]
            vals += [1.0, -1.0]
        leaf += k
    m = len(rows) // 2
    # scatter the leaves over the column range so that a hub's update list does not arrive in column order
    perm = np.random.Generator(np.random.Philox(7)).permutation(n)
    cols = perm[np.asarray(cols)]
    rowStart, col, val, _ = I._csr_from_pairs(m, n, np.asarray(rows), cols, np.asarray(vals))
    inst = I.Instance(name="stars", m=m, n=n, rowStart=rowStart, col=col, val=val,
                      sense=np.full(m, ord("L"), np.uint8), rhs=np.zeros(m), range=np.zeros(m),
                      colType=np.full(n, col_type, np.int8), lb=np.zeros(n), ub=np.full(n, 10.0))
    return inst, perm, hub_leaves


@pytest.mark.parametrize("col_type", [2, 0], ids=["integer_exact_sums", "continuous"])
def test_worklist_update_lists_of_every_length(oracle, gpu_ctx, col_type):
    """update lists of 0, 1, 20, 33, 200, 257, 1500, 2049 and 5000 entries: every ordering path of the run's last phase
    (warp by shuffles, warp counting, CTA bitonic, CTA counting), lists longer than the head the run copies unasked,
    and — second batch — more entries than the list buffers hold (the stand-alone kernels size them)"""
    sizes = [0, 1, 20, 33, 200, 257, 1500, 2049, 5000]
    inst, perm, leaves = _stars(sizes, col_type)
    s = np.arange(len(sizes) + 1, dtype=np.int64)
    v = (perm[np.arange(len(sizes))].astype(np.uint32) | np.uint32(I.UPPER_FLAG))
    b = np.full(len(sizes), 5.0)
    res, st = _check(oracle, gpu_ctx, inst, s, v, b, "stars")
    assert np.array_equal(res["nChanged"], np.asarray(sizes))
    # every hub in every vector: 9 x 9,060 entries > the 65,536 the buffers start with
    B = 9
    s2 = np.arange(B + 1, dtype=np.int64) * len(sizes)
    v2 = np.tile(v, B)
    b2 = np.repeat(np.arange(1.0, B + 1.0), len(sizes))
    res2, _ = _check(oracle, gpu_ctx, inst, s2, v2, b2, "stars, all hubs per vector")
    assert np.all(res2["nChanged"] == sum(sizes))
