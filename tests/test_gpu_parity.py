"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the
same seeded inputs, against the golden micro-instances, and size-independent
properties at larger sizes."""
import numpy as np
import pytest

from cbc_b200 import capi
from cbc_b200 import instances as I
from helpers import assert_same_bounds, assert_same_result, golden_cases

pytestmark = pytest.mark.gpu


def _both(oracle, ctx, inst, lb=None, ub=None, **kw):
    lb = inst.lb if lb is None else lb
    ub = inst.ub if ub is None else ub
    mat = capi.Matrix(ctx, inst)
    rg, lg, ug = mat.propagate(lb, ub, **kw)
    mat.close()
    ro, lo, uo = oracle.propagate(oracle.Problem(inst), lb, ub, **kw)
    return (rg, lg, ug), (ro, lo, uo)


@pytest.mark.parametrize("case", golden_cases(), ids=lambda c: c["name"])
def test_golden_micro(gpu_ctx, case):
    inst = I.micro(case["instance"])
    mat = capi.Matrix(gpu_ctx, inst)
    r, lb, ub = mat.propagate(case["lb_in"], case["ub_in"])
    assert r["feasible"] == case["feasible"]
    assert r["rounds"] == case["rounds"]
    assert r["nChanged"] == case["nChanged"]
    assert r["infeasibleRow"] == case.get("infeasibleRow", -1)
    assert r["infeasibleCol"] == case.get("infeasibleCol", -1)
    assert np.array_equal(lb, np.array(case["lb"], float))
    assert np.array_equal(ub, np.array(case["ub"], float))
    if "singletons" in case:
        nt, nf, _, _ = mat.singletons(case["lb_in"], case["ub_in"])
        assert nt == case["singletons"]
    # batch path, same vector three times
    res, L, U = mat.propagate_batch(np.tile(case["lb_in"], (3, 1)), np.tile(case["ub_in"], (3, 1)))
    for b in range(3):
        assert bool(res["feasible"][b]) == case["feasible"] and res["rounds"][b] == case["rounds"]
        assert res["infeasibleRow"][b] == case.get("infeasibleRow", -1)
        assert np.array_equal(L[b], np.array(case["lb"], float)) and np.array_equal(U[b], np.array(case["ub"], float))
    mat.close()


SV_CASES = {
    "setcover": lambda: I.set_cover(m=2000, n=4000, density=0.01, fix_frac=0.93, seed=21),
    "fast0507": lambda: I.fast0507_shaped(m=101, n=9000, fix_frac=0.985, seed=22),
    "knapassign": lambda: I.knapsack_assignment(m=4000, n=10000, seed=23),
    "mixed": lambda: I.mixed_bigm(m=20000, n=20000, seed=24),
    "mixed_loose": lambda: I.mixed_bigm(m=5000, n=3000, avg_len=20, fix_frac=0.2, seed=25),
    "powerlaw": lambda: I.power_law(n=30000, nnz_target=200000, n_dense=3, dense_len=9000, max_len=3000, seed=26),
    "powerlaw_bad": lambda: I.power_law(n=30000, nnz_target=200000, n_dense=3, dense_len=9000, max_len=3000,
                                        seed=27, infeasible=True),
}


@pytest.mark.parametrize("name", list(SV_CASES))
def test_single_vector_fixpoint(oracle, gpu_ctx, name):
    inst = SV_CASES[name]()
    (rg, lg, ug), (ro, lo, uo) = _both(oracle, gpu_ctx, inst)
    assert_same_result(rg, ro, name)
    assert rg["roundChanged"] == ro["roundChanged"]
    assert_same_bounds(inst.colType, lg, ug, lo, uo, name)
    if inst.xstar is not None:
        assert rg["feasible"] and inst.contains(lg, ug)


def test_max_rounds_and_time_limit(oracle, gpu_ctx):
    inst = I.mixed_bigm(m=20000, n=20000, seed=24)
    for mr in (1, 2, 5):
        (rg, lg, ug), (ro, lo, uo) = _both(oracle, gpu_ctx, inst, max_rounds=mr)
        assert_same_result(rg, ro, f"maxRounds={mr}")
        assert rg["stopReason"] == 2
        assert_same_bounds(inst.colType, lg, ug, lo, uo, f"maxRounds={mr}")
    (rg, lg, ug), (ro, lo, uo) = _both(oracle, gpu_ctx, inst, time_limit=0.0)
    assert_same_result(rg, ro, "timeLimit=0")
    assert rg["stopReason"] == 3 and np.array_equal(lg, inst.lb)


def test_row_activity_counters_exact(oracle, gpu_ctx):
    inst = I.mixed_bigm(m=20000, n=20000, seed=24)
    lb = inst.lb.copy()
    lb[::7] = -inst.inf  # plenty of infinite contributions on both sides
    mat = capi.Matrix(gpu_ctx, inst)
    mn, mx, im, ix = mat.row_activity(lb, inst.ub)
    mat.close()
    omn, omx, oim, oix = oracle.row_activity(oracle.Problem(inst), lb, inst.ub)
    assert np.array_equal(im, oim) and np.array_equal(ix, oix)
    assert np.allclose(mn, omn, rtol=1e-9, atol=1e-12) and np.allclose(mx, omx, rtol=1e-9, atol=1e-12)


def _with_singleton_rows(inst, k, seed):
    """Append k one-nonzero rows that the planted point satisfies (some tight, some 'E', some with zeros)."""
    g = np.random.default_rng(seed)
    cols = g.choice(inst.n, size=k, replace=False)
    a = g.integers(1, 9, size=k).astype(float) * np.where(g.random(k) < 0.4, -1.0, 1.0)
    act = a * inst.xstar[cols]
    kind = g.integers(0, 3, size=k)  # 0: <=, 1: >=, 2: ==
    slack = np.floor(g.random(k) * 3.0)
    sense = np.where(kind == 0, ord("L"), np.where(kind == 1, ord("G"), ord("E"))).astype(np.uint8)
    rhs = np.where(kind == 0, act + slack, np.where(kind == 1, act - slack, act))
    # every appended row also carries an explicit zero entry that must be ignored
    new_cols = np.stack([cols, (cols + 1) % inst.n], axis=1).ravel().astype(np.int32)
    new_vals = np.stack([a, np.zeros(k)], axis=1).ravel()
    rowStart = np.concatenate([inst.rowStart, inst.rowStart[-1] + 2 * np.arange(1, k + 1)])
    return I.Instance(inst.name + "+sing", inst.m + k, inst.n, rowStart, np.concatenate([inst.col, new_cols]),
                      np.concatenate([inst.val, new_vals]), np.concatenate([inst.sense, sense]),
                      np.concatenate([inst.rhs, rhs]), np.concatenate([inst.range, np.zeros(k)]), inst.colType,
                      inst.lb, inst.ub, inst.inf, inst.xstar)


def test_singletons_match(oracle, gpu_ctx):
    inst = _with_singleton_rows(I.mixed_bigm(m=3000, n=3000, seed=31), 900, 32)
    mat = capi.Matrix(gpu_ctx, inst)
    nt, nf, lg, ug = mat.singletons(inst.lb, inst.ub)
    ont, onf, lo, uo = oracle.singletons(oracle.Problem(inst), inst.lb, inst.ub)
    assert (nt, nf) == (ont, onf) and nt > 100 and nf > 10
    assert_same_bounds(inst.colType, lg, ug, lo, uo, "singletons")
    assert inst.contains(lg, ug)
    # and an infeasible singleton: x_j <= lb_j - 1
    j = int(np.flatnonzero(inst.colType == 1)[0])
    bad = _with_singleton_rows(inst, 1, 33)
    bad.col[-2], bad.val[-2], bad.sense[-1], bad.rhs[-1] = j, 1.0, ord("L"), inst.lb[j] - 1.0
    mat2 = capi.Matrix(gpu_ctx, bad)
    assert mat2.singletons(bad.lb, bad.ub)[0] == -1 == oracle.singletons(oracle.Problem(bad), bad.lb, bad.ub)[0]
    mat.close()
    mat2.close()


def test_empty_and_ragged_inputs(oracle, gpu_ctx):
    # no rows at all; rows that are empty, free ('N') or all-zero
    inst = I.Instance("norows", 0, 3, np.zeros(1, np.int64), np.zeros(0, np.int32), np.zeros(0),
                      np.zeros(0, np.uint8), np.zeros(0), np.zeros(0), np.array([1, 2, 0], np.int8),
                      np.zeros(3), np.array([1.0, 2.5, 7.0]))
    (rg, lg, ug), (ro, lo, uo) = _both(oracle, gpu_ctx, inst)
    assert_same_result(rg, ro, "norows")
    assert np.array_equal(ug, uo) and ug[1] == 2.0
    def ragged(rhs2):
        return I.Instance("ragged", 4, 3, np.array([0, 0, 2, 4, 5], np.int64), np.array([0, 1, 0, 2, 1], np.int32),
                          np.array([1.0, 1.0, 0.0, 0.0, 3.0]), np.frombuffer(b"LNLL", np.uint8).copy(),
                          np.array([1.0, 0.5, rhs2, 4.5]), np.zeros(4), np.array([1, 2, 0], np.int8),
                          np.zeros(3), np.array([1.0, 9.0, 7.0]))
    inst = ragged(1.0)  # row 0 empty, row 1 free, row 2 all zeros with 0 <= 1, row 3: 3 x1 <= 4.5
    (rg, lg, ug), (ro, lo, uo) = _both(oracle, gpu_ctx, inst)
    assert_same_result(rg, ro, "ragged")
    assert np.array_equal(lg, lo) and np.array_equal(ug, uo) and ug[1] == 1.0
    inst = ragged(-1.0)  # the all-zero row now says 0 <= -1
    (rg, lg, ug), (ro, lo, uo) = _both(oracle, gpu_ctx, inst)
    assert_same_result(rg, ro, "ragged infeasible")
    assert not rg["feasible"] and rg["infeasibleRow"] == 2 and np.array_equal(ug, inst.ub)


def _check_batch(oracle, ctx, inst, s, v, b, what, **kw):
    mat = capi.Matrix(ctx, inst)
    res, us, uc, ul, uu = mat.propagate_batch_delta(inst.lb, inst.ub, s, v, b, **kw)
    Lg, Ug = mat.batch_fetch_dense()
    ores, Lo, Uo = oracle.propagate_batch_delta(oracle.Problem(inst), inst.lb, inst.ub, s, v, b, threads=4, **kw)
    for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
        bad = np.flatnonzero(res[k] != ores[k])
        assert bad.size == 0, f"{what}: {k} differs for vectors {bad[:8]}: {res[k][bad[:8]]} vs {ores[k][bad[:8]]}"
    for k in range(len(s) - 1):
        assert_same_bounds(inst.colType, Lg[k], Ug[k], Lo[k], Uo[k], f"{what} vector {k}")
    # sparse answer == dense answer restricted to what changed (feasible vectors only)
    L0, U0 = I.expand_deltas(inst.lb, inst.ub, s, v, b)
    for k in range(len(s) - 1):
        cols = uc[us[k]:us[k + 1]]
        if not res["feasible"][k]:
            assert cols.size == 0
            continue
        want = np.flatnonzero((Lg[k] != L0[k]) | (Ug[k] != U0[k]))
        assert np.array_equal(cols, want), f"{what} vector {k}: update list"
        assert np.array_equal(ul[us[k]:us[k + 1]], Lg[k][want]) and np.array_equal(uu[us[k]:us[k + 1]], Ug[k][want])
    # dense entry point gives the same
    res2, L2, U2 = mat.propagate_batch(L0, U0, **kw)
    for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
        assert np.array_equal(res2[k], res[k]), k
    assert np.array_equal(L2, Lg) and np.array_equal(U2, Ug)
    mat.close()
    return res


def test_batch_dives_knapsack_assignment(oracle, gpu_ctx):
    inst = I.knapsack_assignment(m=3000, n=8000, seed=41)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 70, max_depth=48, seed=42)
    res = _check_batch(oracle, gpu_ctx, inst, s, v, b, "dives")
    assert res["feasible"].astype(bool).any() and (~res["feasible"].astype(bool)).any()
    assert res["rounds"].max() >= 3


def test_batch_strong_branching_set_cover(oracle, gpu_ctx):
    inst = I.fast0507_shaped(m=101, n=9000, fix_frac=0.996, seed=22)
    r0, lb0, ub0 = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    inst.lb, inst.ub = lb0, ub0  # node bounds = propagated root, as in strong branching
    s, v, b = I.strong_branch_deltas(inst, lb0, ub0, 48, seed=43)
    res = _check_batch(oracle, gpu_ctx, inst, s, v, b, "strong branching")
    assert res["nChanged"].sum() > 0


def test_batch_probe_mixed(oracle, gpu_ctx):
    inst = I.mixed_bigm(m=6000, n=6000, seed=44)
    s, v, b = I.probe_deltas(inst, inst.lb, inst.ub, 20, seed=45)
    _check_batch(oracle, gpu_ctx, inst, s, v, b, "probes")
    _check_batch(oracle, gpu_ctx, inst, s, v, b, "probes maxRounds=3", max_rounds=3)


def test_batch_long_rows_mixed_signs(oracle, gpu_ctx):
    # rows far longer than a warp's block, 15% negative coefficients, continuous + integer columns
    inst = I.power_law(n=20000, nnz_target=150000, n_dense=3, dense_len=6000, max_len=2500, seed=46)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 40, max_depth=30, seed=47)
    res = _check_batch(oracle, gpu_ctx, inst, s, v, b, "long rows")
    assert res["nChanged"].sum() > 0


def test_batch_base_needs_rounding(oracle, gpu_ctx):
    # base bounds of integer columns are fractional: every vector starts with R11 changes
    inst = I.knapsack_assignment(m=600, n=1500, seed=48)
    inst.colType = inst.colType.copy()
    inst.colType[::3] = 2
    inst.ub = inst.ub.copy()
    inst.ub[::3] = 1.6
    inst.lb = inst.lb.copy()
    inst.lb[1::7] = -0.3
    s, v, b = I.dive_deltas(inst, np.ceil(inst.lb - 1e-7), np.floor(inst.ub + 1e-7), 33, max_depth=20, seed=49)
    _check_batch(oracle, gpu_ctx, inst, s, v, b, "fractional base")


@pytest.mark.parametrize("n,m", [(100_000, 40_000), (300_000, 120_000), (700_000, 280_000)])
def test_batch_wide_matrices_use_narrow_vector_slices(oracle, gpu_ctx, n, m, monkeypatch):
    """Slice widths 16, 8 and 4 of the slice-major layout (bt_rows_narrow_kernel), forced through the
    CBP_L2_BUDGET_MB knob (the default keeps W = 32); results must not depend on the width."""
    import subprocess
    import sys
    import os

    # the knob is read once per process: run this case in a child process with the knob set
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from cbc_b200 import capi, instances as I; from oracle import oracle_py as O\n"
        "inst = I.knapsack_assignment(m=%d, n=%d, seed=55)\n"
        "s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 37, max_depth=40, seed=56)\n"
        "ctx = capi.Context(); mat = capi.Matrix(ctx, inst)\n"
        "res, us, uc, ul, uu = mat.propagate_batch_delta(inst.lb, inst.ub, s, v, b)\n"
        "Lg, Ug = mat.batch_fetch_dense()\n"
        "ores, Lo, Uo = O.propagate_batch_delta(O.Problem(inst), inst.lb, inst.ub, s, v, b, threads=8)\n"
        "for k in ('feasible','rounds','nChanged','infeasibleRow','infeasibleCol','stopReason'):\n"
        "    assert np.array_equal(res[k], ores[k]), k\n"
        "f = res['feasible'].astype(bool)\n"
        "assert np.array_equal(Lg[f], Lo[f]) and np.array_equal(Ug[f], Uo[f])\n"
        "assert res['nChanged'].sum() > 0 and (~f).any()\n"
        "print('OK')\n") % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                            os.path.dirname(os.path.abspath(__file__)), m, n)
    env = dict(os.environ, CBP_L2_BUDGET_MB="48")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "OK" in out.stdout, out.stderr[-2000:]


def test_batch_first_round_from_the_base(oracle, gpu_ctx):
    """A batch with few diffs per vector runs its first round from the shared base bounds and builds the
    per-vector buffers only for groups that go on: vectors that repeat a column in their diffs, vectors
    with no diff at all, a last group that is not full, groups where nothing happens next to groups
    where a lot does, and the dense read-back of groups that were never built."""
    inst = I.fast0507_shaped(m=101, n=9000, fix_frac=0.996, seed=22)
    r0, lb0, ub0 = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    inst.lb, inst.ub = lb0, ub0
    s0, v0, b0 = I.strong_branch_deltas(inst, lb0, ub0, 40, seed=44)
    free = np.flatnonzero(ub0 > lb0)
    start, var, bnd = [0], [], []
    for k in range(100):
        if k < 40:  # strong-branching arms (groups 0 and 1: things happen)
            var += list(v0[s0[k]:s0[k + 1]])
            bnd += list(b0[s0[k]:s0[k + 1]])
        elif k % 3 == 0:  # nothing
            pass
        elif k % 3 == 1:  # the same column twice: ub to 0, then lb to 0 (a no-op after the first)
            j = int(free[(7 * k) % free.size])
            var += [j | I.UPPER_FLAG, j]
            bnd += [0.0, 0.0]
        else:  # ub lowered twice, the second value counts; then raised back: ends where it started
            j = int(free[(11 * k) % free.size])
            var += [j | I.UPPER_FLAG, j | I.UPPER_FLAG]
            bnd += [0.0, float(ub0[j])]
        start.append(len(var))
    s = np.array(start, np.int64)
    v = np.array(var, np.uint32)
    b = np.array(bnd, np.float64)
    res = _check_batch(oracle, gpu_ctx, inst, s, v, b, "from the base")
    assert res["nChanged"][:40].sum() > 0


@pytest.mark.parametrize("env", [{"CBP_LOG_CAP": "7", "CBP_NO_WORKLIST": "1"}, {"CBP_NO_WORKLIST": "1"},
                                 {"CBP_OVERLAY_ROUNDS": "1", "CBP_NO_WORKLIST": "1"},
                                 {"CBP_OVERLAY_ROUNDS": "2", "CBP_NO_WORKLIST": "1"},
                                 {"CBP_NO_MULTI_GROUP": "1", "CBP_NO_WORKLIST": "1"},
                                 {"CBP_NO_BASE_ROUND": "1", "CBP_NO_WORKLIST": "1"},
                                 {"CBP_NO_LAZY_X1": "1", "CBP_NO_WORKLIST": "1"}],
                         ids=["list_overflows", "overlay", "overlay_left_after_1", "overlay_left_after_2",
                              "overlay_one_group_per_warp", "lazy_buffer", "no_lazy_buffer"])
def test_batch_first_round_variants(env):
    """A base+deltas batch normally stays in overlay form (base bounds + touched cells, candidates through a list)
    for all its rounds; a list that is too small, leaving the overlay form after one or two rounds, the older
    lazy-second-buffer path and the plain two-buffer path must all give the same answer."""
    import os
    import subprocess
    import sys

    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "from cbc_b200 import capi, instances as I; from oracle import oracle_py as O\n"
        "inst = I.knapsack_assignment(m=3000, n=8000, seed=41)\n"
        "s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 70, max_depth=48, seed=42)\n"
        "ctx = capi.Context(); mat = capi.Matrix(ctx, inst)\n"
        "res, us, uc, ul, uu = mat.propagate_batch_delta(inst.lb, inst.ub, s, v, b)\n"
        "Lg, Ug = mat.batch_fetch_dense()\n"
        "ores, Lo, Uo = O.propagate_batch_delta(O.Problem(inst), inst.lb, inst.ub, s, v, b, threads=8)\n"
        "for k in ('feasible','rounds','nChanged','infeasibleRow','infeasibleCol','stopReason'):\n"
        "    assert np.array_equal(res[k], ores[k]), k\n"
        "assert np.array_equal(Lg, Lo) and np.array_equal(Ug, Uo)\n"
        "assert res['nChanged'].sum() > 100\n"
        "print('OK')\n") % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),)
    out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env), capture_output=True, text=True,
                         timeout=600)
    assert out.returncode == 0 and "OK" in out.stdout, out.stderr[-2000:]


def test_fixpoint_is_idempotent_at_size(gpu_ctx):
    """Size-independent property on a larger instance: propagating a fixpoint again changes nothing
    and takes exactly one round; the planted point survives."""
    inst = I.mixed_bigm(m=300_000, n=300_000, seed=51)
    mat = capi.Matrix(gpu_ctx, inst)
    r1, l1, u1 = mat.propagate(inst.lb, inst.ub)
    assert r1["feasible"] and r1["stopReason"] == 1 and inst.contains(l1, u1)
    assert np.all(l1 >= inst.lb) and np.all(u1 <= inst.ub)
    r2, l2, u2 = mat.propagate(l1, u1)
    assert r2["rounds"] == 1 and r2["nChanged"] == 0
    assert np.array_equal(l1, l2) and np.array_equal(u1, u2)
    mat.close()


def test_diff_encoding_flags_and_range(oracle, gpu_ctx):
    """Bound diffs come in CbcPartialNodeInfo's encoding: bit 31 = upper bound, bit 30 = the 'can go odd way' flag
    applyBounds() sets (/root/reference/src/CbcPartialNodeInfo.cpp:139,196,208,220) — not part of the column.  A diff
    that names a column outside the matrix is refused, on the host path and on the device-pointer path alike."""
    import torch

    inst = I.knapsack_assignment(m=600, n=1500, seed=41)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 40, max_depth=12, seed=42)
    flagged = v.copy()
    flagged[::3] |= I.ODD_WAY_FLAG
    mat = capi.Matrix(gpu_ctx, inst)
    r0 = mat.propagate_batch_delta(inst.lb, inst.ub, s, v, b)
    r1 = mat.propagate_batch_delta(inst.lb, inst.ub, s, flagged, b)
    for a, c in zip(r0, r1):
        assert np.array_equal(a, c) if a.dtype.names is None else all(
            np.array_equal(a[k], c[k]) for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol"))
    ores, _, _ = oracle.propagate_batch_delta(oracle.Problem(inst), inst.lb, inst.ub, s, flagged, b, want_bounds=False)
    for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
        assert np.array_equal(r1[0][k], ores[k]), k
    bad = v.copy()
    bad[5] = np.uint32(inst.n + 7) | I.UPPER_FLAG
    with pytest.raises(capi.CbpError):
        mat.batch_load_delta(inst.lb, inst.ub, s, bad, b)
    with pytest.raises(ValueError):
        oracle.propagate_batch_delta(oracle.Problem(inst), inst.lb, inst.ub, s, bad, b, want_bounds=False)
    # the same through device pointers
    dl, du = torch.tensor(inst.lb, device="cuda"), torch.tensor(inst.ub, device="cuda")
    ds = torch.tensor(s, device="cuda")
    dv = torch.tensor(bad.astype(np.int64), device="cuda").to(torch.int32) if False else torch.tensor(
        bad.view(np.int32), device="cuda")
    db = torch.tensor(b, device="cuda")
    rc = gpu_ctx.lib.cbp_batch_load_delta(mat.h, capi._ptr(dl.data_ptr()), capi._ptr(du.data_ptr()), len(s) - 1,
                                          capi._ptr(ds.data_ptr()), capi._ptr(dv.data_ptr()), capi._ptr(db.data_ptr()),
                                          capi.FLAG_DEVICE_PTRS)
    assert rc == -1
    mat.close()


def test_negative_zero_bounds(oracle, gpu_ctx):
    """-0.0 in the caller's bounds (the bit pattern the integer-ordered atomics would take for the most negative
    number): a continuous column with ub = -0.0 and lb < 0 whose lower bound is tightened over two rounds."""
    # x0 in [-8, -0.0] continuous, x1 in [0, 4] integer, x2 in [0, 10] continuous
    # row 0: x0 + x1 >= -2 ; row 1: 2 x1 <= 3 ; row 2: x2 - x0 <= 2
    inst = I.Instance("negzero", 3, 3, np.array([0, 2, 3, 5], np.int64), np.array([0, 1, 1, 2, 0], np.int32),
                      np.array([1.0, 1.0, 2.0, 1.0, -1.0]), np.frombuffer(b"GLL", np.uint8).copy(),
                      np.array([-2.0, 3.0, 2.0]), np.zeros(3), np.array([0, 2, 0], np.int8),
                      np.array([-8.0, 0.0, 0.0]), np.array([-0.0, 4.0, 10.0]))
    (rg, lg, ug), (ro, lo, uo) = _both(oracle, gpu_ctx, inst)
    assert_same_result(rg, ro, "negzero")
    assert rg["roundChanged"] == ro["roundChanged"] and rg["nChanged"] >= 3
    assert_same_bounds(inst.colType, lg, ug, lo, uo, "negzero")
    assert ug[2] == 2.0 and ug[1] == 1.0
    # and as a batch (classic form and diffs)
    mat = capi.Matrix(gpu_ctx, inst)
    res, L, U = mat.propagate_batch(np.tile(inst.lb, (2, 1)), np.tile(inst.ub, (2, 1)))
    assert np.array_equal(L[1], lo) and np.array_equal(U[1], uo)
    res, us, uc, ul, uu = mat.propagate_batch_delta(inst.lb, np.array([0.0, 4.0, 10.0]), np.array([0, 1], np.int64),
                                                    np.array([0], np.uint32) | I.UPPER_FLAG, np.array([-0.0]))
    assert int(res["rounds"][0]) == ro["rounds"] and int(res["nChanged"][0]) == ro["nChanged"]
    mat.close()


def test_row_activity_counters_long_free_row(oracle, gpu_ctx):
    """One row of 80,000 columns that are free on both sides: both infinite-bound counters must come back as
    80,000 (a 16-bit packing of the two counters in the warp reduction would wrap)."""
    n = 80_000
    inst = I.Instance("longfree", 1, n, np.array([0, n], np.int64), np.arange(n, dtype=np.int32), np.ones(n),
                      np.frombuffer(b"L", np.uint8).copy(), np.array([5.0]), np.zeros(1), np.zeros(n, np.int8),
                      np.full(n, -I.COIN_INF), np.full(n, I.COIN_INF))
    mat = capi.Matrix(gpu_ctx, inst)
    mn, mx, im, ix = mat.row_activity(inst.lb, inst.ub)
    assert (int(im[0]), int(ix[0])) == (n, n) and mn[0] == 0.0 and mx[0] == 0.0
    r, lb, ub = mat.propagate(inst.lb, inst.ub)
    ro, lo, uo = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    assert_same_result(r, ro, "longfree")
    mat.close()
