"""CPU-only: the oracle against the hand-derived golden vectors, its own
self-checks, and the instance generators."""
import numpy as np
import pytest

from cbc_b200 import instances as I
from helpers import golden_cases


@pytest.mark.parametrize("case", golden_cases(), ids=lambda c: c["name"])
def test_golden_micro(oracle, case):
    inst = I.micro(case["instance"])
    p = oracle.Problem(inst)
    r, lb, ub = oracle.propagate(p, case["lb_in"], case["ub_in"])
    assert r["feasible"] == case["feasible"]
    assert r["rounds"] == case["rounds"]
    assert r["nChanged"] == case["nChanged"]
    assert r["infeasibleRow"] == case.get("infeasibleRow", -1)
    assert r["infeasibleCol"] == case.get("infeasibleCol", -1)
    assert np.array_equal(lb, np.array(case["lb"], float))
    assert np.array_equal(ub, np.array(case["ub"], float))
    if "singletons" in case:
        nt, nf, _, _ = oracle.singletons(p, case["lb_in"], case["ub_in"])
        assert nt == case["singletons"]


def test_stop_reasons_follow_reference_loop(oracle):
    # src/CbcBoundPropagation.cpp:107-201: fixpoint round counts; maxRounds exhausted -> HitMaxRounds;
    # the time test happens before the first round.
    inst = I.mixed_bigm(m=3000, n=3000, seed=7)
    p = oracle.Problem(inst)
    full, _, _ = oracle.propagate(p, inst.lb, inst.ub)
    assert full["stopReason"] == 1 and full["rounds"] >= 3
    assert full["roundChanged"][-1] == 0 and sum(full["roundChanged"]) == full["nChanged"]
    cut, lb2, ub2 = oracle.propagate(p, inst.lb, inst.ub, max_rounds=2)
    assert cut["stopReason"] == 2 and cut["rounds"] == 2
    assert cut["nChanged"] == sum(full["roundChanged"][:2])
    tl, lb3, ub3 = oracle.propagate(p, inst.lb, inst.ub, time_limit=0.0)
    assert tl["stopReason"] == 3 and tl["rounds"] == 0 and tl["feasible"]
    assert np.array_equal(lb3, inst.lb) and np.array_equal(ub3, inst.ub)


@pytest.mark.parametrize("maker", [
    lambda: I.set_cover(m=400, n=800, density=0.02, fix_frac=0.9, seed=3),
    lambda: I.fast0507_shaped(m=60, n=3000, fix_frac=0.97, seed=4),
    lambda: I.knapsack_assignment(m=800, n=2000, seed=5),
    lambda: I.mixed_bigm(m=4000, n=4000, seed=6),
    lambda: I.power_law(n=5000, nnz_target=40000, n_dense=2, dense_len=3000, max_len=500, seed=8),
])
def test_planted_point_survives(oracle, maker):
    inst = maker()
    assert inst.is_feasible_point()
    p = oracle.Problem(inst)
    r, lb, ub = oracle.propagate(p, inst.lb, inst.ub)
    assert r["feasible"] and r["stopReason"] == 1
    assert inst.contains(lb, ub)
    assert np.all(lb >= inst.lb) and np.all(ub <= inst.ub)  # never loosens


def test_jacobi_and_gauss_seidel_agree_on_pure_integer(oracle):
    inst = I.knapsack_assignment(m=800, n=2000, seed=5)
    p = oracle.Problem(inst)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 16, max_depth=40, seed=9)
    L, U = I.expand_deltas(inst.lb, inst.ub, s, v, b)
    for k in range(16):
        rj, lj, uj = oracle.propagate(p, L[k], U[k])
        rg, lg, ug = oracle.propagate(p, L[k], U[k], gauss_seidel=True)
        assert rj["feasible"] == rg["feasible"]
        if rj["feasible"]:
            assert np.array_equal(lj, lg) and np.array_equal(uj, ug)


def test_infeasible_dense_row_detected(oracle):
    inst = I.power_law(n=5000, nnz_target=40000, n_dense=2, dense_len=3000, max_len=500, seed=8, infeasible=True)
    p = oracle.Problem(inst)
    r, _, _ = oracle.propagate(p, inst.lb, inst.ub)
    assert not r["feasible"] and r["infeasibleRow"] == inst.meta["bad_row"] and r["rounds"] == 1


def test_singletons(oracle):
    # rows: 2x0 <= 3 (int -> ub 1), -x1 <= -2 (cont -> lb 2), x2 = 4 (fix), x3 + x4 <= 1 (not a singleton)
    inst = I.Instance("sing", 4, 5, np.array([0, 1, 2, 3, 5], np.int64), np.array([0, 1, 2, 3, 4], np.int32),
                      np.array([2.0, -1.0, 1.0, 1.0, 1.0]), np.frombuffer(b"LLEL", np.uint8).copy(),
                      np.array([3.0, -2.0, 4.0, 1.0]), np.zeros(4), np.array([2, 0, 2, 1, 1], np.int8),
                      np.zeros(5), np.array([10.0, 10.0, 10.0, 1.0, 1.0]))
    p = oracle.Problem(inst)
    nt, nf, lb, ub = oracle.singletons(p, inst.lb, inst.ub)
    assert (nt, nf) == (3, 1)
    assert np.array_equal(lb, [0, 2, 4, 0, 0]) and np.array_equal(ub, [1, 10, 4, 1, 1])


def test_batch_delta_equals_dense(oracle):
    inst = I.knapsack_assignment(m=400, n=1000, seed=11)
    p = oracle.Problem(inst)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 24, max_depth=30, seed=12)
    L, U = I.expand_deltas(inst.lb, inst.ub, s, v, b)
    r1, l1, u1 = oracle.propagate_batch(p, L, U, threads=2)
    r2, l2, u2 = oracle.propagate_batch_delta(p, inst.lb, inst.ub, s, v, b, threads=2)
    for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
        assert np.array_equal(r1[k], r2[k])
    assert np.array_equal(l1, l2) and np.array_equal(u1, u2)
    assert (~r1["feasible"].astype(bool)).any() and r1["feasible"].astype(bool).any()


def test_mps_round_trip(tmp_path, oracle):
    inst = I.mixed_bigm(m=300, n=300, seed=13)
    path = tmp_path / "x.mps"
    I.write_mps(inst, str(path))
    back = I.read_mps(str(path))
    assert back.m == inst.m and back.n == inst.n and back.nnz == inst.nnz
    assert np.array_equal(back.col, inst.col) and np.array_equal(back.val, inst.val)
    assert np.array_equal(back.sense, inst.sense) and np.array_equal(back.rhs, inst.rhs)
    assert np.array_equal(back.lb, inst.lb) and np.array_equal(back.ub, inst.ub)
    assert np.array_equal(back.colType != 0, inst.colType != 0)
    ra, la, ua = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    rb, lb_, ub_ = oracle.propagate(oracle.Problem(back), back.lb, back.ub)
    assert ra["rounds"] == rb["rounds"] and np.array_equal(la, lb_) and np.array_equal(ua, ub_)


def test_diff_encoding_flags_and_range(oracle):
    # CbcPartialNodeInfo's encoding (src/CbcPartialNodeInfo.cpp:139,196,208): bit 31 = upper bound, bit 30 = the
    # 'can go odd way' flag, bits 0-29 = column; a column outside the matrix is an error, not a wild write
    inst = I.knapsack_assignment(m=400, n=1000, seed=11)
    p = oracle.Problem(inst)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 12, max_depth=10, seed=12)
    flagged = v.copy()
    flagged[::2] |= I.ODD_WAY_FLAG
    r1, l1, u1 = oracle.propagate_batch_delta(p, inst.lb, inst.ub, s, v, b)
    r2, l2, u2 = oracle.propagate_batch_delta(p, inst.lb, inst.ub, s, flagged, b)
    assert np.array_equal(l1, l2) and np.array_equal(u1, u2) and np.array_equal(r1["rounds"], r2["rounds"])
    L, U = I.expand_deltas(inst.lb, inst.ub, s, flagged, b)
    L0, U0 = I.expand_deltas(inst.lb, inst.ub, s, v, b)
    assert np.array_equal(L, L0) and np.array_equal(U, U0)
    bad = v.copy()
    bad[3] = np.uint32(inst.n) | I.UPPER_FLAG
    with pytest.raises(ValueError):
        oracle.propagate_batch_delta(p, inst.lb, inst.ub, s, bad, b)
