"""GPU, at BASELINE.json's full sizes, against the CPU oracle: whole instances for the single-vector configs, all
vectors' verdicts plus a sample's bounds for the batches; and size-independent properties beside them."""
import numpy as np
import pytest

from cbc_b200 import capi
from cbc_b200 import instances as I
from helpers import assert_same_bounds, assert_same_result

pytestmark = pytest.mark.gpu


def test_c2_full_batch_4096_vectors(oracle, gpu_ctx):
    """fast0507-shaped 507 x 63,009 with 4096 strong-branching vectors (BASELINE config 2)."""
    inst = I.fast0507_shaped(fix_frac=0.996, seed=I.SEED0 + 2)  # deep node: rows are nearly tight
    r0, lb0, ub0 = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    assert r0["feasible"]
    s, v, b = I.strong_branch_deltas(inst, lb0, ub0, 2048, seed=101)
    mat = capi.Matrix(gpu_ctx, inst)
    res, us, uc, ul, uu = mat.propagate_batch_delta(lb0, ub0, s, v, b)
    assert len(res) == 4096 and res["rounds"].min() >= 1
    # every arm: base + one changed bound; what the batch reports must be what a single-vector run reports
    pick = np.concatenate([np.arange(8), np.flatnonzero(res["nChanged"] > 0)[:8], np.flatnonzero(res["feasible"] == 0)[:4]])
    for k in pick:
        lb, ub = lb0.copy(), ub0.copy()
        j = int(v[k] & 0x3FFFFFFF)
        if v[k] & 0x80000000:
            ub[j] = b[k]
        else:
            lb[j] = b[k]
        r1, l1, u1 = mat.propagate(lb, ub)
        for key in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
            assert int(res[key][k]) == int(r1[key]), (k, key)
        ro, lo, uo = oracle.propagate(oracle.Problem(inst), lb, ub)
        assert_same_result(r1, ro, f"arm {k}")
        assert_same_bounds(inst.colType, l1, u1, lo, uo, f"arm {k}")
        if r1["feasible"]:
            cols = uc[us[k]:us[k + 1]]
            want = np.flatnonzero((l1 != lb) | (u1 != ub))
            assert np.array_equal(cols, want)
            assert np.array_equal(ul[us[k]:us[k + 1]], l1[want]) and np.array_equal(uu[us[k]:us[k + 1]], u1[want])
    # down/up arms of one candidate can never both keep the planted point out of reach: at least one arm feasible
    feas = res["feasible"].astype(bool)
    assert np.all(feas[0::2] | feas[1::2])
    # rerunning the loaded batch gives the same answer (every run starts again from base + deltas)
    mat.batch_run()
    res2 = mat.batch_fetch_results()
    for key in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
        assert np.array_equal(res[key], res2[key])
    mat.close()


KEYS = ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason")


def _batch_against_oracle(oracle, ctx, mat, inst, lb, ub, s, v, b, sample, what, want_mode=None):
    """all vectors' verdicts, and the dense bounds of the first `sample` vectors, against the oracle"""
    import os

    res, us, uc, ul, uu = mat.propagate_batch_delta(lb, ub, s, v, b)
    if want_mode is not None:
        assert ctx.stats()["lastMode"] == want_mode, ctx.stats()
    prob = oracle.Problem(inst)
    threads = max(1, len(os.sched_getaffinity(0)))
    ores, _, _ = oracle.propagate_batch_delta(prob, lb, ub, s, v, b, threads=threads, want_bounds=False)
    for k in KEYS:
        bad = np.flatnonzero(res[k] != ores[k])
        assert bad.size == 0, f"{what}: {k} differs for vectors {bad[:8]}: {res[k][bad[:8]]} vs {ores[k][bad[:8]]}"
    k = sample
    _, Lo, Uo = oracle.propagate_batch_delta(prob, lb, ub, s[:k + 1], v[:s[k]], b[:s[k]], threads=threads)
    for i in range(k):
        if not res["feasible"][i]:
            assert us[i + 1] == us[i]
            continue
        l, u = np.array(lb), np.array(ub)
        for q in range(s[i], s[i + 1]):
            j = int(v[q] & I.COL_MASK)
            if v[q] & I.UPPER_FLAG:
                u[j] = b[q]
            else:
                l[j] = b[q]
        cols = uc[us[i]:us[i + 1]]
        l_in, u_in = l.copy(), u.copy()
        l[cols], u[cols] = ul[us[i]:us[i + 1]], uu[us[i]:us[i + 1]]
        assert_same_bounds(inst.colType, l, u, Lo[i], Uo[i], f"{what} vector {i}")
        assert np.array_equal(cols, np.flatnonzero((l != l_in) | (u != u_in))), f"{what} vector {i}: update list"

    return res


def test_c3_full_single_vector_1m_and_dive_batch(oracle, gpu_ctx):
    """1M x 1M, 10M nnz mixed-integer (BASELINE config 3): the whole fixpoint run against the oracle (round by round
    counts, every bound); sound, monotone, idempotent; then the north-star shape, a batch of dive vectors from the
    propagated root on the same matrix (c3b), all vectors against the oracle."""
    inst = I.mixed_bigm()
    mat = capi.Matrix(gpu_ctx, inst)
    r1, l1, u1 = mat.propagate(inst.lb, inst.ub)
    ro, lo, uo = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    assert_same_result(r1, ro, "c3")
    assert r1["roundChanged"] == ro["roundChanged"]
    assert_same_bounds(inst.colType, l1, u1, lo, uo, "c3")
    assert r1["feasible"] and r1["stopReason"] == 1 and r1["rounds"] > 5
    assert inst.contains(l1, u1)
    assert np.all(l1 >= inst.lb) and np.all(u1 <= inst.ub)
    isint = inst.colType != 0
    fin = np.abs(u1) < inst.inf
    assert np.all(l1[isint] == np.round(l1[isint])) and np.all(u1[isint & fin] == np.round(u1[isint & fin]))
    r2, l2, u2 = mat.propagate(l1, u1)
    assert r2["rounds"] == 1 and r2["nChanged"] == 0 and np.array_equal(l1, l2) and np.array_equal(u1, u2)
    # truncated runs are prefixes of the full run
    r3, l3, u3 = mat.propagate(inst.lb, inst.ub, max_rounds=3)
    assert r3["stopReason"] == 2 and r3["nChanged"] == sum(r1["roundChanged"][:3])
    assert np.all(l3 <= l1) and np.all(u3 >= u1)
    # c3b: node bounds = the oracle's propagated root, 96 dive vectors (worklist form)
    s, v, b = I.dive_deltas(inst, lo, uo, 96, max_depth=16, seed=I.SEED0 + 300)
    res = _batch_against_oracle(oracle, gpu_ctx, mat, inst, lo, uo, s, v, b, 8, "c3b", want_mode=2)
    assert res["rounds"].max() >= 5 and res["nChanged"].sum() > 1000
    mat.close()


def test_c4_full_knapsack_assignment_dive_batches(oracle, gpu_ctx):
    """200k x 500k, 3M nnz (BASELINE config 4): 2,048 dive vectors from the propagated root (worklist form) and 256
    from the unpropagated root (every row can act for every vector: the full-walk overlay kernels)."""
    inst = I.knapsack_assignment(root_fixpoint=True)
    r0, lb0, ub0 = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    assert r0["feasible"] and r0["nChanged"] > 0  # (a few one-column assignment rows fix their column)
    mat = capi.Matrix(gpu_ctx, inst)
    s, v, b = I.dive_deltas(inst, lb0, ub0, 2048, max_depth=64, seed=I.SEED0 + 200)
    res = _batch_against_oracle(oracle, gpu_ctx, mat, inst, lb0, ub0, s, v, b, 64, "c4 node", want_mode=2)
    assert res["nChanged"].sum() > 10000 and res["rounds"].max() >= 3
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 256, max_depth=64, seed=I.SEED0 + 201)
    res = _batch_against_oracle(oracle, gpu_ctx, mat, inst, inst.lb, inst.ub, s, v, b, 32, "c4 root", want_mode=1)
    assert res["nChanged"].min() > 0
    mat.close()


def test_c1_written_as_mps_root_vector(oracle, gpu_ctx, tmp_path):
    """10k x 20k set cover written as MPS and read back (BASELINE config 1): GPU == CPU oracle at the root."""
    inst = I.set_cover(fix_frac=0.93)
    path = str(tmp_path / "c1.mps")
    I.write_mps(inst, path)
    back = I.read_mps(path)
    assert back.m == inst.m and back.n == inst.n and back.nnz == inst.nnz
    mat = capi.Matrix(gpu_ctx, back)
    rg, lg, ug = mat.propagate(back.lb, back.ub)
    ro, lo, uo = oracle.propagate(oracle.Problem(back), back.lb, back.ub)
    assert_same_result(rg, ro, "c1")
    assert np.array_equal(lg, lo) and np.array_equal(ug, uo) and rg["nChanged"] > 0
    assert inst.contains(lg, ug)
    mat.close()


def test_c5_full_power_law_detects_infeasible_dense_row(oracle, gpu_ctx):
    """Power-law rows with eight 100k-nonzero rows (BASELINE config 5), feasible and infeasible variants."""
    for bad in (False, True):
        inst = I.power_law(infeasible=bad)
        mat = capi.Matrix(gpu_ctx, inst)
        rg, lg, ug = mat.propagate(inst.lb, inst.ub)
        ro, lo, uo = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
        assert_same_result(rg, ro, f"c5 infeasible={bad}")
        assert_same_bounds(inst.colType, lg, ug, lo, uo, "c5")
        if bad:
            assert not rg["feasible"] and rg["infeasibleRow"] == inst.meta["bad_row"]
        else:
            assert inst.contains(lg, ug)
        mat.close()
