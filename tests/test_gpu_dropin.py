"""GPU: the C++ drop-ins (CbcBoundPropagation::run, CoinBoundPropagation) driven through the COIN test
double exactly as Cbc's callers drive them, compared with the CPU oracle."""
import json
import os
import struct
import subprocess

import numpy as np
import pytest

from cbc_b200 import instances as I
from helpers import assert_same_bounds

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "cbc_b200", "lib", "cbp_dropin_test")


def write_instance(inst, path):
    with open(path, "wb") as f:
        f.write(struct.pack("<iiq", inst.m, inst.n, inst.nnz))
        f.write(np.ascontiguousarray(inst.rowStart, np.int32).tobytes())
        f.write(np.ascontiguousarray(inst.col, np.int32).tobytes())
        f.write(np.ascontiguousarray(inst.val, np.float64).tobytes())
        f.write(np.ascontiguousarray(inst.sense, np.uint8).tobytes())
        f.write(np.ascontiguousarray(inst.rhs, np.float64).tobytes())
        f.write(np.ascontiguousarray(inst.range, np.float64).tobytes())
        f.write(np.ascontiguousarray(inst.colType, np.int8).tobytes())
        f.write(np.ascontiguousarray(inst.lb, np.float64).tobytes())
        f.write(np.ascontiguousarray(inst.ub, np.float64).tobytes())
        f.write(struct.pack("<d", inst.inf))


def drive(tmp_path, inst, level, max_rounds=100, log=2, time_limit=1e100, dive=False):
    ip, op = str(tmp_path / "inst.bin"), str(tmp_path / "out.bin")
    write_instance(inst, ip)
    cmd = [DRIVER, ip, op, str(level), str(max_rounds), str(log), repr(float(time_limit))] + (["dive"] if dive else [])
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    res = json.loads([ln for ln in out.stdout.splitlines() if ln.startswith("RESULT ")][0][7:])
    again = [json.loads(ln[6:]) for ln in out.stdout.splitlines() if ln.startswith("AGAIN ")]
    b = np.fromfile(op)
    return res, again, b[: inst.n], b[inst.n:], out.stdout


def oracle_run(oracle, inst, level, max_rounds):
    p = oracle.Problem(inst)
    nt, nf, lb, ub = oracle.singletons(p, inst.lb, inst.ub)
    if nt < 0 or level == 1:
        return nt, nf, None, lb, ub
    r, lb, ub = oracle.propagate(p, lb, ub, max_rounds=0 if level == 3 else max_rounds)
    return nt, nf, r, lb, ub


@pytest.mark.parametrize("level,max_rounds", [(3, 0), (2, 100), (2, 3), (1, 100)])
def test_run_matches_oracle(oracle, tmp_path, level, max_rounds):
    from test_gpu_parity import _with_singleton_rows

    inst = _with_singleton_rows(I.mixed_bigm(m=5000, n=5000, seed=61), 400, 62)
    res, again, lb, ub, log = drive(tmp_path, inst, level, max_rounds)
    nt, nf, r, lo, uo = oracle_run(oracle, inst, level, max_rounds)
    assert res["feasible"] == 1
    assert res["nSingletonFixed"] == nf and res["nSingletonTightened"] == nt - nf
    assert_same_bounds(inst.colType, lb, ub, lo, uo, f"level {level}")
    if level == 1:
        assert res["stopReason"] == 1 and res["nRoundsRun"] == 0  # Singletons stops with ReachedFixpoint
        assert f"  Bound propagation fixed {nf} vars in" in log
        return
    assert res["nRoundsRun"] == r["rounds"] and res["nBoundPropFixed"] == r["nChanged"]
    assert res["stopReason"] == r["stopReason"] and res["nFixed"] == nf + r["nChanged"]
    changed = int(np.count_nonzero((lb != inst.lb) | (ub != inst.ub)))
    assert res["nSetLower"] >= changed - nt and res["nSetUpper"] >= changed - nt  # setColLower AND setColUpper per column
    # the reference's output lines (src/CbcBoundPropagation.cpp:70,192,174,205)
    assert f"  Bound propagation (singletons): {nt - nf} tightened, {nf} fixed." in log
    assert f"  Bound propagation: round 1 fixed {r['roundChanged'][0]} variables (total {r['roundChanged'][0]})." in log
    if r["stopReason"] == 1:
        assert f"  Bound propagation: fixpoint reached after {r['rounds']} round(s)." in log
        assert again and again[0]["nRoundsRun"] == 1 and again[0]["nBoundPropFixed"] == 0
    assert f"  Bound propagation fixed {nf + r['nChanged']} vars in" in log


def test_run_infeasible_and_time_limit(oracle, tmp_path):
    inst = I.power_law(n=20000, nnz_target=120000, n_dense=2, dense_len=5000, max_len=2000, seed=63, infeasible=True)
    res, _, lb, ub, log = drive(tmp_path, inst, 3, 0, log=1)
    nt, nf, r, lo, uo = oracle_run(oracle, inst, 3, 0)
    assert res["feasible"] == 0 and res["stopReason"] == 4
    assert res["infeasibleRow"] == r["infeasibleRow"] == inst.meta["bad_row"] and res["infeasibleCol"] == -1
    assert res["nRoundsRun"] == r["rounds"]
    assert f"  Bound propagation: INFEASIBLE in round {r['rounds']} — row {r['infeasibleRow']} (R{r['infeasibleRow']})," in log
    # an expired budget: phase 1 still runs, then the loop stops before its first round and reports feasible
    inst = I.mixed_bigm(m=3000, n=3000, seed=64)
    res, _, lb, ub, log = drive(tmp_path, inst, 3, 0, log=1, time_limit=0.0)
    assert res["feasible"] == 1 and res["stopReason"] == 3 and res["nRoundsRun"] == 0
    assert "TIME LIMIT" in log


def test_dive_step(oracle, tmp_path):
    inst = I.knapsack_assignment(m=2000, n=5000, seed=65)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 6, max_depth=30, seed=66)
    L, U = I.expand_deltas(inst.lb, inst.ub, s, v, b)
    seen = set()
    for k in range(6):
        inst.lb, inst.ub = L[k], U[k]
        res, _, lb, ub, _ = drive(tmp_path, inst, 2, 1, dive=True)
        r, lo, uo = oracle.propagate(oracle.Problem(inst), L[k], U[k], max_rounds=1)
        assert res["ran"] == 1 and res["infeasible"] == (0 if r["feasible"] else 1)
        seen.add(res["infeasible"])
        if r["feasible"]:
            assert res["totalBoundFixed"] == r["nChanged"]
            assert np.array_equal(lb, lo) and np.array_equal(ub, uo)
        else:
            assert res["infeasibleRow"] == r["infeasibleRow"] and res["reasonToStop"] == 1


def test_strong_branching_arms_match_oracle(oracle, tmp_path):
    """CbcStrongPropagation: the 2 x numberToDo arms of a strong-branching step as one batch,
    against the oracle run arm by arm: verdicts, rounds, implied bounds, forced fixings."""
    inst = I.fast0507_shaped(m=101, n=9000, fix_frac=0.996, seed=22)
    r0, lb0, ub0 = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    inst.lb, inst.ub = lb0, ub0  # node bounds = propagated root
    s, v, b = I.strong_branch_deltas(inst, lb0, ub0, 60, seed=71)
    cols = (v[0::2] & 0x3FFFFFFF).astype(np.int32)
    vals = b[0::2] + 0.5  # LP value between the two arms
    ip, op, cp = str(tmp_path / "inst.bin"), str(tmp_path / "out.bin"), str(tmp_path / "cands.bin")
    write_instance(inst, ip)
    with open(cp, "wb") as f:
        f.write(struct.pack("<i", len(cols)))
        f.write(cols.tobytes())
        f.write(np.ascontiguousarray(vals, np.float64).tobytes())
    out = subprocess.run([DRIVER, ip, op, "3", "0", "0", "1e100", "strong", cp], capture_output=True, text=True,
                         timeout=300)
    assert out.returncode == 0, out.stderr
    res = json.loads([ln for ln in out.stdout.splitlines() if ln.startswith("RESULT ")][0][7:])
    assert res["ran"] == 1
    ores, Lo, Uo = oracle.propagate_batch_delta(oracle.Problem(inst), lb0, ub0, s, v, b, threads=4)
    L0, U0 = I.expand_deltas(lb0, ub0, s, v, b)
    raw = np.fromfile(op)
    at = 0
    n_inf = 0
    for k in range(len(cols)):
        for w in range(2):
            a = 2 * k + w
            head = raw[at:at + 8]
            at += 8
            cnt = int(head[4])
            rec = raw[at:at + 3 * cnt].reshape(cnt, 3)
            at += 3 * cnt
            assert int(head[0]) == (0 if ores["feasible"][a] else 1), (k, w)
            assert int(head[3]) == ores["rounds"][a]
            if not ores["feasible"][a]:
                n_inf += 1
                assert int(head[1]) == ores["infeasibleRow"][a] and int(head[2]) == ores["infeasibleCol"][a]
                assert cnt == 0
                continue
            want = np.flatnonzero((Lo[a] != L0[a]) | (Uo[a] != U0[a]))
            assert np.array_equal(rec[:, 0].astype(np.int64), want), (k, w)
            assert np.array_equal(rec[:, 1], Lo[a][want]) and np.array_equal(rec[:, 2], Uo[a][want])
            d_inf, u_inf = not ores["feasible"][2 * k], not ores["feasible"][2 * k + 1]
            assert int(head[5]) == (1 if d_inf != u_inf else 0)
            if d_inf != u_inf:
                assert int(head[6]) == (0 if d_inf else 1)
                assert head[7] == (np.ceil(vals[k]) if d_inf else np.floor(vals[k]))
    assert at == raw.size
    assert res["nInfeasibleArms"] == n_inf
    both = any((not ores["feasible"][2 * k]) and (not ores["feasible"][2 * k + 1]) for k in range(len(cols)))
    assert res["nodeInfeasible"] == (1 if both else 0)
