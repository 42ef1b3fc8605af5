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


@pytest.mark.parametrize("env", [{}, {"CBP_BATCH_BYTES": "1"}], ids=["one_batch", "pieces_of_64_arms"])
def test_strong_branching_arms_match_oracle(oracle, tmp_path, env):
    """CbcStrongPropagation: the 2 x numberToDo arms of a strong-branching step as one batch (or, when
    they would not fit the device as one, in pieces: CBP_BATCH_BYTES), against the oracle run arm by arm: verdicts,
    rounds, implied bounds, forced fixings."""
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
                         timeout=300, env=dict(os.environ, **env))
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


def _drive_mode(tmp_path, inst, mode_args, level=3, max_rounds=0, extra_env=None):
    ip, op = str(tmp_path / "inst.bin"), str(tmp_path / "out.bin")
    write_instance(inst, ip)
    cmd = [DRIVER, ip, op, str(level), str(max_rounds), "0", "1e100"] + mode_args
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, **(extra_env or {})))
    assert out.returncode == 0, out.stderr
    res = json.loads([ln for ln in out.stdout.splitlines() if ln.startswith("RESULT ")][0][7:])
    return res, op


def test_dive_batch_threads_match_oracle(oracle, tmp_path):
    """CbcDiveBatch: 8 dive threads, each walking through its own sequence of bound vectors (one more
    fixing per step), served by one device call per wave; every step of every dive against the oracle's one round."""
    inst = I.knapsack_assignment(m=2000, n=5000, seed=65, root_fixpoint=True)
    r0, lb0, ub0 = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    inst.lb, inst.ub = lb0, ub0
    T, n = 8, inst.n
    g = np.random.default_rng(77)
    seqs = []
    for t in range(T):
        steps = int(g.integers(2, 7))
        s, v, b = I.dive_deltas(inst, lb0, ub0, 1, max_depth=steps, seed=100 + t, follow_xstar=0.7 if t % 3 else 0.2)
        lb, ub = lb0.copy(), ub0.copy()
        seq = []
        for q in range(len(v)):  # one more fixing per dive iteration, bounds carried forward with what propagation found
            j = int(v[q] & I.COL_MASK)
            if v[q] & I.UPPER_FLAG:
                ub[j] = min(ub[j], b[q])
            else:
                lb[j] = max(lb[j], b[q])
            if lb[j] > ub[j]:
                break
            seq.append((lb.copy(), ub.copy()))
            r, lb2, ub2 = oracle.propagate(oracle.Problem(inst), lb, ub, max_rounds=1)
            if not r["feasible"]:
                break
            lb, ub = lb2, ub2
        seqs.append(seq)
    bp = str(tmp_path / "bounds.bin")
    with open(bp, "wb") as f:
        f.write(struct.pack("<i", T))
        for seq in seqs:
            f.write(struct.pack("<i", len(seq)))
            for lb, ub in seq:
                f.write(lb.tobytes())
                f.write(ub.tobytes())
    res, op = _drive_mode(tmp_path, inst, ["divebatch", bp], level=2, max_rounds=1)
    assert res["usable"] == 1 and res["dives"] == T
    assert res["waves"] == max(len(s) for s in seqs)  # one device call per wave, however many dives are still running
    assert res["vectors"] == res["steps"] == sum(len(s) for s in seqs)
    raw = open(op, "rb").read()
    off = 0
    failed = 0
    for t in range(T):
        (steps,) = struct.unpack_from("<i", raw, off)
        off += 4
        assert steps == len(seqs[t])
        for w in range(steps):
            feas, row, col, nupd = struct.unpack_from("<4i", raw, off)
            off += 16
            upd = np.frombuffer(raw, np.float64, 3 * nupd, off).reshape(-1, 3)
            off += 24 * nupd
            lb, ub = seqs[t][w]
            r, lo, uo = oracle.propagate(oracle.Problem(inst), lb, ub, max_rounds=1)
            assert feas == (1 if r["feasible"] else 0), (t, w)
            if not r["feasible"]:
                failed += 1
                assert (row, col) == (r["infeasibleRow"], r["infeasibleCol"]) and nupd == 0
                continue
            want = np.flatnonzero((lo != lb) | (uo != ub))
            assert np.array_equal(upd[:, 0].astype(int), want), (t, w)
            assert np.array_equal(upd[:, 1], lo[want]) and np.array_equal(upd[:, 2], uo[want])
    assert off == len(raw)


@pytest.mark.parametrize("env", [None, {"CBP_BATCH_BYTES": "1"}], ids=["one_batch", "pieces_of_64_arms"])
def test_probe_sweep_consequences(oracle, tmp_path, env):
    """CbcProbePropagation: every free binary probed both ways as one batch; a failed probe fixes the
    column the other way (src/CbcSolverHeuristics.cpp:483-486), what both probes imply holds at the node; all of it
    recomputed here from oracle runs probe by probe."""
    inst = I.knapsack_assignment(m=300, n=700, seed=81, root_fixpoint=True, slack_max=0)
    # planted structure that only probing sees: a + b <= 1, a + c <= 1, b + c >= 1 (a = 1 cannot hold: ub(a) = 0);
    # e + d >= 1, d - e >= 0 (both probes of e give d = 1: lb(d) = 1)
    x0, x1 = np.flatnonzero(inst.xstar == 0), np.flatnonzero(inst.xstar == 1)
    a, e = int(x0[3]), int(x0[11])
    b, c, d = int(x1[2]), int(x1[5]), int(x1[9])
    new_rows = [([a, b], [1.0, 1.0], "L", 1.0), ([a, c], [1.0, 1.0], "L", 1.0), ([b, c], [1.0, 1.0], "G", 1.0),
                ([e, d], [1.0, 1.0], "G", 1.0), ([d, e], [1.0, -1.0], "G", 0.0)]
    k = len(new_rows)
    inst = I.Instance(inst.name + "+probe", inst.m + k, inst.n,
                      np.concatenate([inst.rowStart, inst.rowStart[-1] + 2 * np.arange(1, k + 1)]),
                      np.concatenate([inst.col, np.array([j for r in new_rows for j in r[0]], np.int32)]),
                      np.concatenate([inst.val, np.array([v for r in new_rows for v in r[1]])]),
                      np.concatenate([inst.sense, np.array([ord(r[2]) for r in new_rows], np.uint8)]),
                      np.concatenate([inst.rhs, np.array([r[3] for r in new_rows])]),
                      np.concatenate([inst.range, np.zeros(k)]), inst.colType, inst.lb, inst.ub, inst.inf, inst.xstar)
    assert inst.is_feasible_point()
    r0, lb0, ub0 = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    inst.lb, inst.ub = lb0, ub0
    assert ub0[a] == 1.0 and lb0[d] == 0.0  # propagation alone does not see either conclusion
    res, op = _drive_mode(tmp_path, inst, ["probe", "0"], extra_env=env)
    raw = open(op, "rb").read()
    (np_,) = struct.unpack_from("<i", raw, 0)
    probes = np.frombuffer(raw, np.int32, 3 * np_, 4).reshape(-1, 3)
    off = 4 + 12 * np_
    (nb,) = struct.unpack_from("<i", raw, off)
    nbounds = np.frombuffer(raw, np.float64, 3 * nb, off + 4).reshape(-1, 3)
    off += 4 + 24 * nb
    fin = np.frombuffer(raw, np.float64, 2 * inst.n, off)
    cols = np.flatnonzero((inst.colType != 0) & (ub0 - lb0 >= 1.0))
    assert res["ran"] == 1 and res["numberProbed"] == np_ == cols.size and np.array_equal(probes[:, 0], cols)
    prob = oracle.Problem(inst)
    best_l, best_u = lb0.copy(), ub0.copy()
    n_failed = 0
    for k, j in enumerate(cols):
        outs = []
        for way in (0, 1):
            lb, ub = lb0.copy(), ub0.copy()
            if way == 0:
                ub[j] = lb0[j]
            else:
                lb[j] = lb0[j] + 1.0
            r, lo, uo = oracle.propagate(prob, lb, ub)
            assert probes[k, 1 + way] == (0 if r["feasible"] else 1), (j, way)
            outs.append((r["feasible"], lo, uo))
        (fd, ld, ud), (fu, lu, uu) = outs
        n_failed += (not fd) + (not fu)
        assert fd or fu
        if not fd:
            best_l[j] = max(best_l[j], lb0[j] + 1.0)
        elif not fu:
            best_u[j] = min(best_u[j], lb0[j])
        else:
            both = np.minimum(ld, lu), np.maximum(ud, uu)
            both[0][j], both[1][j] = lb0[j], ub0[j]
            best_l, best_u = np.maximum(best_l, both[0]), np.minimum(best_u, both[1])
    assert res["failedProbes"] == n_failed and res["nodeInfeasible"] == 0
    want = np.flatnonzero((best_l > lb0) | (best_u < ub0))
    assert np.array_equal(nbounds[:, 0].astype(int), want) and best_u[a] == 0.0 and best_l[d] == 1.0
    assert np.array_equal(nbounds[:, 1], best_l[want]) and np.array_equal(nbounds[:, 2], best_u[want])
    assert np.array_equal(fin[:inst.n], best_l) and np.array_equal(fin[inst.n:], best_u) and res["moved"] == want.size
    assert inst.contains(best_l, best_u)  # the planted point survives every conclusion


def test_node_mode_calls_are_cheap_on_the_host(tmp_path):
    """1,000 run() calls on one solver (a B&B node re-presents its matrix every time): the C++ layer — finding the device
    copy again, copying bounds, writing changes back — must stay a small share of the call; the matrix is uploaded once
    and found again by address + light key every other time."""
    inst = I.mixed_bigm(m=100_000, n=100_000, seed=91)  # ~1M nonzeros
    res, _ = _drive_mode(tmp_path, inst, ["bench", "1000"], level=3, max_rounds=0)
    c = res["cache"]
    assert c["uploads"] == 1 and c["pointerHits"] >= 2000 and c["fullHashes"] == 1
    assert res["acquire_share"] < 0.05, res
    print("node-mode bench:", res)
